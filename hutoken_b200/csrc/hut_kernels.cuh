/* sm_100a kernels of the BPE encode engine.  Included only by hut_engine.cu.
 *
 * One fused, persistent "tile" kernel does the whole hot path:
 *
 *   TMA bulk copy of a 4 KiB text tile (+halo) into shared memory        <- parser.c / pretokenizer.c input scan
 *   per-byte class codes, local pretoken-start rule, ballot-free bitmaps <- parser_next_token (parser.c:24-88)
 *   unique-pretoken cache probe (32-byte slots, L2 resident)             <- (no equivalent: reference re-merges)
 *   warp-per-pretoken merge: redux argmin over pair ranks, shuffle compaction
 *                                                                        <- bpe_encode_arena_string (core.c:66-209),
 *                                                                           MinPQ (queue.c), hashmap_get (hashmap.c)
 *   block-cooperative merge for pretokens > 32 symbols / outgrowing the tile
 *   single-pass decoupled look-back scan for output offsets, coalesced int32 stores, per-document offsets
 *                                                                        <- IntVector (vector.c), list building (lib.c)
 *
 * Algorithmic HBM traffic: B_in + 4*T + 8*(N_docs+1) bytes -- the text is read once, ids written once.
 */
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "hut_classify.h"
#include "hut_model.h"

namespace hut {

constexpr int TILE = 4096;            // text bytes owned by one tile
constexpr int BACK = 16;              // look-back halo (needs 4; 16 keeps TMA alignment)
constexpr int FWD = 176;              // look-ahead halo: pretokens ending inside it are handled from shared memory
constexpr int SPAN = BACK + TILE + FWD;
constexpr int LIM = SPAN - 4;         // class codes at/after LIM may lack their look-ahead bytes
constexpr int NT = 128;               // threads per CTA
constexpr int NW = NT / 32;
constexpr int NSEG = SPAN / 32;       // 32-byte segments per span
constexpr int PCAP = 2048;            // pretokens handled per pass over a tile
constexpr int STAGE = 2048;           // ids staged in shared memory per coalesced flush
constexpr int CACHE_KEY_MAX = 23;     // longest pretoken (bytes) the cache keys on
constexpr int CACHE_PROBES = 8;
constexpr int MAX_EVENTS = 1024;
constexpr uint32_t NONE = HUT_UNIT_MASK;
constexpr uint32_t RAWCHAR = 0xFFFFFFFEu;
constexpr uint32_t ENDLINK = 0xFFFFFFFFu;
constexpr uint64_t ESC = 1ull << 63;

static_assert(SPAN % 32 == 0 && (SPAN % 16) == 0, "span must be TMA / segment aligned");

enum : uint32_t {
    FLAG_NUL = 1u,        // a 0x00 byte inside a document
    FLAG_OUT_CAP = 2u,    // caller's id buffer too small
    FLAG_BIG_CAP = 4u,    // pretoken longer than the configured working area (no-limit mode)
    FLAG_EVENTS = 8u      // more over-long pretokens than MAX_EVENTS
};

struct CacheSlot {  // 32 bytes = one L2 sector
    unsigned long long k0, k1, k2, val;
};

struct CutEvent {  // a document cut at an over-long pretoken (core.c:402-407, 503)
    unsigned long long tok_index;  // global index of the first id to drop
    unsigned long long pos;        // byte position of the pretoken
};

struct Counters {
    unsigned int ticket;
    unsigned int flags;
    unsigned int n_events;
    unsigned int pad;
    unsigned long long total_ids;
    unsigned long long st_pretokens, st_hits, st_warp, st_block, st_big, st_cut;
    unsigned int pool_lock[32];
    CutEvent events[MAX_EVENTS];
};

struct DevModel {
    const unsigned long long* pair_tab;
    uint32_t pair_mask;
    const int32_t* out_id;
    const int32_t* rank_of;
    uint32_t n_vocab, n_units, inert;
    int distinct_values, identity_ids, byte_mode, multi_sym, has_prefix, cache_ok;
    const uint32_t* sym_off;   // [257]
    const uint32_t* sym_pool;
    const uint32_t* byte1;     // [256] the single unit of a byte / replaced lead byte; RAWCHAR; NONE = use the pool
    const uint32_t* ascii_unit;  // [128]
    const unsigned long long* char_tab;
    uint32_t char_mask;        // 0 if no table
    const uint32_t* prefix_syms;
    uint32_t n_prefix_syms;
    const int32_t* prefix_alone;
    uint32_t n_prefix_alone;
    uint32_t kmax;
    unsigned long long limit;  // longest pretoken merged; beyond: cut (compat) or error (no-limit)
    int compat_cut;
};

struct Job {
    const uint8_t* text;
    unsigned long long nbytes;
    const unsigned long long* doc_off;
    unsigned long long ndocs;
    int32_t* out;
    unsigned long long out_cap;
    unsigned long long* tok_off;
    uint32_t ntiles;
    unsigned long long* tile_state;
    uint32_t* first_doc;
    Counters* ctr;
    CacheSlot* cache;
    uint32_t cache_bits;
    // per-CTA scratch in global memory (L2 resident): medium merge arrays + overflow ids
    uint32_t* cta_scratch;
    unsigned long long cta_scratch_words;  // per CTA
    uint32_t med_cap;                      // symbols per medium array
    uint32_t ovf_cap;                      // overflow ids per CTA
    // pool of working areas for pretokens that outgrow the tile
    uint32_t* big_pool;
    unsigned long long big_slot_words;
    uint32_t big_cap;  // symbols per big array
    uint32_t big_slots;
};

// ------------------------------------------------------------------------------------------------
// small PTX helpers (mbarrier + 1-D TMA bulk copy)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// streaming store for ids: written once, never re-read by this kernel
__device__ __forceinline__ void st_stream_s32(int32_t* p, int32_t v) {
    asm volatile("st.global.cs.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ------------------------------------------------------------------------------------------------
// table probes
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    return x;
}

// (L,R) -> merged unit, NONE if the concatenation is not a vocab entry.  Replaces
// get_pair_rank_from_strings (core.c:700-722) + hashmap_get (hashmap.c:232-257) + murmur (hash.c).
__device__ __forceinline__ uint32_t pair_probe(const DevModel& M, uint32_t l, uint32_t r) {
    if (l >= M.inert || r >= M.inert) return NONE;
    uint64_t key = ((uint64_t)l << 21) | r;
    uint32_t i = (uint32_t)(mix64(key) >> 32) & M.pair_mask;
    for (;;) {
        unsigned long long s = __ldg(M.pair_tab + i);
        if (s == HUT_PAIR_EMPTY) return NONE;
        if ((s >> 21) == key) return (uint32_t)(s & HUT_UNIT_MASK);
        i = (i + 1) & M.pair_mask;
    }
}

__device__ __forceinline__ uint32_t rank_of_unit(const DevModel& M, uint32_t mg) {
    if (mg == NONE) return 0xFFFFFFFFu;
    return M.distinct_values ? mg : (uint32_t)__ldg(M.rank_of + mg);
}

__device__ __forceinline__ int32_t id_of_unit(const DevModel& M, uint32_t u) {
    if (u >= M.n_vocab) return -1;  // virtual unit or inert: core.c:205-207
    return M.identity_ids ? (int32_t)u : __ldg(M.out_id + u);
}

__device__ __forceinline__ uint32_t char_probe(const DevModel& M, uint32_t key) {
    if (M.char_mask == 0) return M.inert;
    uint32_t i = (uint32_t)(mix64(key) >> 32) & M.char_mask;
    for (;;) {
        unsigned long long s = __ldg(M.char_tab + i);
        if (s == HUT_CHAR_EMPTY) return M.inert;
        if ((uint32_t)(s >> 32) == key) return (uint32_t)s;
        i = (i + 1) & M.char_mask;
    }
}

// ------------------------------------------------------------------------------------------------
// packed values: up to three ids in 3 x 21 bits (id + 2; 0 = empty) -- both the cache value and the
// per-pretoken record in shared memory.  ESC records point into the CTA's overflow ids instead.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t val_count(unsigned long long v) {
    if (v & ESC) return (uint32_t)((v >> 32) & 0x3FFFFFFFu);
    return (uint32_t)((v & 0x1FFFFFull) != 0) + (uint32_t)(((v >> 21) & 0x1FFFFFull) != 0) +
           (uint32_t)(((v >> 42) & 0x1FFFFFull) != 0);
}
__device__ __forceinline__ int32_t val_id(unsigned long long v, uint32_t i) {
    return (int32_t)((v >> (21 * i)) & 0x1FFFFFull) - 2;
}
__device__ __forceinline__ unsigned long long esc_make(uint32_t cnt, uint32_t off) {
    return ESC | ((unsigned long long)cnt << 32) | off;
}

__device__ __forceinline__ uint32_t cache_slot_of(unsigned long long k0, unsigned long long k1, unsigned long long k2,
                                                  uint32_t bits) {
    uint64_t h = k0 * 0x9E3779B97F4A7C15ull;
    h ^= h >> 32;
    h += k1 * 0xC2B2AE3D27D4EB4Full + k2 * 0x165667B19E3779F9ull;
    h ^= h >> 29;
    h *= 0xD6E8FEB86659FD93ull;
    return (uint32_t)(h >> (64 - bits));
}

// 0 = miss (or entry still being written)
__device__ __forceinline__ unsigned long long cache_lookup(const Job& J, unsigned long long k0, unsigned long long k1,
                                                           unsigned long long k2) {
    uint32_t mask = (1u << J.cache_bits) - 1;
    uint32_t i = cache_slot_of(k0, k1, k2, J.cache_bits);
#pragma unroll 1
    for (int p = 0; p < CACHE_PROBES; ++p) {
        const ulonglong2* s = reinterpret_cast<const ulonglong2*>(J.cache + i);
        ulonglong2 a = s[0], b = s[1];
        if (a.x == 0) return 0;
        if (a.x == k0 && a.y == k1 && b.x == k2) return b.y;
        i = (i + 1) & mask;
    }
    return 0;
}

// Called by one thread.  Slots are claimed by CAS on k0 (never zero for a valid key), the value is
// published last; readers treat val == 0 as a miss, so a half-written slot is harmless.
__device__ __forceinline__ void cache_insert(const Job& J, unsigned long long k0, unsigned long long k1,
                                             unsigned long long k2, unsigned long long val) {
    uint32_t mask = (1u << J.cache_bits) - 1;
    uint32_t i = cache_slot_of(k0, k1, k2, J.cache_bits);
#pragma unroll 1
    for (int p = 0; p < CACHE_PROBES; ++p) {
        CacheSlot* s = J.cache + i;
        unsigned long long old = atomicCAS(&s->k0, 0ull, k0);
        if (old == 0) {
            s->k1 = k1;
            s->k2 = k2;
            __threadfence();
            atomicExch(&s->val, val);
            return;
        }
        if (old == k0) {
            unsigned long long o1 = ld_volatile_u64(&s->k1), o2 = ld_volatile_u64(&s->k2);
            if (o1 == k1 && o2 == k2) return;  // same key already there (or being written by another warp)
        }
        i = (i + 1) & mask;
    }
}

// ------------------------------------------------------------------------------------------------
// shared memory of one CTA
// ------------------------------------------------------------------------------------------------
struct __align__(16) Smem {
    uint8_t text[SPAN];
    uint8_t code[SPAN];
    uint32_t docmask[NSEG + 2];
    uint32_t startmask[NSEG + 2];  // pretoken starts
    uint32_t bndmask[NSEG + 2];    // starts | END bytes
    uint16_t wcount[NSEG + 2];
    uint16_t wprefix[NSEG + 2];    // exclusive count of tile-proper starts before this word
    uint16_t pstart[TILE + 4];
    unsigned long long pval[PCAP];
    int32_t stage[STAGE];          // ids staged for a coalesced flush; reused as the miss queue before that
    uint32_t byte1[256];
    uint32_t tbase[NT];
    uint32_t wscratch[NW][32];     // per-warp symbol scatter
    unsigned long long red64[NW];
    uint32_t red32[NW];
    unsigned long long mbar;
    // scalars
    uint32_t tile;
    uint32_t np;
    uint32_t nmiss;
    uint32_t nmed;
    uint32_t ovf_used;
    uint32_t dfirst;
    unsigned long long base;       // exclusive prefix of ids before this tile
    unsigned long long found;      // find_run_end result
    unsigned long long doc_end;
    uint32_t slot;
    uint32_t medq[160];            // pretokens for the block-cooperative path
    uint32_t bigout[96];           // ESC records copied cooperatively: k
    uint32_t nbigout;
    unsigned long long st[6];      // stats accumulators
};

// block-wide exclusive scan of one uint32 per thread; returns exclusive prefix, *total = sum
__device__ __forceinline__ uint32_t block_scan_excl(Smem& sm, uint32_t v, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t n = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += n;
    }
    __syncthreads();  // protect red32 from a previous use
    if (lane == 31) sm.red32[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        uint32_t s = sm.red32[w];
        if (w < warp) wbase += s;
        tot += s;
    }
    *total = tot;
    return wbase + inc - v;
}

// ------------------------------------------------------------------------------------------------
// text access beyond the staged span (block-cooperative / big paths)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t gbyte(const Job& J, unsigned long long p, unsigned long long end) {
    return p < end ? (uint32_t)__ldg(J.text + p) : 0u;
}

// class of the char starting at p (bytes at/after `end` read as 0); continuation bytes covered by a valid lead
// come back as HC_CONT.  Same rule as the tile pass, evaluated from global memory.
__device__ uint32_t gclass(const Job& J, unsigned long long p, unsigned long long lo, unsigned long long end) {
    uint32_t b0 = gbyte(J, p, end);
    if ((b0 & 0xC0u) == 0x80u) {
        for (int k = 1; k <= 3; ++k) {
            if (p < lo + k) break;
            unsigned long long q = p - k;
            uint32_t bq = gbyte(J, q, end);
            if ((bq & 0xC0u) == 0x80u) continue;
            uint32_t r = hut_lead_class(bq, gbyte(J, q + 1, end), gbyte(J, q + 2, end), gbyte(J, q + 3, end));
            if ((r >> 4) > (uint32_t)k) return HC_CONT;
            break;
        }
        return HC_BAD;
    }
    return hut_lead_class(b0, gbyte(J, p + 1, end), gbyte(J, p + 2, end), gbyte(J, p + 3, end)) & 15u;
}

// First position >= from where the run of class `cls` (HC_SPACE = run of spaces) ends; at most `end`.
// Stops early (returns a value > stop) once the scan passes `stop`.  All threads call; result broadcast.
__device__ unsigned long long find_run_end(const Job& J, Smem& sm, unsigned long long from, unsigned long long lo,
                                           unsigned long long end, uint32_t cls, unsigned long long stop) {
    unsigned long long p0 = from;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) sm.found = ~0ull;
        __syncthreads();
        unsigned long long p = p0 + threadIdx.x;
        if (p < end) {
            uint32_t c = gclass(J, p, lo, end);
            bool bnd = (cls == HC_SPACE) ? (c != HC_SPACE) : (c != HC_CONT && c != cls);
            if (bnd) atomicMin(&sm.found, p);
        } else {
            atomicMin(&sm.found, end);
        }
        __syncthreads();
        unsigned long long f = sm.found;
        if (f != ~0ull) return f;
        p0 += NT;
        if (p0 > stop) return p0;
    }
}

// ------------------------------------------------------------------------------------------------
// warp-per-pretoken merge (<= 32 symbols).  Every lane holds one symbol; one merge per round:
// redux.min over ranks + ballot for the leftmost minimum = the (rank, left index) heap order of
// queue.c:152-199; the survivors shift left with one shuffle.  Returns the number of tokens;
// lane i < count holds unit u.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t warp_merge(const DevModel& M, uint32_t& u, uint32_t m) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t un = __shfl_down_sync(0xFFFFFFFFu, u, 1);
    uint32_t mg = (lane + 1 < m) ? pair_probe(M, u, un) : NONE;
    uint32_t rk = rank_of_unit(M, mg);
    for (;;) {
        uint32_t rmin = __reduce_min_sync(0xFFFFFFFFu, rk);
        if (rmin == 0xFFFFFFFFu) break;
        uint32_t j = __ffs(__ballot_sync(0xFFFFFFFFu, rk == rmin)) - 1;
        uint32_t mj = __shfl_sync(0xFFFFFFFFu, mg, j);
        // lane j takes the merged unit; lanes > j take their right neighbour's state
        uint32_t u_r = __shfl_down_sync(0xFFFFFFFFu, u, 1);
        uint32_t mg_r = __shfl_down_sync(0xFFFFFFFFu, mg, 1);
        uint32_t rk_r = __shfl_down_sync(0xFFFFFFFFu, rk, 1);
        if (lane == j) u = mj;
        if (lane > j) {
            u = u_r;
            mg = mg_r;
            rk = rk_r;
        }
        m--;
        if (lane >= m) {
            mg = NONE;
            rk = 0xFFFFFFFFu;
        }
        // only the pairs touching the new token change: (j-1, j) and (j, j+1)
        uint32_t nb = __shfl_down_sync(0xFFFFFFFFu, u, 1);
        if (lane == j || lane + 1 == j) {
            mg = (lane + 1 < m) ? pair_probe(M, u, nb) : NONE;
            rk = rank_of_unit(M, mg);
        }
    }
    return m;
}

// ------------------------------------------------------------------------------------------------
// initial symbols of the char starting at text position p inside pretoken [gs, ge)   (A.3)
//   byte mode : every byte is a char; units = pool[sym_off[b] .. sym_off[b+1])
//   plain mode: chars start at gs and at every non-continuation byte; a replaced lead byte maps to its pool,
//               otherwise the raw char is one unit (ASCII table / char table)
// Returns the number of symbols (0 if p is not a char start); *first = unit if exactly one.
// ------------------------------------------------------------------------------------------------
template <class ByteAt>
__device__ __forceinline__ uint32_t char_symbols(const DevModel& M, const uint32_t* byte1, ByteAt at,
                                                 unsigned long long p, unsigned long long gs, unsigned long long ge,
                                                 uint32_t* first, uint32_t* pool_begin) {
    uint32_t b = at(p);
    if (!M.byte_mode && p != gs && (b & 0xC0u) == 0x80u) return 0;
    uint32_t one = byte1[b];
    if (one == RAWCHAR) {
        if (b < 0x80u) {
            *first = __ldg(M.ascii_unit + b);
            return 1;
        }
        uint32_t key = b;
#pragma unroll
        for (int k = 1; k < 4; ++k) {
            if (p + k >= ge) break;
            uint32_t c = at(p + k);
            if ((c & 0xC0u) != 0x80u) break;
            key |= c << (8 * k);
        }
        *first = char_probe(M, key);
        return 1;
    }
    if (one != NONE) {
        *first = one;
        return 1;
    }
    uint32_t lo = __ldg(M.sym_off + b), hi = __ldg(M.sym_off + b + 1);
    *pool_begin = lo;
    if (hi - lo == 1) *first = __ldg(M.sym_pool + lo);
    return hi - lo;
}

// 64-bit min across the block (all threads get the result)
__device__ __forceinline__ unsigned long long block_min64(Smem& sm, unsigned long long v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, v, d);
        v = o < v ? o : v;
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm.red64[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long r = sm.red64[0];
#pragma unroll
    for (int w = 1; w < NW; ++w) r = sm.red64[w] < r ? sm.red64[w] : r;
    return r;
}

// ------------------------------------------------------------------------------------------------
// block-cooperative merge of one pretoken [gs, ge) with its symbol arrays in global memory (L2):
// U unit, MG merged unit of pair (i, next(i)), NX / PV linked list.  One merge per round, chosen by a
// block-wide min over (rank, index) -- the same order as the reference heap.  Ids go to dst.
// Returns the token count (0xFFFFFFFF if the symbol count exceeds cap).
// ------------------------------------------------------------------------------------------------
__device__ uint32_t cta_merge(const DevModel& M, const Job& J, Smem& sm, unsigned long long gs,
                              unsigned long long ge, bool with_prefix, bool alone_first, uint32_t* U, uint32_t* MG,
                              uint32_t* NX, uint32_t* PV, uint32_t cap, int32_t* dst, unsigned long long dst_cap) {
    const uint32_t tid = threadIdx.x;
    auto at = [&](unsigned long long p) -> uint32_t { return (uint32_t)__ldg(J.text + p); };
    const unsigned long long len = ge - gs;
    const unsigned long long chunk = (len + NT - 1) / NT;
    unsigned long long a = gs + chunk * tid, b = a + chunk;
    if (a > ge) a = ge;
    if (b > ge) b = ge;
    // pass 1: symbols per thread range
    uint32_t cnt = 0;
    for (unsigned long long p = a; p < b; ++p) {
        uint32_t f, pb;
        cnt += char_symbols(M, sm.byte1, at, p, gs, ge, &f, &pb);
    }
    uint32_t total;
    uint32_t base = block_scan_excl(sm, cnt, &total);
    const uint32_t npfx = with_prefix ? M.n_prefix_syms : 0;
    const unsigned long long m64 = (unsigned long long)total + npfx;
    if (m64 > cap) return 0xFFFFFFFFu;
    const uint32_t m = (uint32_t)m64;
    // pass 2: write units
    if (tid < npfx) U[tid] = M.prefix_syms[tid];
    for (uint32_t i = tid + NT; i < npfx; i += NT) U[i] = M.prefix_syms[i];
    uint32_t w = base + npfx;
    for (unsigned long long p = a; p < b; ++p) {
        uint32_t f = M.inert, pb = 0;
        uint32_t c = char_symbols(M, sm.byte1, at, p, gs, ge, &f, &pb);
        if (c == 1) U[w++] = f;
        else
            for (uint32_t k = 0; k < c; ++k) U[w++] = __ldg(M.sym_pool + pb + k);
    }
    __syncthreads();
    for (uint32_t i = tid; i < m; i += NT) {
        MG[i] = (i + 1 < m) ? pair_probe(M, U[i], U[i + 1]) : NONE;
        NX[i] = (i + 1 < m) ? i + 1 : ENDLINK;
        PV[i] = i ? i - 1 : ENDLINK;
    }
    __syncthreads();
    for (;;) {
        unsigned long long best = ~0ull;
        for (uint32_t i = tid; i < m; i += NT) {
            uint32_t mg = MG[i];
            if (mg != NONE) {
                unsigned long long key = ((unsigned long long)rank_of_unit(M, mg) << 32) | i;
                best = key < best ? key : best;
            }
        }
        best = block_min64(sm, best);
        if (best == ~0ull) break;
        if (tid == 0) {
            uint32_t i = (uint32_t)best, j = NX[i];
            U[i] = MG[i];
            U[j] = ENDLINK;  // dead
            MG[j] = NONE;
            uint32_t nj = NX[j];
            NX[i] = nj;
            if (nj != ENDLINK) PV[nj] = i;
            uint32_t p = PV[i];
            if (p != ENDLINK) MG[p] = pair_probe(M, U[p], U[i]);
            MG[i] = (nj != ENDLINK) ? pair_probe(M, U[i], U[nj]) : NONE;
        }
        __syncthreads();
    }
    // survivors in order -> ids
    const uint32_t chunk2 = (m + NT - 1) / NT;
    uint32_t i0 = chunk2 * tid, i1 = i0 + chunk2;
    if (i0 > m) i0 = m;
    if (i1 > m) i1 = m;
    uint32_t live = 0;
    for (uint32_t i = i0; i < i1; ++i) live += (U[i] != ENDLINK);
    uint32_t tot2;
    uint32_t ob = block_scan_excl(sm, live, &tot2);
    const uint32_t nal = alone_first ? M.n_prefix_alone : 0;
    if (tid < nal && tid < dst_cap) dst[tid] = M.prefix_alone[tid];
    ob += nal;
    for (uint32_t i = i0; i < i1; ++i) {
        uint32_t u = U[i];
        if (u != ENDLINK) {
            if (ob < dst_cap) dst[ob] = id_of_unit(M, u);
            ob++;
        }
    }
    __syncthreads();
    return tot2 + nal;
}

// ------------------------------------------------------------------------------------------------
// decoupled look-back over tile_state (status in bits 62-63: 1 aggregate, 2 inclusive prefix).
// Called by warp 0.  publish_aggregate == false: wait-only variant for tiles that cannot know their
// count early (they publish their inclusive prefix themselves when done).
// ------------------------------------------------------------------------------------------------
constexpr unsigned long long ST_AGG = 1ull << 62, ST_INC = 2ull << 62, ST_MASK = 3ull << 62;

__device__ unsigned long long lookback(const Job& J, uint32_t tile, unsigned long long aggregate,
                                       bool publish_aggregate) {
    const uint32_t lane = threadIdx.x & 31;
    if (tile == 0) {
        if (publish_aggregate && lane == 0) {
            __threadfence();
            st_volatile_u64(J.tile_state, ST_INC | aggregate);
        }
        return 0;
    }
    if (publish_aggregate && lane == 0) {
        __threadfence();
        st_volatile_u64(J.tile_state + tile, ST_AGG | aggregate);
    }
    unsigned long long excl = 0;
    long long top = (long long)tile - 1;
    for (;;) {
        long long j = top - lane;
        unsigned long long s;
        if (j >= 0) {
            s = ld_volatile_u64(J.tile_state + j);
            while ((s & ST_MASK) == 0) {
                __nanosleep(40);
                s = ld_volatile_u64(J.tile_state + j);
            }
        } else {
            s = ST_INC;  // before tile 0: inclusive prefix 0
        }
        uint32_t inc_mask = __ballot_sync(0xFFFFFFFFu, (s & ST_MASK) == ST_INC);
        unsigned long long v = s & ~ST_MASK;
        if (inc_mask) {
            uint32_t first = __ffs(inc_mask) - 1;
            if (lane > first) v = 0;
        }
#pragma unroll
        for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
        excl += v;
        if (inc_mask) break;
        top -= 32;
    }
    if (publish_aggregate && lane == 0) {
        __threadfence();
        st_volatile_u64(J.tile_state + tile, ST_INC | (excl + aggregate));
    }
    return excl;
}

}  // namespace hut
