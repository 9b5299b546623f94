/* sm_100a device code of the BPE encode engine, part 1: data layout, PTX helpers, table probes, the
 * unique-pretoken cache and the block-cooperative merge of long pretokens.  Included only by hut_engine.cu (through
 * hut_tile.cuh / hut_pipeline.cuh).
 *
 * Pipeline per shard (hut_pipeline.cuh), all on one stream, no host round trip, no inter-CTA waiting:
 *
 *   doc_kernel      one thread per document: first document starting in each 960-byte tile
 *   tile_kernel     (hut_tile.cuh) persistent, WARP per tile, no block barrier:
 *                   TMA bulk copy of the tile (+ 32-byte halos) into the warp's shared-memory slice; bit-sliced class
 *                   masks + local start rule (hut_bitslice.h)           <- parser_next_token (parser.c:24-88)
 *                   lane-per-pretoken probe of the unique-pretoken cache (32-byte slots, one L2 sector per probe)
 *                                                       <- (no equivalent: the reference re-merges every occurrence)
 *                   cache misses reserve room in the slot and are merged 32 at a time, one pretoken per lane
 *                   (merge_entry)                       <- bpe_encode_arena_string (core.c:66-209), MinPQ (queue.c),
 *                                                          get_pair_rank_from_strings (core.c:700-722), hashmap.c
 *                   ids -> the tile's slot of the staging buffer, in order; per-tile counts
 *   mid_kernel      deferred pretokens of up to 1024 / 4096 symbols: one warp each, arrays in shared memory
 *   long_kernel     the rest up to 4160 symbols: one CTA each, tournament tree over (rank, position)
 *   huge_kernel     longer ones: cooperative rounds (all pairs of the minimum rank at once, checked per round)
 *   scan kernels    exclusive scan of the per-tile id counts      <- IntVector growth (vector.c:41-55)
 *   compact_kernel  staging slots -> dense int32 ids, coalesced   <- list building (lib.c:699-715, 810-860)
 *   fixdoc_kernel   per-document id offsets
 *   small_kernel    all of the above in one launch of one CTA for a small batch (hutoken.encode of a sentence)
 *
 * (A single-pass variant -- ids kept in shared memory until a decoupled look-back delivers the tile's base -- was
 * built and measured in round 2 and is in the history: the in-order hand-over of bases stalls warps behind every
 * slow tile and the shared memory it needs leaves the SM without L1; see profiles/README.md.)
 *
 * Algorithmic HBM traffic of the path: B_in + 4*T + 8*(N_docs+1) bytes.
 */
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "hut_bitslice.h"
#include "hut_classify.h"
#include "hut_model.h"

namespace hut {

constexpr int OWN = 960;              // text bytes owned by one tile = 30 chunks of 32 bytes
constexpr int CHUNK = 32;
constexpr int BACK = 32;              // chunk 0 of the span: look-back halo (its own results are discarded)
constexpr int SPAN = 1024;            // chunks 0..31; chunk 31 is the look-ahead halo
constexpr int LOADB = SPAN + 16;      // bytes staged per tile (16 more: look-ahead of the halo's last chars)
constexpr int TEXTB = LOADB + 16;     // shared bytes reserved (unaligned 20-byte key reads stay inside)
constexpr int TRUST = SPAN - 3;       // boundaries at span positions >= TRUST lack their look-ahead
#ifndef HUT_WPB
#define HUT_WPB 8
#endif
constexpr int WPB = HUT_WPB;          // warps per CTA (they never synchronise with each other)
constexpr int NT = WPB * 32;
constexpr int MAXW = 32;              // longest pretoken (bytes) / most symbols handled inside tile_kernel
constexpr int CACHE_KEY_MAX = 15;     // longest pretoken (bytes) the cache keys on
constexpr int CACHE2_KEY_MAX = 31;    // ... the second cache (64-byte slots) keys on
constexpr int CACHE_PROBES = 8;
constexpr int CACHE_IDS = 5;          // most ids a cache value holds
constexpr int MAX_EVENTS = 1024;
constexpr int SCAN_BLK = 4096;        // tiles per scan block
constexpr int LNT = 128;              // threads per CTA of long_kernel (pretokens of up to med_cap symbols); measured on the
                                      // heavy-tail mix: 8 CTAs of 128 per SM 8.6 ms, 16 of 32 10.1 ms, 2 of 128 13.2 ms
constexpr int LSM_CAP = 4096;         // symbols of a pretoken whose working arrays long_kernel keeps in shared memory
constexpr int HNT = 1024;             // threads per CTA of huge_kernel (longer ones: cooperative rounds)
constexpr int LMAXW = 32;             // most warps a CTA of the long path has
constexpr int MISS_SMALL = 32;        // most initial symbols of a cache miss tile_kernel merges itself (one lane each)
constexpr int32_t HOLE = (int32_t)0x80000000;  // staging filler behind the ids of a cache miss merged later
constexpr uint32_t CNT_CHAIN = 0x80000000u;    // cnt[tile] flag: the tile has pretokens deferred to long_kernel
constexpr uint32_t CNT_HOLES = 0x40000000u;    // cnt[tile] flag: the tile's slot holds holes (misses merged later)
constexpr uint32_t CNT_MASK = 0x3FFFFFFFu;
constexpr uint32_t NONE = HUT_UNIT_MASK;
constexpr uint32_t RAWCHAR = 0xFFFFFFFEu;
constexpr uint32_t ENDLINK = 0xFFFFFFFFu;
constexpr uint32_t NODOC = 0xFFFFFFFFu;
constexpr uint32_t BEYOND = 0xFFFFu;  // pstart sentinel: the pretoken's end lies beyond the halo

enum : uint32_t {
    FLAG_NUL = 1u,        // a 0x00 byte inside a document
    FLAG_OUT_CAP = 2u,    // caller's id buffer too small
    FLAG_BIG_CAP = 4u,    // pretoken longer than the configured working area (no-limit mode)
    FLAG_EVENTS = 8u,     // more over-long pretokens than MAX_EVENTS
    FLAG_DEF_CAP = 16u,   // deferred-pretoken list full   (host retries with a larger one)
    FLAG_LONG_CAP = 32u,  // ids of deferred pretokens do not fit (host retries with a larger buffer)
    FLAG_BIG_NEED = 64u   // a pretoken needs a working area of the big pool, which does not exist yet (host retries)
};

struct __align__(16) CacheSlot {  // 32 bytes = one L2 sector
    unsigned long long k0, k1;    // text bytes 0..14, meta (0x80 | first << 4 | len) in byte 15
    uint32_t w[4];                // ids 0..3 as signed 21-bit fields in bits 0..20 of w[0..3]; count in w[0] bits
                                  // 21..23; id 4 spread over w[0] bits 24..31, w[1] bits 21..31, w[2] bits 21..22
};

struct __align__(16) CacheSlot2 {  // 64 bytes: the same for pretokens of 16..31 bytes
    uint32_t k[8];                 // text bytes 0..30, meta (0x80 | first << 5 | len) in byte 31
    uint32_t w[4];                 // as CacheSlot::w
    uint32_t pad[4];
};

struct CutEvent {  // a document cut at an over-long pretoken (core.c:402-407, 503)
    unsigned long long tok_index;  // global index of the first id to drop
    unsigned long long pos;        // byte position of the pretoken
};

struct DefEntry {  // a pretoken tile_kernel hands to long_kernel
    unsigned long long pos;   // global byte position of its first byte
    uint32_t tile;
    uint32_t local_off;       // ids of its tile that precede it in the staging slot
    uint32_t len;             // bytes, 0 = unknown (runs past the halo)
    uint32_t flags;           // bit 0: first pretoken of its document, bit 1: cut (set by long_kernel)
    uint32_t count;           // ids produced (long_kernel)
    uint32_t pad;
    unsigned long long ids_off;  // into long_ids
};

struct Counters {
    unsigned int flags;
    unsigned int n_events;
    unsigned int n_def;
    unsigned int ticket;           // next deferred entry to hand out: mid_kernel<1024>
    unsigned int ticket2;          // ... mid_kernel<4096>
    unsigned int tile_ticket;      // next tile to hand out: tile_kernel
    unsigned long long total_ids;
    unsigned long long long_used;  // bump pointer into long_ids
    unsigned long long st_pretokens, st_hits, st_warp, st_block, st_big, st_cut;
#ifdef HUT_LONGSTAT
    unsigned long long lp[4];      // huge_kernel: rounds applied at once, rounds replaced by a stretch, merges applied at
                                   // once, rounds that applied a prefix only (build switch, not shipped)
#endif
    CutEvent events[MAX_EVENTS];
};

struct DevModel {
    const unsigned long long* pair_tab;
    uint32_t pair_mask;
    const int32_t* out_id;
    const int32_t* rank_of;
    uint32_t n_vocab, n_units, inert;
    int distinct_values, identity_ids, byte_mode, has_prefix, cache_ok;
    int monotone;              // every pair containing a merged token ranks above that token (hut_model.c)
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
    uint32_t miss_len_max;     // longest pretoken (bytes) tile_kernel merges itself (batches of 32, one per lane)
    uint32_t first_extra;      // extra ids the first pretoken of a document can produce (prefix symbols / ids)
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
    uint32_t stride;           // ids per staging slot of a tile in which no document starts
    uint32_t slot_extra;       // more ids per document starting in the tile (prefix symbols / ids); 0 = no prefix
    const unsigned long long* slot_off;  // [ntiles + 1] start of each tile's slot: t * stride + docs_before(t) * slot_extra
    int32_t* stage;
    uint32_t* cnt;             // [ntiles] entries used in the slot (CNT_MASK) | CNT_CHAIN | CNT_HOLES
    uint32_t* tot;             // [ntiles] ids of the tile incl. deferred ones; after scan1: exclusive in-block prefix
    unsigned long long* bsum;  // [nblocks] block sums; after scan2: exclusive prefix
    uint32_t nblocks;
    uint32_t* first_doc;       // [ntiles + 1]
    uint32_t* def_first;       // [ntiles] index of the tile's first DefEntry (valid if CNT_CHAIN)
    DefEntry* def;
    uint32_t def_cap;
    int32_t* long_ids;
    unsigned long long long_cap;
    Counters* ctr;
    CacheSlot* cache;
    uint32_t cache_bits;
    CacheSlot2* cache2;
    uint32_t cache2_bits;
    // long_kernel working areas (global memory, L2 resident)
    uint32_t* cta_scratch;
    unsigned long long cta_scratch_words;  // per CTA
    uint32_t med_cap;                      // symbols per medium array
    uint32_t* big_pool;
    unsigned long long big_slot_words;
    uint32_t big_cap;  // symbols per big array
    uint32_t big_slots;
};

// staging slot of a tile: room for one id per symbol of the pretokens starting in it, plus the prefix ids of the
// documents starting in it (so a prefix model does not pay the prefix room on every tile)
__device__ __forceinline__ int32_t* stage_slot(const Job& J, unsigned long long tile) {
    return J.stage + (J.slot_extra ? __ldg(J.slot_off + tile) : tile * J.stride);  // the table exists for prefix models only
}

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
// one 32-byte cache slot = one L2 sector, fetched by a single 256-bit load (LDG.E.256, sm_100+) that bypasses L1
// (a slot is not re-read soon by the same SM) and is kept in L2 against the streamed text and ids.  Measured on
// the 1 GiB bench: -2 % kernel time against two 128-bit default loads; evict_first hints on the text TMA and the
// staging stores made no difference and are not used.
__device__ __forceinline__ void ld_slot256(const void* p, uint4& a, uint4& b) {
    asm volatile("ld.global.L1::no_allocate.L2::evict_last.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
                 : "l"(p));
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// streaming accesses for data touched exactly once
__device__ __forceinline__ void st_stream_s32(int32_t* p, int32_t v) {
    asm volatile("st.global.cs.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_stream_s32(const int32_t* p) {
    int32_t v;
    asm volatile("ld.global.cs.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void set_flag(const Job& J, uint32_t f) { atomicOr(&J.ctr->flags, f); }

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
    const uint64_t key = ((uint64_t)l << 21) | r;
    uint32_t i = hut_pair_hash(l, r, M.pair_mask);
    for (;;) {
        unsigned long long s = __ldg(M.pair_tab + i);
        if (s == HUT_PAIR_EMPTY) return NONE;
        if ((s >> 21) == key) return (uint32_t)(s & HUT_UNIT_MASK);
        i = (i + 1) & M.pair_mask;
    }
}

// both new neighbour pairs of a merge, (l, c) and (c, r), with their first loads issued together; l / r == NONE
// means there is no such neighbour
__device__ __forceinline__ void pair_probe2(const DevModel& M, uint32_t l, uint32_t c, uint32_t r, uint32_t* ml,
                                            uint32_t* mr) {
    const bool okc = c < M.inert;
    const bool ok1 = okc && l < M.inert, ok2 = okc && r < M.inert;
    uint32_t i1 = hut_pair_hash(l, c, M.pair_mask), i2 = hut_pair_hash(c, r, M.pair_mask);
    unsigned long long s1 = HUT_PAIR_EMPTY, s2 = HUT_PAIR_EMPTY;
    if (ok1) s1 = __ldg(M.pair_tab + i1);
    if (ok2) s2 = __ldg(M.pair_tab + i2);
    const uint64_t key1 = ((uint64_t)l << 21) | c, key2 = ((uint64_t)c << 21) | r;
    while (s1 != HUT_PAIR_EMPTY && (s1 >> 21) != key1) {
        i1 = (i1 + 1) & M.pair_mask;
        s1 = __ldg(M.pair_tab + i1);
    }
    while (s2 != HUT_PAIR_EMPTY && (s2 >> 21) != key2) {
        i2 = (i2 + 1) & M.pair_mask;
        s2 = __ldg(M.pair_tab + i2);
    }
    *ml = s1 == HUT_PAIR_EMPTY ? NONE : (uint32_t)(s1 & HUT_UNIT_MASK);
    *mr = s2 == HUT_PAIR_EMPTY ? NONE : (uint32_t)(s2 & HUT_UNIT_MASK);
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
// unique-pretoken cache.  Key = the pretoken's bytes (<= 15) + meta byte; value = up to 5 ids (CacheSlot).
// ------------------------------------------------------------------------------------------------
// 16 key bytes of text[s .. s+len) read from shared memory (len <= 15), meta in byte 15.  kmask = the
// per-length byte masks (kmask[len] = 4 words), also in shared memory.
__device__ __forceinline__ uint32_t build_key(const uint8_t* text, const uint4* kmask, int s, int len, uint32_t first,
                                              uint32_t x[4]) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(text + (s & ~3));
    const uint32_t sh = (uint32_t)(s & 3) * 8u;
    const uint32_t r0 = w[0], r1 = w[1], r2 = w[2], r3 = w[3], r4 = w[4];
    const uint4 m = kmask[len];
    const uint32_t x0 = __funnelshift_r(r0, r1, sh);
    x[0] = x0 & m.x;
    x[1] = __funnelshift_r(r1, r2, sh) & m.y;
    x[2] = __funnelshift_r(r2, r3, sh) & m.z;
    x[3] = (__funnelshift_r(r3, r4, sh) & m.w) | ((0x80u | (first << 4) | (uint32_t)len) << 24);
    return x0 & 0xFFu;  // the first text byte
}

__device__ __forceinline__ uint32_t cache_hash_raw(const uint32_t x[4]) {
    uint32_t h = x[0] * 0x9E3779B1u;
    h += x[1] * 0x85EBCA77u;
    h += x[2] * 0xC2B2AE3Du;
    h += x[3] * 0x27D4EB2Fu;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    return h;
}
__device__ __forceinline__ uint32_t cache_hash(const uint32_t x[4], uint32_t bits) { return cache_hash_raw(x) >> (32 - bits); }

__device__ __forceinline__ int32_t sext21(uint32_t w) {
    int32_t r;
    asm("bfe.s32 %0, %1, 0, 21;" : "=r"(r) : "r"(w));
    return r;
}
__device__ __forceinline__ uint32_t val_count(uint32_t w0) { return (w0 >> 21) & 7u; }
__device__ __forceinline__ int32_t val_id4(const uint32_t w[4]) {
    return sext21((w[0] >> 24) | ((w[1] >> 21) << 8) | (((w[2] >> 21) & 3u) << 19));
}
// packs up to 5 ids (each in [-2^20, 2^20)) + their count into a slot value
__device__ __forceinline__ void val_pack(const int32_t t[5], uint32_t total, uint32_t w[4]) {
    const uint32_t F = 0x1FFFFFu;
    const uint32_t f4 = total > 4 ? ((uint32_t)t[4] & F) : 0u;
    w[0] = (total > 0 ? ((uint32_t)t[0] & F) : 0u) | (total << 21) | ((f4 & 0xFFu) << 24);
    w[1] = (total > 1 ? ((uint32_t)t[1] & F) : 0u) | (((f4 >> 8) & 0x7FFu) << 21);
    w[2] = (total > 2 ? ((uint32_t)t[2] & F) : 0u) | (((f4 >> 19) & 3u) << 21);
    w[3] = (total > 3 ? ((uint32_t)t[3] & F) : 0u);
}

// true on a hit; a slot whose value is still being written (count field 0) reads as a miss
__device__ __forceinline__ bool cache_lookup(const Job& J, const uint32_t x[4], uint32_t w[4]) {
    const uint32_t mask = (1u << J.cache_bits) - 1;
    uint32_t i = cache_hash(x, J.cache_bits);
#pragma unroll 1
    for (int p = 0; p < CACHE_PROBES; ++p) {
        uint4 a, b;
        ld_slot256(J.cache + i, a, b);
        if ((a.x | a.y) == 0) return false;
        if (a.x == x[0] && a.y == x[1] && a.z == x[2] && a.w == x[3]) {
            w[0] = b.x;
            w[1] = b.y;
            w[2] = b.z;
            w[3] = b.w;
            return val_count(b.x) != 0;
        }
        i = (i + 1) & mask;
    }
    return false;
}

// the probes after the first one (tile_kernel issues the first probe of several pretokens together)
__device__ __forceinline__ bool cache_lookup_from(const Job& J, const uint32_t x[4], uint32_t w[4], uint32_t i,
                                                  int probes) {
    const uint32_t mask = (1u << J.cache_bits) - 1;
#pragma unroll 1
    for (int p = 0; p < probes; ++p) {
        uint4 a, b;
        ld_slot256(J.cache + i, a, b);
        if ((a.x | a.y) == 0) return false;
        if (a.x == x[0] && a.y == x[1] && a.z == x[2] && a.w == x[3]) {
            w[0] = b.x;
            w[1] = b.y;
            w[2] = b.z;
            w[3] = b.w;
            return val_count(b.x) != 0;
        }
        i = (i + 1) & mask;
    }
    return false;
}

// ---- second cache: keys of 16..31 text bytes ----
__device__ __forceinline__ void build_key32(const uint8_t* text, int s, int len, uint32_t first, uint32_t x[8]) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(text + (s & ~3));
    const uint32_t sh = (uint32_t)(s & 3) * 8u;
    uint32_t r[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) r[i] = w[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) {  // byte masks worked out here (rare path): no table in shared memory
        const int rem = len - 4 * i;
        const uint32_t m = rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : ((1u << (8 * rem)) - 1u));
        x[i] = __funnelshift_r(r[i], r[i + 1], sh) & m;
    }
    x[7] |= (0x80u | (first << 5) | (uint32_t)len) << 24;
}

__device__ __forceinline__ uint32_t cache2_hash(const uint32_t x[8], uint32_t bits) {
    uint32_t h = x[0] * 0x9E3779B1u;
    h += x[1] * 0x85EBCA77u;
    h += x[2] * 0xC2B2AE3Du;
    h += x[3] * 0x27D4EB2Fu;
    h = (h ^ (h >> 15)) * 0x2C1B3C6Du;
    h += x[4] * 0x9E3779B1u;
    h += x[5] * 0x85EBCA77u;
    h += x[6] * 0xC2B2AE3Du;
    h += x[7] * 0x27D4EB2Fu;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    return h >> (32 - bits);
}

__device__ __forceinline__ bool cache2_lookup(const Job& J, const uint32_t x[8], uint32_t w[4]) {
    const uint32_t mask = (1u << J.cache2_bits) - 1;
    uint32_t i = cache2_hash(x, J.cache2_bits);
#pragma unroll 1
    for (int p = 0; p < CACHE_PROBES; ++p) {
        const uint4* s = reinterpret_cast<const uint4*>(J.cache2 + i);
        const uint4 a = s[0], b = s[1];
        if ((a.x | a.y) == 0) return false;
        if (a.x == x[0] && a.y == x[1] && a.z == x[2] && a.w == x[3] && b.x == x[4] && b.y == x[5] && b.z == x[6] &&
            b.w == x[7]) {
            const uint4 v = s[2];
            w[0] = v.x;
            w[1] = v.y;
            w[2] = v.z;
            w[3] = v.w;
            return val_count(v.x) != 0;
        }
        i = (i + 1) & mask;
    }
    return false;
}

__device__ __forceinline__ void cache2_insert(const Job& J, const uint32_t x[8], const uint32_t w[4]) {
    const unsigned long long k0 = x[0] | ((unsigned long long)x[1] << 32);
    const uint32_t mask = (1u << J.cache2_bits) - 1;
    uint32_t i = cache2_hash(x, J.cache2_bits);
#pragma unroll 1
    for (int p = 0; p < CACHE_PROBES; ++p) {
        CacheSlot2* s = J.cache2 + i;
        unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long*>(s->k), 0ull, k0);
        if (old == 0) {
#pragma unroll
            for (int z = 2; z < 8; ++z) s->k[z] = x[z];
            s->w[1] = w[1];
            s->w[2] = w[2];
            s->w[3] = w[3];
            __threadfence();
            atomicExch(&s->w[0], w[0]);
            return;
        }
        if (old == k0) {  // the same key (inserted or being inserted by someone else)?  Otherwise keep probing.
            bool same = true;
#pragma unroll
            for (int z = 2; z < 8; ++z) same = same && *reinterpret_cast<volatile uint32_t*>(&s->k[z]) == x[z];
            if (same) return;
        }
        i = (i + 1) & mask;
    }
}

// Called by one thread.  Slots are claimed by CAS on k0 (never zero: the first text byte is not NUL); the value
// word carrying the count is published last, so a half-written slot reads as a miss.
__device__ __forceinline__ void cache_insert(const Job& J, const uint32_t x[4], const uint32_t w[4]) {
    const unsigned long long k0 = x[0] | ((unsigned long long)x[1] << 32);
    const unsigned long long k1 = x[2] | ((unsigned long long)x[3] << 32);
    const uint32_t mask = (1u << J.cache_bits) - 1;
    uint32_t i = cache_hash(x, J.cache_bits);
#pragma unroll 1
    for (int p = 0; p < CACHE_PROBES; ++p) {
        CacheSlot* s = J.cache + i;
        unsigned long long old = atomicCAS(&s->k0, 0ull, k0);
        if (old == 0) {
            s->k1 = k1;
            s->w[1] = w[1];
            s->w[2] = w[2];
            s->w[3] = w[3];
            __threadfence();
            atomicExch(&s->w[0], w[0]);
            return;
        }
        if (old == k0 && ld_volatile_u64(&s->k1) == k1) {  // already there (or being written)
#ifdef HUT_EXP_DUPCOUNT  // experiment: merges whose result another warp had inserted meanwhile, reported as big_merges
            atomicAdd(&J.ctr->st_big, 1ull);
#endif
            return;
        }
        i = (i + 1) & mask;
    }
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

// ------------------------------------------------------------------------------------------------
// lane-per-pretoken merge.  ONE LANE runs the whole merge loop of one pretoken of at most S symbols (arrays in
// local memory, which the hardware interleaves by lane): per merge one scan for the leftmost minimal rank -- the
// (rank, left index) order of the reference's heap (queue.c:162-164) -- one shift to close the gap and two pair-table
// probes (issued together) for the new neighbours.  Replaces bpe_encode_arena_string (core.c:66-209) for the
// pretokens the cache has not seen; 32 independent pretokens per warp keep the probe latency overlapped.
// The ids go into the `room` entries tile_kernel reserved at dst (the rest of the room becomes HOLEs) and, when the
// pretoken is cacheable, into the cache.  Returns room - ids written.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

template <bool WIDE> struct AliveMask { using type = uint32_t; };
template <> struct AliveMask<true> { using type = unsigned long long; };
__device__ __forceinline__ int al_ffs(uint32_t x) { return __ffs((int)x); }
__device__ __forceinline__ int al_ffs(unsigned long long x) { return __ffsll((long long)x); }
__device__ __forceinline__ int al_top(uint32_t x) { return 31 - __clz((int)x); }
__device__ __forceinline__ int al_top(unsigned long long x) { return 63 - __clzll((long long)x); }
__device__ __forceinline__ int al_popc(uint32_t x) { return __popc(x); }
__device__ __forceinline__ int al_popc(unsigned long long x) { return __popcll(x); }

template <int S>
__device__ __forceinline__ uint32_t merge_entry(const DevModel& M, const Job& J, const uint32_t* byte1,
                                                unsigned long long pos, uint32_t len, uint32_t first, uint32_t room,
                                                int32_t* dst) {
    static_assert(S <= 64 && S % 4 == 0, "positions are packed into 6 bits of the scan keys, which are read four at a time");
    using AL = typename AliveMask<(S > 32)>::type;  // one bit per initial symbol
    uint32_t U[S], MG[S];
    __align__(16) uint32_t KEY[S];
    auto at = [&](unsigned long long p) -> uint32_t { return (uint32_t)__ldg(J.text + p); };
    const uint32_t b0 = at(pos);
    const bool with_prefix = first && b0 != ' ';
    const bool alone_first = first && b0 == ' ';
    const uint32_t npfx = with_prefix ? M.n_prefix_syms : 0u;
    const uint32_t nal = alone_first ? M.n_prefix_alone : 0u;
    // ---- initial symbols (A.3) + the cache key ----
    uint32_t m = 0;
#pragma unroll 1
    for (uint32_t k = 0; k < npfx; ++k) U[m++] = M.prefix_syms[k];
    unsigned long long k0 = 0, k1 = 0, k2 = 0, k3 = 0;
#pragma unroll 1
    for (uint32_t p = 0; p < len; ++p) {
        uint32_t unit = M.inert, pb = 0;
        const uint32_t cnt = char_symbols(M, byte1, at, pos + p, pos, pos + len, &unit, &pb);
        if (cnt == 1) U[m++] = unit;
        else
#pragma unroll 1
            for (uint32_t k = 0; k < cnt; ++k) U[m++] = __ldg(M.sym_pool + pb + k);
        const unsigned long long b = at(pos + p);
        if (p < 8) k0 |= b << (8 * p);
        else if (p < 16) k1 |= b << (8 * (p - 8));
        else if (S > 16 && p < 24) k2 |= b << (8 * (p - 16));
        else if (S > 16 && p < 31) k3 |= b << (8 * (p - 24));
    }
    // ---- merge loop (A.4).  The first slot of every initial pair is prefetched so that the probes overlap.
    // Symbols never move: a merged-away symbol is cleared in `alive` and neighbours are found with bit scans, so the
    // array index of the scans is the same in every lane (coalesced local memory) ----
    const uint32_t m0 = m;
#pragma unroll 1
    for (uint32_t k = 0; k + 1 < m0; ++k) {
        const uint32_t l = U[k], r = U[k + 1];
        if (l < M.inert && r < M.inert) prefetch_l1(M.pair_tab + hut_pair_hash(l, r, M.pair_mask));
    }
    // KEY[k] = rank of the pair starting at position k in bits 6.., k below (ranks are dense, < 2^21: hut_model.c), all
    // ones if the pair does not merge: the minimum over KEY is the reference's next merge, leftmost among equal ranks
    // with distinct values a unit's rank is the unit itself (rank_of_unit): the key then holds the merged unit, and the
    // array of merged units is neither written nor read (a third less local memory in the SM's L1)
    const bool dv = M.distinct_values != 0;
    auto keyof = [&](uint32_t mg, uint32_t k) -> uint32_t { return mg == NONE ? 0xFFFFFFFFu : ((rank_of_unit(M, mg) << 6) | k); };
#pragma unroll 1
    for (uint32_t k = 0; k < m0; ++k) {
        const uint32_t mg = k + 1 < m0 ? pair_probe(M, U[k], U[k + 1]) : NONE;
        if (!dv) MG[k] = mg;
        KEY[k] = keyof(mg, k);
    }
#pragma unroll 1
    for (uint32_t k = m0; k < ((m0 + 3u) & ~3u); ++k) KEY[k] = 0xFFFFFFFFu;  // the scan reads groups of four
    AL alive = m0 >= (uint32_t)(8 * sizeof(AL)) ? ~(AL)0 : (((AL)1 << m0) - (AL)1);
    for (;;) {
        uint32_t bk = 0xFFFFFFFFu;
#pragma unroll 1
        for (uint32_t k = 0; k < m0; k += 4) {
            const uint4 v = *reinterpret_cast<const uint4*>(&KEY[k]);
            bk = min(min(bk, v.x), min(min(v.y, v.z), v.w));
        }
        if (bk == 0xFFFFFFFFu) break;
        const uint32_t j = bk & 63u, mu = dv ? (bk >> 6) : MG[j];
        const uint32_t r = j + 1 + (uint32_t)(al_ffs(alive >> (j + 1)) - 1);  // right partner of j
        alive &= ~((AL)1 << r);
        U[j] = mu;
        if (!dv) MG[r] = NONE;
        KEY[r] = 0xFFFFFFFFu;
        const AL below = alive & (((AL)1 << j) - (AL)1), above = r + 1 < (uint32_t)(8 * sizeof(AL)) ? (alive >> (r + 1)) : (AL)0;
        const int l = below ? al_top(below) : -1;
        const int r2 = above ? (int)(r + 1) + (al_ffs(above) - 1) : -1;
        const uint32_t ul = l >= 0 ? U[l] : NONE, ur = r2 >= 0 ? U[r2] : NONE;
        uint32_t ml, mr;
        pair_probe2(M, ul, mu, ur, &ml, &mr);
        if (l >= 0) {
            if (!dv) MG[l] = ml;
            KEY[l] = keyof(ml, (uint32_t)l);
        }
        if (!dv) MG[j] = mr;  // NONE when j is the last symbol now
        KEY[j] = keyof(mr, j);
    }
    // ---- ids into the reserved room, HOLEs behind them ----
    const uint32_t total = nal + (uint32_t)al_popc(alive);  // <= room by construction (tile_kernel reserves the worst case)
    int32_t t5[CACHE_IDS] = {0, 0, 0, 0, 0};
    {
        uint32_t k = 0;
#pragma unroll 1
        for (; k < nal; ++k) {
            const int32_t id = M.prefix_alone[k];
#pragma unroll
            for (int z = 0; z < CACHE_IDS; ++z)
                if (k == (uint32_t)z) t5[z] = id;
            dst[k] = id;
        }
#pragma unroll 1
        for (uint32_t i = 0; i < m0; ++i) {
            if (!((alive >> i) & (AL)1)) continue;
            const int32_t id = id_of_unit(M, U[i]);
#pragma unroll
            for (int z = 0; z < CACHE_IDS; ++z)
                if (k == (uint32_t)z) t5[z] = id;
            dst[k++] = id;
        }
#pragma unroll 1
        for (; k < room; ++k) dst[k] = HOLE;
    }
    if (M.cache_ok && total >= 1 && total <= (uint32_t)CACHE_IDS) {
        uint32_t pw[4];
        val_pack(t5, total, pw);
        if (len <= (uint32_t)CACHE_KEY_MAX) {
            uint32_t key[4];
            key[0] = (uint32_t)k0;
            key[1] = (uint32_t)(k0 >> 32);
            key[2] = (uint32_t)k1;
            key[3] = ((uint32_t)(k1 >> 32) & 0x00FFFFFFu) | ((0x80u | (first << 4) | len) << 24);
            cache_insert(J, key, pw);
        } else if (S > 16 && len <= (uint32_t)CACHE2_KEY_MAX) {
            uint32_t key[8];
            key[0] = (uint32_t)k0;
            key[1] = (uint32_t)(k0 >> 32);
            key[2] = (uint32_t)k1;
            key[3] = (uint32_t)(k1 >> 32);
            key[4] = (uint32_t)k2;
            key[5] = (uint32_t)(k2 >> 32);
            key[6] = (uint32_t)k3;
            key[7] = (uint32_t)(k3 >> 32) | ((0x80u | (first << 5) | len) << 24);
            cache2_insert(J, key, pw);
        }
    }
    return room - total;
}

// ------------------------------------------------------------------------------------------------
// long_kernel / huge_kernel helpers: one CTA (blockDim.x threads) per deferred pretoken
// ------------------------------------------------------------------------------------------------
constexpr int TREE_SM_KEYS = 1024 + 32 + 8;  // tournament-tree levels of up to 1024 keys stay in shared memory
constexpr int TREE_MAX_LEVELS = 7;           // 32^7 positions

struct LongSm {
    unsigned long long tree[TREE_SM_KEYS];
    uint32_t byte1[256];
    unsigned long long red64[LMAXW];
    uint32_t red32[LMAXW];
    unsigned long long found;
    unsigned long long doc_end;
    unsigned long long ids_off;
    uint32_t slot;
    uint32_t na, nb, nd, chain;  // cooperative rounds: list lengths, "a chain was too long to walk" flag
    uint32_t first_bad, nkeep;   // leftmost occurrence whose merge would create a pair that jumps the queue; applied ones
};

// block-wide exclusive scan of one uint32 per thread; returns exclusive prefix, *total = sum
__device__ __forceinline__ uint32_t block_scan_excl(LongSm& sm, uint32_t v, uint32_t* total) {
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
    const int nwarps = blockDim.x >> 5;
    for (int w = 0; w < nwarps; ++w) {
        uint32_t s = sm.red32[w];
        if (w < warp) wbase += s;
        tot += s;
    }
    *total = tot;
    return wbase + inc - v;
}

// 64-bit min across the block (all threads get the result)
__device__ __forceinline__ unsigned long long block_min64(LongSm& sm, unsigned long long v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, v, d);
        v = o < v ? o : v;
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sm.red64[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long r = sm.red64[0];
    const int nwarps = blockDim.x >> 5;
    for (int w = 1; w < nwarps; ++w) r = sm.red64[w] < r ? sm.red64[w] : r;
    return r;
}

__device__ __forceinline__ uint32_t gbyte(const Job& J, unsigned long long p, unsigned long long end) {
    return p < end ? (uint32_t)__ldg(J.text + p) : 0u;
}

// class of the char starting at p (bytes at/after `end` read as 0); continuation bytes covered by a valid lead
// come back as HC_CONT.  The scalar rule of hut_classify.h, evaluated from global memory.
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
__device__ unsigned long long find_run_end(const Job& J, LongSm& sm, unsigned long long from, unsigned long long lo,
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
        p0 += blockDim.x;
        if (p0 > stop) return p0;
    }
}

// ------------------------------------------------------------------------------------------------
// block-cooperative merge of one pretoken [gs, ge) with its symbol arrays in global memory (L2):
// U unit, MG merged unit of pair (i, next(i)), NX / PV linked list, T tournament-tree levels (fan-out 32) over the
// keys (rank << 32 | position): the next merge of the reference -- the pair with the smallest (rank, left index),
// queue.c:162-164 -- is the tree's root.
//   X == nullptr : one merge per step, applied by one lane, the touched tree groups refreshed by its warp.
//   X != nullptr (huge_kernel; the model's ranks are monotone: every pair containing a merged token ranks above it,
//     checked at hut_create): the order "smallest rank first, leftmost first" then applies ALL pairs of the current
//     minimum rank before anything else, and the pairs they create rank higher.  So a round takes the minimum rank r
//     from the root, walks the tree down to every position whose pair has rank r (all warps, level by level), drops
//     every second one of a chain of overlapping occurrences (the leftmost wins, core.c:111-176), applies the rest at
//     once, re-probes the new neighbour pairs and refreshes the touched tree groups level by level.  A round of one or
//     two occurrences is cheaper done by the single-lane step, which therefore goes first (up to 4 pops).
// cta_merge_run leaves the survivors in U (dead = ENDLINK) and returns their number (+ the ids emitted alone
// before the first pretoken); 0xFFFFFFFF if the symbol count exceeds cap.  cta_merge_emit writes the ids.
// ------------------------------------------------------------------------------------------------
__device__ uint32_t cta_merge_run(const DevModel& M, const Job& J, LongSm& sm, unsigned long long gs,
                                  unsigned long long ge, bool with_prefix, bool alone_first, uint32_t* U, uint32_t* MG,
                                  uint32_t* NX, uint32_t* PV, unsigned long long* T, uint32_t cap, uint32_t* m_out,
                                  uint32_t* X) {
    const uint32_t tid = threadIdx.x;
    const uint32_t LNT = blockDim.x, LNW = blockDim.x >> 5;
    auto at = [&](unsigned long long p) -> uint32_t { return (uint32_t)__ldg(J.text + p); };
    const unsigned long long len = ge - gs;
    const unsigned long long chunk = (len + LNT - 1) / LNT;
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
    for (uint32_t i = tid; i < npfx; i += LNT) U[i] = M.prefix_syms[i];
    uint32_t w = base + npfx;
    for (unsigned long long p = a; p < b; ++p) {
        uint32_t f = M.inert, pb = 0;
        uint32_t c = char_symbols(M, sm.byte1, at, p, gs, ge, &f, &pb);
        if (c == 1) U[w++] = f;
        else
            for (uint32_t k = 0; k < c; ++k) U[w++] = __ldg(M.sym_pool + pb + k);
    }
    __syncthreads();
    for (uint32_t i = tid; i < m; i += LNT) {
        MG[i] = (i + 1 < m) ? pair_probe(M, U[i], U[i + 1]) : NONE;
        NX[i] = (i + 1 < m) ? i + 1 : ENDLINK;
        PV[i] = i ? i - 1 : ENDLINK;
    }
    __syncthreads();
    // ---- tournament tree over the keys (rank << 32 | position), fan-out 32: level l holds the minimum of every group
    // of 32 entries of the level below (level -1 = the positions themselves, keyed from MG).  The merge order is the
    // reference heap's (rank, left index) order (queue.c:162-164); picking the next merge costs one read of the root
    // and refreshing it after a merge costs one warp-wide minimum per level for the (at most three) touched groups,
    // instead of a pass over all m positions.  Small levels live in shared memory, the rest in the scratch area.
    const uint32_t lane = tid & 31u, warp = tid >> 5;
    unsigned long long* lv[TREE_MAX_LEVELS];
    uint32_t lcnt[TREE_MAX_LEVELS];
    int nl = 0;
    {
        uint32_t c = m, sm_used = 0;
        unsigned long long* g = T;
        do {
            c = (c + 31u) >> 5;
            lcnt[nl] = c;
            if (c <= (uint32_t)TREE_SM_KEYS - sm_used && c <= 1024u) {
                lv[nl] = sm.tree + sm_used;
                sm_used += c;
            } else {
                lv[nl] = g;
                g += c;
            }
            ++nl;
        } while (c > 1u && nl < TREE_MAX_LEVELS);
    }
    auto pos_key = [&](uint32_t pos) -> unsigned long long {
        if (pos >= m) return ~0ull;
        const uint32_t mg = MG[pos];
        return mg == NONE ? ~0ull : (((unsigned long long)rank_of_unit(M, mg) << 32) | pos);
    };
    // minimum key of the warp.  Lane order is position order at every level (lane l of group g stands for position /
    // child group g * 32 + l), so among equal ranks the lowest lane holds the smallest key: one hardware reduction
    // of the ranks, one ballot, one shuffle instead of five 64-bit shuffle-and-compare steps.
    auto warp_min = [&](unsigned long long k) -> unsigned long long {
        const uint32_t r = (uint32_t)(k >> 32);
        const uint32_t rm = __reduce_min_sync(0xFFFFFFFFu, r);
        const uint32_t b = __ballot_sync(0xFFFFFFFFu, r == rm);
        return __shfl_sync(0xFFFFFFFFu, k, __ffs(b) - 1);
    };
    for (uint32_t g = warp; g < lcnt[0]; g += LNW) {
        const unsigned long long k = warp_min(pos_key(g * 32u + lane));
        if (lane == 0) lv[0][g] = k;
    }
    __syncthreads();
    for (int l = 1; l < nl; ++l) {
        for (uint32_t g = warp; g < lcnt[l]; g += LNW) {
            const uint32_t c = g * 32u + lane;
            const unsigned long long k = warp_min(c < lcnt[l - 1] ? lv[l - 1][c] : ~0ull);
            if (lane == 0) lv[l][g] = k;
        }
        __syncthreads();
    }
    const int top = nl - 1;
    // one merge: the pair at position i (the tree's root).  Called by warp 0; lane 0 applies it, the warp refreshes
    // the (at most three) touched groups of every level.
    auto pop_one = [&](uint32_t i) {
        uint32_t p = ENDLINK, j = ENDLINK;
        if (lane == 0) {
            // loads first, as few in a row as the links allow (every one of them is an L2 round trip), then the stores
            j = NX[i];
            const uint32_t mi = MG[i];
            p = PV[i];
            const uint32_t nj = NX[j];
            const uint32_t up = p != ENDLINK ? U[p] : NONE;
            const uint32_t unj = nj != ENDLINK ? U[nj] : NONE;
            U[i] = mi;
            U[j] = ENDLINK;  // dead
            MG[j] = NONE;
            NX[i] = nj;
            if (nj != ENDLINK) PV[nj] = i;
            uint32_t ml, mr;
            pair_probe2(M, up, mi, unj, &ml, &mr);  // both new neighbour pairs, probed together
            if (p != ENDLINK) MG[p] = ml;
            MG[i] = mr;  // NONE when i is the last symbol now
        }
        p = __shfl_sync(0xFFFFFFFFu, p, 0);
        j = __shfl_sync(0xFFFFFFFFu, j, 0);
        __syncwarp();
        uint32_t a0 = i >> 5, a1 = j >> 5, a2 = p != ENDLINK ? p >> 5 : a0;
        {  // the three groups side by side (equal ones are simply done again: cheaper than a branch between them)
            const unsigned long long k0 = pos_key(a0 * 32u + lane), k1 = pos_key(a1 * 32u + lane),
                                     k2 = pos_key(a2 * 32u + lane);
            const unsigned long long m0 = warp_min(k0), m1 = warp_min(k1), m2 = warp_min(k2);
            if (lane == 0) {
                lv[0][a0] = m0;
                lv[0][a1] = m1;
                lv[0][a2] = m2;
            }
        }
        __syncwarp();
        for (int l = 1; l <= top; ++l) {
            a0 >>= 5;
            a1 >>= 5;
            a2 >>= 5;
            const unsigned long long* below = lv[l - 1];
            const uint32_t nb = lcnt[l - 1];
            const uint32_t c0 = a0 * 32u + lane, c1 = a1 * 32u + lane, c2 = a2 * 32u + lane;
            const unsigned long long k0 = c0 < nb ? below[c0] : ~0ull, k1 = c1 < nb ? below[c1] : ~0ull,
                                     k2 = c2 < nb ? below[c2] : ~0ull;
            const unsigned long long m0 = warp_min(k0), m1 = warp_min(k1), m2 = warp_min(k2);
            if (lane == 0) {
                lv[l][a0] = m0;
                lv[l][a1] = m1;
                lv[l][a2] = m2;
            }
            __syncwarp();
        }
    };
    auto root = [&]() -> unsigned long long { return *reinterpret_cast<volatile unsigned long long*>(lv[top]); };
    if (X == nullptr) {
        if (warp == 0)
            for (;;) {
                const unsigned long long best = root();
                if (best == ~0ull) break;
                pop_one((uint32_t)best);
            }
    } else {
        // scratch behind the tree: two position / group lists, the dirty-group list and one dirty bitmap per level
        uint32_t* LA = X;
        uint32_t* LB = LA + cap;
        uint32_t* STAMP = LB + cap;              // per position: the last round that selected it
        uint32_t* DL = STAMP + cap;              // <= lcnt[0] groups
        uint32_t* BM = DL + (cap / 32u + 64u);   // bitmaps of all levels, one after the other
        uint32_t bm_off[TREE_MAX_LEVELS];
        {
            uint32_t o = 0;
            for (int l = 0; l < nl; ++l) {
                bm_off[l] = o;
                o += (lcnt[l] + 31u) / 32u + 1u;
            }
            for (uint32_t i = tid; i < o; i += LNT) BM[i] = 0u;
            for (uint32_t i = tid; i < m; i += LNT) STAMP[i] = 0u;
        }
        const uint32_t lt = (1u << lane) - 1u;
        uint32_t round = 0;
        uint32_t stretch = 64;  // single-lane steps after a round that could not be applied at once (doubles up to 64 Ki
                                // while that keeps happening: a model whose ranks are far from monotone stays serial)
        for (;;) {
            ++round;
            __syncthreads();
            unsigned long long best = root();
            if (best == ~0ull) break;
            const uint32_t rmin = (uint32_t)(best >> 32);
            // ---- a handful of occurrences: the single-lane step ----
            if (warp == 0) {
                for (int it = 0; it < 4; ++it) {
                    const unsigned long long b = root();
                    if (b == ~0ull || (uint32_t)(b >> 32) != rmin) break;
                    pop_one((uint32_t)b);
                }
            }
            __syncthreads();
            best = root();
            if (best == ~0ull) break;
            if ((uint32_t)(best >> 32) != rmin) continue;
            // ---- many: every position whose pair has rank rmin, found by walking the tree down, all warps ----
            if (tid == 0) {
                LA[0] = 0u;
                sm.na = 1u;
                sm.chain = 0u;
                sm.first_bad = ENDLINK;
                sm.nkeep = 0u;
            }
            __syncthreads();
            for (int l = top; l >= 0; --l) {
                if (tid == 0) sm.nb = 0u;
                __syncthreads();
                const uint32_t na = sm.na;
                for (uint32_t k = warp; k < na; k += LNW) {
                    const uint32_t c = LA[k] * 32u + lane;
                    unsigned long long key;
                    if (l > 0) key = c < lcnt[l - 1] ? lv[l - 1][c] : ~0ull;
                    else key = pos_key(c);
                    const bool mine = (uint32_t)(key >> 32) == rmin;
                    const uint32_t hm = __ballot_sync(0xFFFFFFFFu, mine);
                    if (hm) {
                        uint32_t base = 0;
                        if (lane == 0) base = atomicAdd(&sm.nb, (uint32_t)__popc(hm));
                        base = __shfl_sync(0xFFFFFFFFu, base, 0);
                        if (mine) LB[base + __popc(hm & lt)] = c;
                    }
                }
                __syncthreads();
                if (tid == 0) sm.na = sm.nb;
                uint32_t* t = LA;
                LA = LB;
                LB = t;
                __syncthreads();
            }
            const uint32_t np = sm.na;  // LA[0 .. np): the positions
            // ---- overlapping occurrences (x, x, x: pairs at p and next(p)): the leftmost wins, so of a chain every
            // second one stays.  A position is dropped iff the number of chain members left of it is odd. ----
            for (uint32_t k = tid; k < np; k += LNT) {
                const uint32_t p = LA[k];
                uint32_t q = PV[p], steps = 0;
                while (q != ENDLINK && steps < 64u) {
                    const uint32_t mg = MG[q];
                    if (mg == NONE || rank_of_unit(M, mg) != rmin) break;
                    ++steps;
                    q = PV[q];
                }
                if (steps >= 64u) sm.chain = 1u;
                if (steps & 1u) LA[k] = ENDLINK;  // dropped
            }
            __syncthreads();
            // ---- which of them can be applied at once?  Going left to right, the single-lane order would create for
            // each of them the pairs (left neighbour, merged) and (merged, right neighbour); as long as these rank above
            // rmin nothing else is touched before the next occurrence is merged, so the result is the same.  With
            // monotone ranks that always holds; otherwise it is checked here (the left neighbour is the merged unit of
            // the occurrence before it when that one ends right before this one) and the round applies only the
            // occurrences LEFT of the first one that fails -- a prefix of the reference's own order -- and then hands
            // that one to the single-lane step. ----
            if (!M.monotone && !sm.chain) {
                for (uint32_t k = tid; k < np; k += LNT) {
                    const uint32_t p = LA[k];
                    if (p != ENDLINK) STAMP[p] = round;
                }
                __syncthreads();
                for (uint32_t k = tid; k < np; k += LNT) {
                    const uint32_t p = LA[k];
                    if (p == ENDLINK) continue;
                    const uint32_t mu = MG[p];
                    const uint32_t nj = NX[NX[p]];
                    const uint32_t right = nj != ENDLINK ? U[nj] : NONE;
                    uint32_t left = NONE;
                    const uint32_t pv = PV[p];
                    if (pv != ENDLINK) {
                        const uint32_t pp = PV[pv];
                        left = (pp != ENDLINK && STAMP[pp] == round) ? MG[pp] : U[pv];
                    }
                    uint32_t ml, mr;
                    pair_probe2(M, left, mu, right, &ml, &mr);
                    if (rank_of_unit(M, ml) <= rmin || rank_of_unit(M, mr) <= rmin) atomicMin(&sm.first_bad, p);
                }
                __syncthreads();
                const uint32_t fb = sm.first_bad;
                if (fb != ENDLINK) {
                    uint32_t kept = 0;
                    for (uint32_t k = tid; k < np; k += LNT) {
                        const uint32_t p = LA[k];
                        if (p == ENDLINK) continue;
                        if (p >= fb) LA[k] = ENDLINK;
                        else ++kept;
                    }
                    if (kept) atomicAdd(&sm.nkeep, kept);
                    __syncthreads();
                }
            }
            const bool partial = sm.first_bad != ENDLINK;
            const uint32_t nkeep = sm.nkeep;
            if (sm.chain || (partial && nkeep == 0u)) {
                // a very long chain (e.g. a run of one letter), or the leftmost occurrence itself would create a pair
                // that jumps the queue: a stretch of exact single-lane steps, then the next round looks again
                if (warp == 0)
                    for (uint32_t it = 0; it < stretch; ++it) {
                        const unsigned long long b = root();
                        if (b == ~0ull) break;
                        pop_one((uint32_t)b);
                    }
                if (stretch < 65536u) stretch *= 2u;
#ifdef HUT_LONGSTAT
                if (tid == 0) atomicAdd(&J.ctr->lp[1], 1ull);
#endif
                continue;
            }
            if (!partial || nkeep >= 32u) stretch = 64;
#ifdef HUT_LONGSTAT
            if (tid == 0) {
                atomicAdd(&J.ctr->lp[0], 1ull);
                atomicAdd(&J.ctr->lp[2], (unsigned long long)(partial ? nkeep : np));
                if (partial) atomicAdd(&J.ctr->lp[3], 1ull);
            }
#endif
            // ---- apply all at once (no two of them share a symbol) ----
            for (uint32_t k = tid; k < np; k += LNT) {
                const uint32_t p = LA[k];
                if (p == ENDLINK) continue;
                const uint32_t j = NX[p];
                U[p] = MG[p];
                U[j] = ENDLINK;  // dead
                MG[j] = NONE;
                const uint32_t nj = NX[j];
                NX[p] = nj;
                if (nj != ENDLINK) PV[nj] = p;
                LB[k] = j;
            }
            if (tid == 0) sm.nd = 0u;
            __syncthreads();
            // ---- new neighbour pairs (both ends may have changed in this round, so this comes after the barrier);
            // the leaf groups of everything touched become dirty ----
            auto dirty = [&](int l, uint32_t g) {
                const uint32_t old = atomicOr(&BM[bm_off[l] + (g >> 5)], 1u << (g & 31u));
                if (!(old & (1u << (g & 31u)))) DL[atomicAdd(&sm.nd, 1u)] = g;
            };
            for (uint32_t k = tid; k < np; k += LNT) {
                const uint32_t p = LA[k];
                if (p == ENDLINK) continue;
                const uint32_t nj = NX[p], pp = PV[p];
                MG[p] = nj != ENDLINK ? pair_probe(M, U[p], U[nj]) : NONE;
                dirty(0, p >> 5);
                dirty(0, LB[k] >> 5);
                if (pp != ENDLINK) {
                    MG[pp] = pair_probe(M, U[pp], U[p]);  // the same value if pp merged in this round too
                    dirty(0, pp >> 5);
                }
            }
            __syncthreads();
            // ---- refresh the tree, level by level, dirty groups only ----
            for (int l = 0; l <= top; ++l) {
                const uint32_t nd = sm.nd;
                // the dirty groups of this level move to LA (the position list is not needed any more)
                for (uint32_t k = tid; k < nd; k += LNT) LA[k] = DL[k];
                __syncthreads();
                if (tid == 0) sm.nd = 0u;
                __syncthreads();
                for (uint32_t k = warp; k < nd; k += LNW) {
                    const uint32_t g = LA[k];
                    const uint32_t c = g * 32u + lane;
                    unsigned long long key;
                    if (l > 0) key = c < lcnt[l - 1] ? lv[l - 1][c] : ~0ull;
                    else key = pos_key(c);
                    key = warp_min(key);
                    if (lane == 0) {
                        lv[l][g] = key;
                        BM[bm_off[l] + (g >> 5)] = 0u;  // cleared as a word: every group of the word is in the list
                        if (l < top) dirty(l + 1, g >> 5);
                    }
                }
                __syncthreads();
            }
            if (partial && warp == 0) {  // the occurrence that failed the check is the root now: the single-lane step
                const unsigned long long b = root();
                if (b != ~0ull && (uint32_t)(b >> 32) == rmin) pop_one((uint32_t)b);
            }
        }
    }
    __syncthreads();
    uint32_t live = 0;
    for (uint32_t i = tid; i < m; i += LNT) live += (U[i] != ENDLINK);
    uint32_t tot2;
    block_scan_excl(sm, live, &tot2);
    *m_out = m;
    return tot2 + (alone_first ? M.n_prefix_alone : 0u);
}

__device__ void cta_merge_emit(const DevModel& M, LongSm& sm, const uint32_t* U, uint32_t m, bool alone_first,
                               int32_t* dst) {
    const uint32_t tid = threadIdx.x;
    const uint32_t LNT = blockDim.x;
    const uint32_t chunk2 = (m + LNT - 1) / LNT;
    uint32_t i0 = chunk2 * tid, i1 = i0 + chunk2;
    if (i0 > m) i0 = m;
    if (i1 > m) i1 = m;
    uint32_t live = 0;
    for (uint32_t i = i0; i < i1; ++i) live += (U[i] != ENDLINK);
    uint32_t tot2;
    uint32_t ob = block_scan_excl(sm, live, &tot2);
    const uint32_t nal = alone_first ? M.n_prefix_alone : 0;
    for (uint32_t i = tid; i < nal; i += LNT) dst[i] = M.prefix_alone[i];
    ob += nal;
    for (uint32_t i = i0; i < i1; ++i) {
        uint32_t u = U[i];
        if (u != ENDLINK) dst[ob++] = id_of_unit(M, u);
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// warp-cooperative merge of one pretoken of at most CAP symbols (mid_kernel): the same order as cta_merge_run --
// the pair with the smallest (rank, left index) first, queue.c:162-164 -- without a tree.  Position q belongs to lane
// q & 31; every lane keeps the minimum key (rank << log2 CAP | position) of its positions in a register, the next merge is
// one hardware reduction over the 32 registers, and after a merge only the lanes that own the (at most three)
// changed positions look at their positions again.  All lanes read the links of the merge themselves (broadcast reads
// of shared memory), lanes 0 and 1 probe the two new neighbour pairs side by side, lane 0 stores.
// W: 4 * CAP words of shared memory of this warp: unit, key, merged unit, links (next | previous << 16).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t MID_END = 0xFFFFu;

// First position >= from where the run of class `cls` ends (find_run_end for one warp); a value > stop once the scan
// has passed `stop` without finding it.
__device__ __forceinline__ unsigned long long warp_find_run_end(const Job& J, unsigned long long from,
                                                                 unsigned long long lo, unsigned long long end,
                                                                 uint32_t cls, unsigned long long stop) {
    const uint32_t lane = threadIdx.x & 31u;
    for (unsigned long long p0 = from;; p0 += 32) {
        const unsigned long long p = p0 + lane;
        bool bnd = true;  // the end of the document ends the run
        if (p < end) {
            const uint32_t c = gclass(J, p, lo, end);
            bnd = (cls == HC_SPACE) ? (c != HC_SPACE) : (c != HC_CONT && c != cls);
        }
        const uint32_t mask = __ballot_sync(0xFFFFFFFFu, bnd);
        if (mask) {
            const unsigned long long f = p0 + (uint32_t)(__ffs(mask) - 1);
            return f < end ? f : end;
        }
        if (p0 + 32 > stop) return p0 + 32;
    }
}

// CAP: 1024 or 4096 symbols.  key = rank << log2(CAP) | position in 32 bits: CAP 4096 needs ranks below 2^20 (the
// host launches that variant only then).  Every lane keeps one minimum per 1024 positions (32 of its own), so looking
// at a changed position again costs 32 shared-memory loads whatever the capacity.
// returns the number of ids (the survivors + the ids emitted alone before the first pretoken); 0xFFFFFFFF if the
// pretoken has more than CAP symbols (nothing done).
template <uint32_t CAP>
__device__ __forceinline__ uint32_t warp_merge_count(const DevModel& M, const Job& J, const uint32_t* byte1,
                                                     unsigned long long gs, unsigned long long ge, bool with_prefix,
                                                     bool alone_first, uint32_t* W, uint32_t* m_out) {
    constexpr uint32_t SUBS = CAP / 1024u;
    constexpr uint32_t POSB = CAP == 1024u ? 10u : 12u;
    static_assert(CAP == 1024u || CAP == 4096u, "capacities with a key layout");
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t* const U = W;
    uint32_t* const KEY = W + CAP;
    uint32_t* const MGv = W + 2 * CAP;
    uint32_t* const LK = W + 3 * CAP;
    auto at = [&](unsigned long long p) -> uint32_t { return (uint32_t)__ldg(J.text + p); };
    const uint32_t len = (uint32_t)(ge - gs);
    const uint32_t chunk = (len + 31u) >> 5;
    unsigned long long a = gs + (unsigned long long)chunk * lane, b = a + chunk;
    if (a > ge) a = ge;
    if (b > ge) b = ge;
    uint32_t cnt = 0;
    for (unsigned long long p = a; p < b; ++p) {
        uint32_t f, pb;
        cnt += char_symbols(M, byte1, at, p, gs, ge, &f, &pb);
    }
    uint32_t inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, inc, d);
        if (lane >= d) inc += t;
    }
    const uint32_t npfx = with_prefix ? M.n_prefix_syms : 0u;
    const uint32_t m = __shfl_sync(FULL, inc, 31) + npfx;
    if (m > CAP) return 0xFFFFFFFFu;
    for (uint32_t i = lane; i < npfx; i += 32) U[i] = M.prefix_syms[i];
    uint32_t w = inc - cnt + npfx;
    for (unsigned long long p = a; p < b; ++p) {
        uint32_t f = M.inert, pb = 0;
        const uint32_t c = char_symbols(M, byte1, at, p, gs, ge, &f, &pb);
        if (c == 1) U[w++] = f;
        else
            for (uint32_t k = 0; k < c; ++k) U[w++] = __ldg(M.sym_pool + pb + k);
    }
    __syncwarp();
    auto key_of = [&](uint32_t mg, uint32_t pos) -> uint32_t {
        return mg == NONE ? 0xFFFFFFFFu : ((rank_of_unit(M, mg) << POSB) | pos);
    };
    uint32_t lmin[SUBS];
#pragma unroll
    for (uint32_t sb = 0; sb < SUBS; ++sb) lmin[sb] = 0xFFFFFFFFu;
    for (uint32_t i = lane; i < m; i += 32) {
        const uint32_t mg = (i + 1 < m) ? pair_probe(M, U[i], U[i + 1]) : NONE;
        const uint32_t k = key_of(mg, i);
        MGv[i] = mg;
        KEY[i] = k;
        LK[i] = ((i + 1 < m) ? i + 1 : MID_END) | ((i ? i - 1 : MID_END) << 16);
#pragma unroll
        for (uint32_t sb = 0; sb < SUBS; ++sb)
            if ((i >> 10) == sb) lmin[sb] = k < lmin[sb] ? k : lmin[sb];
    }
    __syncwarp();
    // A merge needs the two new neighbour pairs probed in the pair table: one or two L2 round trips, by far the longest
    // part of a step.  So the probes are made ahead of time, 32 candidates at once: every lane probes for the best pair
    // of its own positions as if it were merged now (s_*: the position, the three units it saw, the two results), then
    // merges are applied in key order for as long as the best pair's lane holds a guess made from the units that are
    // still there.  A lane whose positions changed guesses again in the next pass; the best pair of all is always
    // freshly guessed at the start of a pass, so every pass applies at least one merge.
    uint32_t s_i = 0xFFFFFFFFu, s_up = 0, s_mi = 0, s_unj = 0, s_ml = NONE, s_mr = NONE;
    for (;;) {
        uint32_t mine = lmin[0];
#pragma unroll
        for (uint32_t sb = 1; sb < SUBS; ++sb) mine = lmin[sb] < mine ? lmin[sb] : mine;
        if (__reduce_min_sync(FULL, mine) == 0xFFFFFFFFu) break;
        // ---- guesses ----
        if (mine != 0xFFFFFFFFu) {
            const uint32_t ci = mine & (CAP - 1u);
            const uint32_t lk = LK[ci];
            const uint32_t cj = lk & 0xFFFFu, cp = lk >> 16;
            const uint32_t cmi = MGv[ci];
            const uint32_t cnj = LK[cj] & 0xFFFFu;
            const uint32_t cup = cp != MID_END ? U[cp] : NONE;
            const uint32_t cunj = cnj != MID_END ? U[cnj] : NONE;
            if (!(s_i == ci && s_up == cup && s_mi == cmi && s_unj == cunj)) {
                pair_probe2(M, cup, cmi, cunj, &s_ml, &s_mr);
                s_i = ci;
                s_up = cup;
                s_mi = cmi;
                s_unj = cunj;
            }
        }
        __syncwarp();
        // ---- merges, in key order, while the guesses hold ----
        for (;;) {
            mine = lmin[0];
#pragma unroll
            for (uint32_t sb = 1; sb < SUBS; ++sb) mine = lmin[sb] < mine ? lmin[sb] : mine;
            const uint32_t best = __reduce_min_sync(FULL, mine);
            if (best == 0xFFFFFFFFu) break;
            const uint32_t i = best & (CAP - 1u);
            const uint32_t lki = LK[i];
            const uint32_t j = lki & 0xFFFFu, p = lki >> 16;
            const uint32_t mi = MGv[i];
            const uint32_t nj = LK[j] & 0xFFFFu;
            const uint32_t up = p != MID_END ? U[p] : NONE;
            const uint32_t unj = nj != MID_END ? U[nj] : NONE;
            const uint32_t owner = i & 31u;
            const uint32_t kj = KEY[j], kp = p != MID_END ? KEY[p] : 0xFFFFFFFFu;  // before the merge
            const bool good = s_i == i && s_up == up && s_mi == mi && s_unj == unj;  // meaningful on the owner
            if (!__shfl_sync(FULL, (int)good, owner)) break;
            const uint32_t ml = __shfl_sync(FULL, s_ml, owner), mr = __shfl_sync(FULL, s_mr, owner);
            if (lane == 0) {
                U[i] = mi;
                U[j] = ENDLINK;  // dead
                MGv[j] = NONE;
                KEY[j] = 0xFFFFFFFFu;
                LK[i] = nj | (p << 16);
                if (nj != MID_END) LK[nj] = (LK[nj] & 0xFFFFu) | (i << 16);
                if (p != MID_END) {
                    MGv[p] = ml;
                    KEY[p] = key_of(ml, p);
                }
                MGv[i] = mr;  // NONE when i is the last symbol now
                KEY[i] = key_of(mr, i);
            }
            __syncwarp();
            // the lanes that own i, j and p look at the 32 positions of theirs around the changed one again
            uint32_t todo = 0;
            if (lane == owner) todo |= 1u << (i >> 10);  // its minimum was this pair
            if (lane == (j & 31u) && kj != 0xFFFFFFFFu) {  // j's pair is gone: matters only if it was the lane's minimum
#pragma unroll
                for (uint32_t t = 0; t < SUBS; ++t)
                    if (t == (j >> 10) && lmin[t] == kj) todo |= 1u << t;
            }
            if (p != MID_END && lane == (p & 31u)) {
                const uint32_t kn = KEY[p];
#pragma unroll
                for (uint32_t t = 0; t < SUBS; ++t)
                    if (t == (p >> 10)) {
                        if (kn <= lmin[t]) lmin[t] = kn;
                        else if (lmin[t] == kp) todo |= 1u << t;
                    }
            }
            while (todo) {
                const uint32_t sb = (uint32_t)__ffs(todo) - 1u;
                todo &= todo - 1u;
                // all 32 loads in flight before the first compare: the step is a chain of latencies, not of instructions
                uint32_t v[32];
                const uint32_t q0 = (sb << 10) + lane;
#pragma unroll
                for (uint32_t r = 0; r < 32; ++r) {
                    const uint32_t q = q0 + r * 32u;
                    v[r] = q < m ? KEY[q] : 0xFFFFFFFFu;
                }
#pragma unroll
                for (uint32_t h = 16; h; h >>= 1)
#pragma unroll
                    for (uint32_t r = 0; r < h; ++r) v[r] = v[r + h] < v[r] ? v[r + h] : v[r];
#pragma unroll
                for (uint32_t t = 0; t < SUBS; ++t)
                    if (t == sb) lmin[t] = v[0];
            }
        }
    }
    __syncwarp();
    *m_out = m;
    uint32_t live = 0;
    for (uint32_t q = lane; q < m; q += 32) live += (U[q] != ENDLINK);
#pragma unroll
    for (int d = 16; d; d >>= 1) live += __shfl_xor_sync(FULL, live, d);
    return live + (alone_first ? M.n_prefix_alone : 0u);
}

// the ids of the survivors, in order (warp_merge_count left them in W)
__device__ __forceinline__ void warp_merge_emit(const DevModel& M, const uint32_t* W, uint32_t m, bool alone_first,
                                                int32_t* dst) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t nal = alone_first ? M.n_prefix_alone : 0u;
    for (uint32_t i = lane; i < nal; i += 32) dst[i] = M.prefix_alone[i];
    uint32_t ob = nal;
    for (uint32_t q0 = 0; q0 < m; q0 += 32) {  // 32 positions at a time, compacted by a ballot
        const uint32_t q = q0 + lane;
        const uint32_t u = q < m ? W[q] : ENDLINK;
        const uint32_t mask = __ballot_sync(FULL, u != ENDLINK);
        if (u != ENDLINK) dst[ob + __popc(mask & ((1u << lane) - 1u))] = id_of_unit(M, u);
        ob += __popc(mask);
    }
}

}  // namespace hut
