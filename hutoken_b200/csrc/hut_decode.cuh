// Decode path on the device: ids -> text.
//
// Replaces decode() (core.c:513-581: bounds check, concatenation of vocab_decode[id]) and pretokenizer_decode
// (pretokenizer.c:197-296: strip the prefix once, then left to right: longest special-chars replacement found by
// descending the trie of ac.c:27-50 -> the byte it stands for; otherwise byte mode: one UTF-8 character -> the byte
// with that code point, '?' above 255; plain mode: the character is copied).
//
// The scan is greedy and sequential (where a step starts depends on where the previous one ended), so it is done
// in two parts per 1024-byte tile of a document's raw text: every position computes the step it WOULD take
// (independent, parallel), then the positions really visited are found by binary lifting over those steps
// (next^(2^k) tables in shared memory) from the tile's entry offset, which the previous tile hands over.
//
// Data layout: ids int32[n] + tok_off uint64[ndocs+1] -> sub_len uint32[n/256] (token bytes per 256 ids) -> (scan)
// sub_off uint64[n/256+1] -> raw bytes (documents back to back, 64 zero bytes of padding) -> dec bytes written at the
// same base (a step never emits more bytes than it consumes) + doc_raw uint64[ndocs] + doc_len uint32[ndocs] ->
// (scan) text_off uint64[ndocs+1] -> text packed.
#pragma once
#include <cstdint>

#include "hut_model.h"

namespace hut {

struct DecModel {
    const uint32_t* dec_off;   // [vocab_lines + 1]
    const uint32_t* dec_len;   // [vocab_lines], HUT_DEC_MISSING if no line carries the id
    const uint8_t* dec_pool;
    uint32_t vocab_lines;
    const unsigned long long* trie_edge;
    uint32_t trie_mask;        // 0xFFFFFFFF when the trie is empty
    const int16_t* trie_out;
    const uint8_t* trie_depth;
    uint32_t trie_nodes;
    const uint32_t* root_child;  // [256]
    const uint8_t* prefix;
    uint32_t prefix_len;
    int byte_mode;
    // ---- token-wise path (below): the decoded bytes of every token, worked out once per model ----
    const uint8_t* trie_kids;  // [trie_nodes] node has children
    const uint8_t* t_len8;     // [vocab_lines] decoded length; 255: look at t_meta (long, missing or not independent)
    const uint32_t* t_meta;    // [vocab_lines] T_* flags | bytes the prefix stands for << 24 | decoded length
    const uint4* t_slot;       // [vocab_lines] the first 15 decoded bytes, byte 15 = length (255: t_pool)
    const uint32_t* t_off;     // [vocab_lines + 1] offsets into t_pool
    const uint8_t* t_pool;
};

constexpr uint32_t T_MISSING = 1u << 31;       // no vocab line carries the id
constexpr uint32_t T_UNSAFE = 1u << 30;        // the scan of the token alone can differ from its scan inside a text
constexpr uint32_t T_UNSAFE_START = 1u << 29;  // ... when it is the first token of a document (prefix stripping)
constexpr uint32_t T_PS_SHIFT = 24, T_PS_MASK = 31u, T_LEN_MASK = 0x7FFFFFu;
constexpr uint32_t T_SOFT = 1u << 23;  // with T_UNSAFE: only because a trie walk reaches the end of the token; whether it
                                       // goes on to a longer replacement is looked up in the tokens that follow

constexpr int DEC_TILE = 1024;   // raw bytes per tile
constexpr int DEC_LOOK = 32;     // look-ahead: the longest replacement is 26 bytes (32-byte line buffer, lib.c:483)
constexpr int DEC_NT = 128;
constexpr int DEC_LEVELS = 11;   // 2^10 hops cover a tile; one more level keeps the loop uniform
constexpr int SCAN_NT = 256;
constexpr int SCAN_ITEMS = 16;   // per thread
constexpr int SCAN_BLOCK = SCAN_NT * SCAN_ITEMS;

// ------------------------------------------------------------------------------------------------
// exclusive scan uint32[n] -> uint64[n + 1] (three kernels: block sums, scan of the sums, local scan + base)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long block_exclusive_u64(unsigned long long v, unsigned long long* total,
                                                                  unsigned long long* sm /* [32] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    unsigned long long inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) sm[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        unsigned long long w = lane < nw ? sm[lane] : 0ull, winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, winc, d);
            if (lane >= d) winc += t;
        }
        sm[lane] = winc - w;
        if (lane == 31) *total = winc;
    }
    __syncthreads();
    const unsigned long long r = sm[warp] + inc - v;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_NT) scan_sums_kernel(const uint32_t* in, unsigned long long n,
                                                            unsigned long long* bsum) {
    __shared__ unsigned long long sm[32];
    __shared__ unsigned long long tot;
    const unsigned long long base = (unsigned long long)blockIdx.x * SCAN_BLOCK;
    unsigned long long s = 0;
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const unsigned long long i = base + (unsigned long long)k * SCAN_NT + threadIdx.x;
        if (i < n) s += in[i];
    }
    block_exclusive_u64(s, &tot, sm);
    if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) scan_top_kernel(unsigned long long* bsum, uint32_t nblocks) {
    __shared__ unsigned long long sm[32];
    __shared__ unsigned long long tot;
    unsigned long long carry = 0;
    for (uint32_t b0 = 0; b0 < nblocks; b0 += 1024) {
        const uint32_t i = b0 + threadIdx.x;
        const unsigned long long v = i < nblocks ? bsum[i] : 0ull;
        const unsigned long long ex = block_exclusive_u64(v, &tot, sm);
        if (i < nblocks) bsum[i] = carry + ex;
        carry += tot;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SCAN_NT) scan_apply_kernel(const uint32_t* in, unsigned long long n,
                                                             const unsigned long long* bsum, unsigned long long* out) {
    __shared__ unsigned long long sm[32];
    __shared__ unsigned long long tot;
    // thread t owns SCAN_ITEMS consecutive items, so its partial sums are a plain running sum
    const unsigned long long first = (unsigned long long)blockIdx.x * SCAN_BLOCK + (unsigned long long)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        v[k] = first + k < n ? in[first + k] : 0u;
        s += v[k];
    }
    unsigned long long run = bsum[blockIdx.x] + block_exclusive_u64(s, &tot, sm);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        if (first + k <= n) out[first + k] = run;  // index n receives the grand total
        run += v[k];
    }
}

// ------------------------------------------------------------------------------------------------
// ids -> token lengths (bounds check of core.c:520-531) summed per sub-block of DEC_SUB ids; after the scan of the
// sub-block sums, token bytes -> raw text.  No per-id offset array is kept: a position inside a sub-block is
// recomputed from the ids (the length table is 4 bytes per vocab line and stays in L1 / L2).
// ------------------------------------------------------------------------------------------------
constexpr int DEC_SUB = 256;

// HUT_DEC_MISSING for an id outside the vocabulary or carried by no line
__device__ __forceinline__ uint32_t dec_token_len(const DecModel& D, int32_t id) {
    return (id >= 0 && (uint32_t)id < D.vocab_lines) ? __ldg(D.dec_len + id) : (uint32_t)HUT_DEC_MISSING;
}

// warp per sub-block, 8 consecutive ids per lane
__global__ void __launch_bounds__(256) dec_sub_kernel(DecModel D, const int32_t* ids, unsigned long long n,
                                                      uint32_t* sub_len, unsigned int* bad) {
    const int lane = threadIdx.x & 31;
    const unsigned long long blk = (unsigned long long)blockIdx.x * 8u + (threadIdx.x >> 5);
    if (blk * DEC_SUB >= n) return;
    const unsigned long long i0 = blk * DEC_SUB + (unsigned long long)lane * 8u;
    int32_t id[8];
    if (i0 + 8 <= n && (reinterpret_cast<uintptr_t>(ids) & 15u) == 0) {
        const int4 a = reinterpret_cast<const int4*>(ids + i0)[0], b = reinterpret_cast<const int4*>(ids + i0)[1];
        id[0] = a.x; id[1] = a.y; id[2] = a.z; id[3] = a.w;
        id[4] = b.x; id[5] = b.y; id[6] = b.z; id[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) id[k] = i0 + k < n ? ids[i0 + k] : 0;
    }
    uint32_t sum = 0;
    bool missing = false;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (i0 + k >= n) continue;
        const uint32_t l = dec_token_len(D, id[k]);
        if (l == HUT_DEC_MISSING) missing = true;
        else sum += l;
    }
    if (missing) atomicExch(bad, 1u);
#pragma unroll
    for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, d);
    if (lane == 0) sub_len[blk] = sum;
}

// Warp per sub-block (n / DEC_SUB + 1 of them: the last one also writes the 64 zero bytes the last tile's
// look-ahead reads).  Every lane takes 8 consecutive ids, the warp scans their lengths, the token bytes are put
// together in shared memory at the 16-byte phase of their destination and leave in 16-byte stores.
constexpr int GATHER_WARPS = 8;
constexpr int GATHER_STAGE = 2048;  // bytes of one sub-block that fit the staging buffer; larger ones are written directly

__global__ void __launch_bounds__(GATHER_WARPS * 32) dec_gather_kernel(DecModel D, const int32_t* ids,
                                                                       unsigned long long n,
                                                                       const unsigned long long* sub_off,
                                                                       uint8_t* raw) {
    __shared__ uint4 stage[GATHER_WARPS][(GATHER_STAGE + 32) / 16];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long blk = (unsigned long long)blockIdx.x * GATHER_WARPS + warp;
    if (blk > n / DEC_SUB) return;  // warps never synchronise with each other
    const unsigned long long i0 = blk * DEC_SUB + (unsigned long long)lane * 8u;
    int32_t id[8];
    if (i0 + 8 <= n && (reinterpret_cast<uintptr_t>(ids) & 15u) == 0) {
        const int4 a = reinterpret_cast<const int4*>(ids + i0)[0], b = reinterpret_cast<const int4*>(ids + i0)[1];
        id[0] = a.x; id[1] = a.y; id[2] = a.z; id[3] = a.w;
        id[4] = b.x; id[5] = b.y; id[6] = b.z; id[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) id[k] = i0 + k < n ? ids[i0 + k] : -1;
    }
    uint32_t l[8], mine = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        l[k] = dec_token_len(D, id[k]);
        if (l[k] == HUT_DEC_MISSING) l[k] = 0;
        mine += l[k];
    }
    uint32_t inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += t;
    }
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, inc, 31), before = inc - mine;
    uint8_t* const gdst = raw + sub_off[blk];
    if (total <= (uint32_t)GATHER_STAGE) {
        const uint32_t ph = (uint32_t)(reinterpret_cast<uintptr_t>(gdst) & 15u);
        uint8_t* sp = reinterpret_cast<uint8_t*>(stage[warp]) + ph + before;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (!l[k]) continue;
            const uint8_t* src = D.dec_pool + __ldg(D.dec_off + id[k]);
            for (uint32_t c = 0; c < l[k]; ++c) *sp++ = __ldg(src + c);
        }
        __syncwarp();
        const uint32_t end = ph + total;
        const uint8_t* const sb = reinterpret_cast<const uint8_t*>(stage[warp]);
        uint8_t* const gb = gdst - ph;  // 16-byte aligned
        for (uint32_t w = lane; w * 16u < end; w += 32) {
            const uint32_t lo = w * 16u, hi = lo + 16u;
            if (lo >= ph && hi <= end) {
                reinterpret_cast<uint4*>(gb)[w] = stage[warp][w];
            } else {
                const uint32_t a = lo > ph ? lo : ph, b = hi < end ? hi : end;
                for (uint32_t q = a; q < b; ++q) gb[q] = sb[q];
            }
        }
    } else {
        uint8_t* dst = gdst + before;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (!l[k]) continue;
            const uint8_t* src = D.dec_pool + __ldg(D.dec_off + id[k]);
            for (uint32_t c = 0; c < l[k]; ++c) *dst++ = __ldg(src + c);
        }
    }
    if (blk == n / DEC_SUB)
        for (int k = lane; k < 64; k += 32) gdst[total + k] = 0;
}

// ------------------------------------------------------------------------------------------------
// raw text -> decoded text, one CTA per document (grid-stride), tile by tile
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t trie_child(const unsigned long long* edge, uint32_t mask, uint32_t node,
                                               uint32_t byte) {
    const uint32_t key = (node << 8) | byte;
    uint32_t i = hut_trie_hash(key, mask);
    for (;;) {
        const unsigned long long s = edge[i];
        if (s == HUT_TRIE_EMPTY) return 0u;
        if ((uint32_t)(s >> 32) == key) return (uint32_t)s;
        i = (i + 1) & mask;
    }
}

constexpr int DEC_D1 = 4;        // depth-1 trie nodes with a direct second-byte table (GPT-2's table has three)
constexpr int DEC_TRIE_SM = 512;  // tries of up to this many edge slots / nodes are staged in shared memory

struct DecSm {
    uint4 rawbuf[(DEC_TILE + DEC_LOOK + 16 + 16 + 48) / 16];  // the tile at its global 16-byte phase + look-ahead
    uint16_t nxt[DEC_LEVELS][DEC_TILE];  // nxt[k][i] = position after 2^k steps from i; >= tile_len: left the tile
    uint8_t eb[DEC_TILE];                // byte emitted by a replacement / byte-mode step
    uint8_t el[DEC_TILE];                // 1, or 0x80 | n: copy n raw bytes (plain mode)
    uint32_t root_child[256];
    unsigned long long tedge[DEC_TRIE_SM];
    int16_t tout[DEC_TRIE_SM];
    uint8_t tdepth[DEC_TRIE_SM];
    uint8_t tkid[DEC_TRIE_SM];           // node has children
    uint8_t d1slot[DEC_TRIE_SM];         // 1 + row of pair2 for the children of the root, 0 otherwise
    uint16_t pair2[DEC_D1][256];         // child of a depth-1 node by the second byte: no hashing for 2-byte patterns
    uint32_t use_d1;
    uint4 outbuf[(DEC_TILE + 32) / 16];  // one-pass path: the tile's output at the 16-byte phase of its destination
    unsigned long long scan[32];
    unsigned long long tot;
    uint32_t entry;
};

__global__ void __launch_bounds__(DEC_NT) dec_doc_kernel(DecModel D, const int32_t* ids, const unsigned long long* tok_off,
                                                         unsigned long long ndocs, const unsigned long long* sub_off,
                                                         const uint8_t* raw, uint8_t* dec, uint32_t* doc_len,
                                                         unsigned long long* doc_raw) {
    __shared__ DecSm sm;
    const int tid = threadIdx.x;
    for (int i = tid; i < 256; i += DEC_NT) sm.root_child[i] = D.root_child[i];
    // small tries (every practical special-chars table) are walked in shared memory
    const bool trie_sm = D.trie_mask < (uint32_t)DEC_TRIE_SM;
    if (trie_sm) {
        for (uint32_t i = tid; i <= D.trie_mask; i += DEC_NT) sm.tedge[i] = D.trie_edge[i];
        for (uint32_t i = tid; i < (uint32_t)DEC_TRIE_SM; i += DEC_NT) {
            sm.tout[i] = i < D.trie_nodes ? D.trie_out[i] : (int16_t)-1;
            sm.tdepth[i] = i < D.trie_nodes ? D.trie_depth[i] : (uint8_t)0;
        }
    }
    const unsigned long long* const edge = trie_sm ? sm.tedge : D.trie_edge;
    for (uint32_t i = tid; i < (uint32_t)DEC_TRIE_SM; i += DEC_NT) {
        sm.tkid[i] = 0;
        sm.d1slot[i] = 0;
    }
    for (uint32_t i = tid; i < (uint32_t)DEC_D1 * 256u; i += DEC_NT) sm.pair2[i >> 8][i & 255u] = 0;
    __syncthreads();
    if (tid == 0) {
        uint32_t n1 = 0;
        bool ok = trie_sm;
        for (int b = 0; b < 256 && ok; ++b) {
            const uint32_t n = sm.root_child[b];
            if (!n || sm.d1slot[n]) continue;
            if (n1 == (uint32_t)DEC_D1) ok = false;
            else sm.d1slot[n] = (uint8_t)(++n1);
        }
        sm.use_d1 = ok ? 1u : 0u;
    }
    __syncthreads();
    if (sm.use_d1) {
        for (uint32_t i = tid; i <= D.trie_mask; i += DEC_NT) {
            const unsigned long long e = sm.tedge[i];
            if (e == HUT_TRIE_EMPTY) continue;
            const uint32_t key = (uint32_t)(e >> 32), parent = key >> 8;
            sm.tkid[parent] = 1;
            if (sm.d1slot[parent]) sm.pair2[sm.d1slot[parent] - 1][key & 255u] = (uint16_t)e;
        }
    }
    __syncthreads();
    const bool use_d1 = sm.use_d1 != 0;
    for (unsigned long long d = blockIdx.x; d < ndocs; d += gridDim.x) {
        // raw offset of id index t: the scanned sum of its sub-block + the lengths of the ids before it inside the
        // sub-block, added up by the whole CTA
        auto raw_at = [&](unsigned long long t) -> unsigned long long {
            unsigned long long part = 0;
            for (unsigned long long j = (t & ~(unsigned long long)(DEC_SUB - 1)) + tid; j < t; j += DEC_NT) {
                const uint32_t l = dec_token_len(D, ids[j]);
                if (l != HUT_DEC_MISSING) part += l;
            }
            block_exclusive_u64(part, &sm.tot, sm.scan);
            const unsigned long long r = sub_off[t / DEC_SUB] + sm.tot;
            __syncthreads();  // sm.tot is written again by the next call
            return r;
        };
        const unsigned long long r0 = raw_at(tok_off[d]), r1 = raw_at(tok_off[d + 1]);
        const unsigned long long rlen = r1 - r0;
        if (tid == 0) doc_raw[d] = r0;
        unsigned long long written = 0;
        // ---- prefix, stripped once when the raw text starts with it (pretokenizer.c:209-215) ----
        if (tid == 0) {
            uint32_t e = 0;
            if (D.prefix_len && rlen >= D.prefix_len) {
                bool same = true;
                for (uint32_t k = 0; k < D.prefix_len; ++k) same = same && raw[r0 + k] == D.prefix[k];
                if (same) e = D.prefix_len;
            }
            sm.entry = e;
        }
        __syncthreads();
        for (unsigned long long t0 = 0; t0 < rlen; t0 += DEC_TILE) {
            const uint32_t tile_len = rlen - t0 < DEC_TILE ? (uint32_t)(rlen - t0) : (uint32_t)DEC_TILE;
            const uint32_t avail = rlen - t0 < DEC_TILE + DEC_LOOK ? (uint32_t)(rlen - t0) : (uint32_t)(DEC_TILE + DEC_LOOK);
            // ---- stage the tile with 16-byte loads at its global phase; bytes past the document read as 0 ----
            const unsigned long long gbase = r0 + t0;
            const uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(raw) + gbase) & 15u);
            const uint4* gsrc = reinterpret_cast<const uint4*>(raw + gbase - mis);
            const uint32_t nwords = (mis + avail + 15u) >> 4;
            for (uint32_t w = tid; w < nwords; w += DEC_NT) sm.rawbuf[w] = gsrc[w];
            uint8_t* const rw = reinterpret_cast<uint8_t*>(sm.rawbuf) + mis;
            uint32_t entry = sm.entry;
            __syncthreads();
            if (avail < (uint32_t)(DEC_TILE + DEC_LOOK) && tid < 40) rw[avail + tid] = 0;
            __syncthreads();
            if (entry >= tile_len) {  // the previous step jumped over this whole tile (prefix / replacement at its end)
                if (tid == 0) sm.entry = entry - tile_len;
                __syncthreads();
                continue;
            }
            // ---- one-pass path (direct second-byte table): thread t owns positions [8t, 8t + 8) and keeps the step and
            //      the output of each in registers.  If every visited position is the start of a well-formed character
            //      and steps exactly over it (an ordinary character, or a replacement as long as its first
            //      character), the visited positions are the character starts from the entry on: one block scan of
            //      the emitted lengths places the bytes and nothing else is needed. ----
            if (use_d1) {
                const uint32_t p0 = (uint32_t)tid * 8u;
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(sm.rawbuf) + ((mis + p0) >> 2);
                const uint32_t sh = ((mis + p0) & 3u) * 8u;
                const uint32_t r0_ = wp[0], r1_ = wp[1], r2_ = wp[2], r3_ = wp[3];
                const uint32_t x[3] = {__funnelshift_r(r0_, r1_, sh), __funnelshift_r(r1_, r2_, sh),
                                       __funnelshift_r(r2_, r3_, sh)};
                auto B = [&](int k) -> uint32_t { return (x[k >> 2] >> ((k & 3) * 8)) & 0xFFu; };
                bool ok = true;
                uint32_t mine = 0;
                uint32_t outs[2] = {0u, 0u};  // byte k: the byte position p0 + k emits (replacement / byte mode)
                uint32_t emits = 0;           // nibble k: bytes position p0 + k emits; bit 3 = they are copied from the text
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t i = p0 + k;
                    const uint32_t c0 = B(k), c1 = B(k + 1), c2 = B(k + 2), c3 = B(k + 3), c4 = B(k + 4);
                    // branch-free classification: neighbouring lanes sit on characters of different classes
                    const bool live = (c0 & 0xC0u) != 0x80u && i >= entry && i < tile_len;
                    uint32_t cl = c0 < 0x80u ? 1u : ((0x43221111u >> ((c0 >> 2) & 0x1Cu)) & 7u);
                    if (c0 >= 0xF8u) cl = 1;  // not a lead byte: one byte, U+FFFD -> '?'
                    const uint32_t cont = ((c1 & 0xC0u) == 0x80u ? 1u : 0u) | ((c2 & 0xC0u) == 0x80u ? 2u : 0u) |
                                          ((c3 & 0xC0u) == 0x80u ? 4u : 0u) | ((c4 & 0xC0u) == 0x80u ? 8u : 0u);
                    // continuation bytes in place and the byte after the character is not one; three- and four-byte
                    // forms with a zero lead payload can be over-long encodings of code points below 256, which the
                    // reference decodes to that byte: left to the general path
                    bool wf = (cont & ((1u << cl) - 1u)) == (1u << (cl - 1u)) - 1u && !(cl >= 3u && (c0 & 0x0Fu) == 0u);
                    uint32_t out = c0 < 0x80u ? c0
                                              : (cl == 2u && (c0 & 0x1Cu) == 0u) ? (((c0 & 3u) << 6) | (c1 & 0x3Fu))
                                                                                 : (uint32_t)'?';  // pretokenizer.c:247-254
                    uint32_t step = cl, emit = D.byte_mode ? 1u : (8u | cl);  // plain mode: copied (pretokenizer.c:283-287)
                    const uint32_t node = live ? sm.root_child[c0] : 0u;
                    if (node) {  // longest replacement starting here: one or two bytes without hashing
                        int o = sm.tout[node];
                        if (o >= 0) { step = 1; out = (uint32_t)o; emit = 1; }
                        const uint32_t n2 = i + 1 < avail ? (uint32_t)sm.pair2[sm.d1slot[node] - 1][c1] : 0u;
                        if (n2) {
                            if (sm.tkid[n2]) wf = false;  // a longer replacement may start here: general path
                            else if ((o = sm.tout[n2]) >= 0) { step = 2; out = (uint32_t)o; emit = 1; }
                        }
                    }
                    ok = ok && (!live || (wf && step == cl));
                    if (live) {
                        outs[k >> 2] |= (out & 0xFFu) << ((k & 3) * 8);
                        emits |= emit << (4 * k);
                        mine += emit & 7u;
                    }
                }
                const bool one_pass = __syncthreads_and(ok && (rw[entry] & 0xC0u) != 0x80u) != 0;
                if (one_pass) {
                    const unsigned long long ex = block_exclusive_u64(mine, &sm.tot, sm.scan);
                    uint8_t* const dst = dec + r0 + written;
                    const uint32_t ph = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
                    uint8_t* ob = reinterpret_cast<uint8_t*>(sm.outbuf) + ph + (uint32_t)ex;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const uint32_t e = (emits >> (4 * k)) & 15u;
                        if (!e) continue;
                        if (e & 8u) {
                            for (uint32_t c = 0; c < (e & 7u); ++c) *ob++ = (uint8_t)B(k + (int)c);
                        } else {
                            *ob++ = (uint8_t)(outs[k >> 2] >> ((k & 3) * 8));
                        }
                    }
                    const uint32_t total = (uint32_t)sm.tot;
                    if (tid == 0) {  // hand-over: the first character start at or after the end of the tile
                        uint32_t k = 0;
                        while (k < 4 && (rw[tile_len + k] & 0xC0) == 0x80) ++k;
                        sm.entry = k;
                    }
                    __syncthreads();
                    // staged bytes -> global: 16-byte stores where the destination is aligned, bytes at the two ends
                    const uint32_t end = ph + total;
                    const uint8_t* const sb = reinterpret_cast<const uint8_t*>(sm.outbuf);
                    uint8_t* const gb = dst - ph;  // 16-byte aligned
                    for (uint32_t w = tid; w * 16u < end; w += DEC_NT) {
                        const uint32_t lo = w * 16u, hi = lo + 16u;
                        if (lo >= ph && hi <= end) {
                            reinterpret_cast<uint4*>(gb)[w] = sm.outbuf[w];
                        } else {
                            const uint32_t a = lo > ph ? lo : ph, b = hi < end ? hi : end;
                            for (uint32_t q = a; q < b; ++q) gb[q] = sb[q];
                        }
                    }
                    written += total;
                    __syncthreads();
                    continue;
                }
            }
            // ---- the step every position would take ----
            // `simple` stays true while every character start steps exactly over its own well-formed character: then
            // the positions visited are the character starts and no lifting is needed
            bool simple = true;
            for (uint32_t i = tid; i < tile_len; i += DEC_NT) {
                const uint32_t b0 = rw[i];
                uint32_t step = 0, out = 0, emit = 1;
                const uint32_t b1 = rw[i + 1], b2 = rw[i + 2], b3 = rw[i + 3];
                uint32_t node = sm.root_child[b0];
                if (node) {  // longest replacement starting here (never past the end of the document)
                    auto walk = [&](uint32_t nd, uint32_t q) {
                        for (;;) {
                            const int o = trie_sm ? (int)sm.tout[nd] : (int)__ldg(D.trie_out + nd);
                            if (o >= 0) {
                                step = trie_sm ? (uint32_t)sm.tdepth[nd] : (uint32_t)__ldg(D.trie_depth + nd);
                                out = (uint32_t)o;
                            }
                            if (q >= avail) break;
                            nd = trie_child(edge, D.trie_mask, nd, rw[q]);
                            if (!nd) break;
                            ++q;
                        }
                    };
                    if (use_d1) {  // one- and two-byte replacements without hashing
                        int o = sm.tout[node];
                        if (o >= 0) {
                            step = 1;
                            out = (uint32_t)o;
                        }
                        const uint32_t n2 = i + 1 < avail ? (uint32_t)sm.pair2[sm.d1slot[node] - 1][b1] : 0u;
                        if (n2) {
                            if (sm.tkid[n2]) {
                                walk(n2, i + 2);
                            } else if ((o = sm.tout[n2]) >= 0) {
                                step = 2;
                                out = (uint32_t)o;
                            }
                        }
                    } else {
                        walk(node, i + 1);
                    }
                }
                uint32_t cl, cp;
                if (b0 < 0x80) { cl = 1; cp = b0; }
                else if ((b0 & 0xE0) == 0xC0) { cl = 2; cp = ((b0 & 0x1F) << 6) | (b1 & 0x3F); }
                else if ((b0 & 0xF0) == 0xE0) { cl = 3; cp = ((b0 & 0x0F) << 12) | ((b1 & 0x3F) << 6) | (b2 & 0x3F); }
                else if ((b0 & 0xF8) == 0xF0) { cl = 4; cp = ((b0 & 0x07) << 18) | ((b1 & 0x3F) << 12) | ((b2 & 0x3F) << 6) | (b3 & 0x3F); }
                else { cl = 1; cp = 0xFFFD; }
                if (!step) {
                    step = cl;
                    if (D.byte_mode) {
                        out = cp < 256 ? cp : (uint32_t)'?';  // pretokenizer.c:247-254
                    } else {  // the character is copied (pretokenizer.c:283-287), clamped to the document
                        const unsigned long long left = rlen - t0 - i;
                        emit = left < cl ? (uint32_t)left : cl;
                        out = 0x100;  // marker: copy raw bytes
                    }
                }
                if ((b0 & 0xC0) != 0x80) {  // a character start
                    bool ok = step == cl && (rw[i + cl] & 0xC0) != 0x80;
                    if (cl > 1) ok = ok && (b1 & 0xC0) == 0x80;
                    if (cl > 2) ok = ok && (b2 & 0xC0) == 0x80;
                    if (cl > 3) ok = ok && (b3 & 0xC0) == 0x80;
                    simple = simple && ok;
                }
                sm.nxt[0][i] = (uint16_t)(i + step);
                sm.eb[i] = (uint8_t)out;
                sm.el[i] = (uint8_t)(out == 0x100 ? 0x80 | emit : 1);
            }
            simple = __syncthreads_and(simple && (rw[entry] & 0xC0) != 0x80) != 0;
            if (simple) {
                // ---- fast path: thread t owns positions [8t, 8t + 8); visited = character starts from the entry on ----
                const uint32_t p0 = (uint32_t)tid * (DEC_TILE / DEC_NT);
                uint32_t mine = 0;
#pragma unroll
                for (uint32_t k = 0; k < DEC_TILE / DEC_NT; ++k) {
                    const uint32_t i = p0 + k;
                    if (i >= entry && i < tile_len && (rw[i] & 0xC0) != 0x80) mine += sm.el[i] & 0x7Fu;
                }
                const unsigned long long ex = block_exclusive_u64(mine, &sm.tot, sm.scan);
                uint8_t* o = dec + r0 + written + ex;
#pragma unroll
                for (uint32_t k = 0; k < DEC_TILE / DEC_NT; ++k) {
                    const uint32_t i = p0 + k;
                    if (i >= entry && i < tile_len && (rw[i] & 0xC0) != 0x80) {
                        const uint32_t e = sm.el[i];
                        if (e & 0x80u) {
                            for (uint32_t c = 0; c < (e & 0x7Fu); ++c) *o++ = rw[i + c];
                        } else {
                            *o++ = sm.eb[i];
                        }
                    }
                }
                if (tid == 0) {  // hand-over: the first character start at or after the end of the tile
                    uint32_t k = 0;
                    while (k < 4 && (rw[tile_len + k] & 0xC0) == 0x80) ++k;
                    sm.entry = k;
                }
                written += sm.tot;
                __syncthreads();
                continue;
            }
            // ---- general path: binary lifting ----
            for (int k = 1; k < DEC_LEVELS; ++k) {
                for (uint32_t i = tid; i < tile_len; i += DEC_NT) {
                    const uint32_t v = sm.nxt[k - 1][i];
                    sm.nxt[k][i] = v >= tile_len ? (uint16_t)v : sm.nxt[k - 1][v];
                }
                __syncthreads();
            }
            // hops that stay inside the tile, counted from the entry
            uint32_t H = 0;
            {
                uint32_t pos = entry;
                for (int k = DEC_LEVELS - 1; k >= 0; --k) {
                    const uint32_t v = sm.nxt[k][pos];
                    if (v < tile_len) {
                        pos = v;
                        H += 1u << k;
                    }
                }
                if (tid == 0) sm.entry = sm.nxt[0][pos] - tile_len;  // hand-over to the next tile
            }
            const uint32_t nodes = H + 1;
            const uint32_t per = (nodes + DEC_NT - 1) / DEC_NT;
            const uint32_t h0 = tid * per;
            uint32_t pos = entry, mine = 0;
            if (h0 < nodes) {
                for (int k = DEC_LEVELS - 1; k >= 0; --k)
                    if (h0 & (1u << k)) pos = sm.nxt[k][pos];
                uint32_t p = pos;
                for (uint32_t h = h0; h < nodes && h < h0 + per; ++h) {
                    mine += sm.el[p] & 0x7Fu;
                    p = sm.nxt[0][p];
                }
            }
            const unsigned long long ex = block_exclusive_u64(mine, &sm.tot, sm.scan);
            if (h0 < nodes) {
                uint8_t* o = dec + r0 + written + ex;
                uint32_t p = pos;
                for (uint32_t h = h0; h < nodes && h < h0 + per; ++h) {
                    const uint32_t e = sm.el[p];
                    if (e & 0x80u) {
                        for (uint32_t k = 0; k < (e & 0x7Fu); ++k) *o++ = rw[p + k];
                    } else {
                        *o++ = sm.eb[p];
                    }
                    p = sm.nxt[0][p];
                }
            }
            written += sm.tot;
            __syncthreads();
        }
        if (tid == 0) doc_len[d] = (uint32_t)written;
        __syncthreads();
    }
}

// decoded bytes of each document -> packed text.  Warp per document; 16-byte stores at the destination's alignment,
// the source read as aligned 32-bit words and shifted into place (both ends of the document byte by byte)
__global__ void dec_pack_kernel(unsigned long long ndocs, const unsigned long long* doc_raw, const uint8_t* dec,
                                const unsigned long long* text_off, uint8_t* text) {
    const unsigned long long w = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const unsigned long long nw = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    for (unsigned long long d = w; d < ndocs; d += nw) {
        const uint8_t* src = dec + doc_raw[d];
        uint8_t* dst = text + text_off[d];
        const unsigned long long n = text_off[d + 1] - text_off[d];
        unsigned long long head = (16u - (reinterpret_cast<uintptr_t>(dst) & 15u)) & 15u;
        if (head > n) head = n;
        for (unsigned long long i = lane; i < head; i += 32) dst[i] = src[i];
        const uint8_t* s2 = src + head;
        const unsigned long long nwords = (n - head) >> 4;
        const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(s2) & 3u), sh = mis * 8u;
        const uint32_t* sw = reinterpret_cast<const uint32_t*>(s2 - mis);
        uint4* dw = reinterpret_cast<uint4*>(dst + head);
        for (unsigned long long q = lane; q < nwords; q += 32) {
            const uint32_t* p = sw + 4 * q;
            const uint32_t r0 = p[0], r1 = p[1], r2 = p[2], r3 = p[3], r4 = p[4];  // d_dec has 128 spare bytes
            dw[q] = make_uint4(__funnelshift_r(r0, r1, sh), __funnelshift_r(r1, r2, sh), __funnelshift_r(r2, r3, sh),
                               __funnelshift_r(r3, r4, sh));
        }
        for (unsigned long long i = head + (nwords << 4) + lane; i < n; i += 32) dst[i] = src[i];
    }
}

// ================================================================================================
// Token-wise path.  pretokenizer_decode (pretokenizer.c:197-296) scans the concatenated token strings greedily; the
// scan of a text is the scan of its tokens one after the other whenever no step straddles a token boundary, i.e.
// when, scanning each token ALONE, (a) every character is complete inside the token and (b) no walk down the
// replacement trie reaches the end of the token at a node that still has children (inside a text it would go on
// into the next token).  Every token of a GPT-2 / Llama style vocabulary is like that, so the decoded bytes of each
// token are worked out once per model (dec2_token_kernel, which restates the same scan for one token) and decoding
// becomes: lengths per sub-block of 256 ids -> scan -> copy of the tokens' decoded bytes to their final place, with
// text_off read off the same scan.  The prefix is stripped from the first token of a document by skipping the bytes
// its characters stand for.  A batch that uses a token without the property (T_UNSAFE / T_UNSAFE_START) goes
// through the general kernels above instead, so the results are the same by construction.
// ================================================================================================
__global__ void dec2_kids_kernel(const unsigned long long* edge, uint32_t cap, uint8_t* kids) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += gridDim.x * blockDim.x) {
        const unsigned long long e = edge[i];
        if (e != HUT_TRIE_EMPTY) kids[(uint32_t)(e >> 32) >> 8] = 1;
    }
}

// Thread per vocab line.  EMIT == false: t_meta[t] (flags, prefix bytes, decoded length).  EMIT == true: the decoded
// bytes at pool + off[t].
template <bool EMIT>
__global__ void dec2_token_kernel(DecModel D, const uint8_t* kids, uint32_t* meta, const uint32_t* off, uint8_t* pool) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= D.vocab_lines) return;
    const uint32_t len = D.dec_len[t];
    if (len == HUT_DEC_MISSING) {
        if (!EMIT) meta[t] = T_MISSING;
        return;
    }
    const uint8_t* const R = D.dec_pool + D.dec_off[t];
    uint8_t* const out = EMIT ? pool + off[t] : nullptr;
    auto B = [&](uint32_t i) -> uint32_t { return i < len ? (uint32_t)R[i] : 0u; };
    bool with_prefix = D.prefix_len && len >= D.prefix_len;
    if (with_prefix)
        for (uint32_t k = 0; k < D.prefix_len; ++k) with_prefix = with_prefix && R[k] == D.prefix[k];
    bool part_of_prefix = D.prefix_len && len < D.prefix_len;  // the prefix could continue in the next token
    if (part_of_prefix)
        for (uint32_t k = 0; k < len; ++k) part_of_prefix = part_of_prefix && R[k] == D.prefix[k];
    constexpr uint32_t NONE = 0xFFFFFFFFu;
    uint32_t p = 0, em = 0, ps = NONE;
    bool unsafe = false, soft = false;
    while (p < len) {
        if (with_prefix && p == D.prefix_len) ps = em;
        const uint32_t b0 = R[p];
        uint32_t step = 0, outb = 0, q = p + 1;
        uint32_t node = D.root_child[b0];
        while (node) {  // longest replacement starting here
            const int o = D.trie_out[node];
            if (o >= 0) {
                step = D.trie_depth[node];
                outb = (uint32_t)o;
            }
            if (q >= len) {
                if (kids[node]) soft = true;  // (b)
                break;
            }
            node = trie_child(D.trie_edge, D.trie_mask, node, R[q]);
            ++q;
        }
        if (step) {
            if (EMIT) out[em] = (uint8_t)outb;
            em += 1;
        } else {
            const uint32_t b1 = B(p + 1), b2 = B(p + 2), b3 = B(p + 3);
            uint32_t cl, cp;
            if (b0 < 0x80) { cl = 1; cp = b0; }
            else if ((b0 & 0xE0) == 0xC0) { cl = 2; cp = ((b0 & 0x1F) << 6) | (b1 & 0x3F); }
            else if ((b0 & 0xF0) == 0xE0) { cl = 3; cp = ((b0 & 0x0F) << 12) | ((b1 & 0x3F) << 6) | (b2 & 0x3F); }
            else if ((b0 & 0xF8) == 0xF0) { cl = 4; cp = ((b0 & 0x07) << 18) | ((b1 & 0x3F) << 12) | ((b2 & 0x3F) << 6) | (b3 & 0x3F); }
            else { cl = 1; cp = 0xFFFD; }
            if (p + cl > len) {  // (a)
                unsafe = true;
                cl = len - p;
            }
            step = cl;
            if (D.byte_mode) {
                if (EMIT) out[em] = (uint8_t)(cp < 256 ? cp : (uint32_t)'?');  // pretokenizer.c:247-254
                em += 1;
            } else {  // pretokenizer.c:283-287
                if (EMIT)
                    for (uint32_t k = 0; k < cl; ++k) out[em + k] = R[p + k];
                em += cl;
            }
        }
        p += step;
    }
    if (EMIT) return;
    if (with_prefix && p == D.prefix_len) ps = em;  // the token is the prefix
    uint32_t m = em & T_LEN_MASK;
    if (unsafe || em > T_LEN_MASK) m |= T_UNSAFE;
    else if (soft) m |= T_UNSAFE | T_SOFT;
    if (part_of_prefix || (with_prefix && (ps == NONE || ps > T_PS_MASK))) m |= T_UNSAFE_START;
    else if (with_prefix) m |= ps << T_PS_SHIFT;
    meta[t] = m;
}

// first_doc[b] = the first entry d of tok_off[0..ndocs] with tok_off[d] / DEC_SUB == b (0xFFFFFFFF, set beforehand: no
// document starts in the sub-block).  tok_off does not decrease, so exactly one entry per sub-block writes.
constexpr uint32_t DEC2_NO_DOC = 0xFFFFFFFFu;

__global__ void __launch_bounds__(256) dec2_first_doc_kernel(const unsigned long long* tok_off, unsigned long long ndocs,
                                                             uint32_t* first_doc) {
    const unsigned long long d = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (d > ndocs) return;
    const unsigned long long b = __ldg(tok_off + d) / DEC_SUB;
    if (d == 0 || __ldg(tok_off + d - 1) / DEC_SUB != b) first_doc[b] = (uint32_t)d;
}

// Token ids[i] is T_SOFT: scanning it alone, some walk down the trie stops at the end of the token with children left.
// Inside the document the walk goes on into the following tokens; the token still decodes as it does alone unless
// that finds a longer replacement.  Rare (a token that ends like the beginning of a special-chars pattern).
__device__ __noinline__ bool dec2_context_ok(const DecModel& D, const int32_t* ids, unsigned long long i,
                                             const unsigned long long* tok_off, unsigned long long ndocs) {
    unsigned long long lo = 0, hi = ndocs;  // first entry of tok_off beyond i = end of the token's document
    while (lo < hi) {
        const unsigned long long m = (lo + hi) >> 1;
        if (__ldg(tok_off + m) <= i) lo = m + 1;
        else hi = m;
    }
    const unsigned long long doc_end = __ldg(tok_off + lo);
    const int32_t id = ids[i];
    const uint32_t len = __ldg(D.dec_len + id);
    const uint8_t* const R = D.dec_pool + __ldg(D.dec_off + id);
    uint32_t p = 0;
    while (p < len) {
        const uint32_t b0 = R[p];
        uint32_t node = D.root_child[b0], step = 0, q = p + 1;
        while (node) {
            if (D.trie_out[node] >= 0) step = D.trie_depth[node];
            if (q >= len) break;
            node = trie_child(D.trie_edge, D.trie_mask, node, R[q]);
            ++q;
        }
        if (node && D.trie_kids[node]) {  // on into the document
            unsigned long long j = i + 1;
            const uint8_t* src = nullptr;
            uint32_t c = 0, l2 = 0;
            for (;;) {
                while (c >= l2 && j < doc_end) {
                    const int32_t id2 = ids[j++];
                    l2 = (uint32_t)id2 < D.vocab_lines ? __ldg(D.dec_len + id2) : (uint32_t)HUT_DEC_MISSING;
                    if (l2 == HUT_DEC_MISSING) return false;  // bad[0] is raised by the warp that owns the id
                    src = D.dec_pool + __ldg(D.dec_off + id2);
                    c = 0;
                }
                if (c >= l2) break;  // end of the document
                node = trie_child(D.trie_edge, D.trie_mask, node, src[c++]);
                if (!node) break;
                if (D.trie_out[node] >= 0) return false;
            }
        }
        if (!step) step = b0 < 0x80 ? 1u : (b0 & 0xE0) == 0xC0 ? 2u : (b0 & 0xF0) == 0xE0 ? 3u : (b0 & 0xF8) == 0xF0 ? 4u : 1u;
        p += step;
    }
    return true;
}

// Only for batches in which dec2_sub_kernel met T_SOFT tokens: each of them against the tokens that follow it.
__global__ void __launch_bounds__(256) dec2_soft_kernel(DecModel D, const int32_t* ids, unsigned long long n,
                                                        const unsigned long long* tok_off, unsigned long long ndocs,
                                                        unsigned int* bad) {
    const unsigned long long step = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
        const int32_t t = ids[i];
        if ((uint32_t)t >= D.vocab_lines || __ldg(D.t_len8 + t) != 255u) continue;
        const uint32_t m = __ldg(D.t_meta + t);
        if ((m & T_UNSAFE) && (m & T_SOFT) && !dec2_context_ok(D, ids, i, tok_off, ndocs)) atomicExch(bad + 1, 1u);
    }
}

__device__ __forceinline__ void dec2_load_ids(const int32_t* ids, unsigned long long i0, unsigned long long n,
                                              int32_t (&id)[8]) {
    if (i0 + 8 <= n && (reinterpret_cast<uintptr_t>(ids) & 15u) == 0) {
        const int4 a = __ldg(reinterpret_cast<const int4*>(ids + i0)), b = __ldg(reinterpret_cast<const int4*>(ids + i0) + 1);
        id[0] = a.x; id[1] = a.y; id[2] = a.z; id[3] = a.w;
        id[4] = b.x; id[5] = b.y; id[6] = b.z; id[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) id[k] = i0 + k < n ? ids[i0 + k] : 0;
    }
}

// Warp per sub-block of DEC_SUB ids: decoded bytes of the sub-block.  bad[0]: an id outside the vocabulary
// (core.c:520-531); bad[1]: the batch needs the general kernels; bad[2]: T_SOFT tokens seen.  PFX: the bytes the prefix stands for come off the
// first token of every document that starts in the sub-block.
template <bool PFX>
__global__ void __launch_bounds__(256) dec2_sub_kernel(DecModel D, const int32_t* ids, unsigned long long n,
                                                       const unsigned long long* tok_off, unsigned long long ndocs,
                                                       const uint32_t* first_doc, uint32_t* sub_len,
                                                       unsigned int* bad) {
    const int lane = threadIdx.x & 31;
    const unsigned long long blk = (unsigned long long)blockIdx.x * 8u + (threadIdx.x >> 5);
    if (blk == 0 && lane == 0 && (tok_off[0] != 0 || tok_off[ndocs] != n)) atomicExch(bad + 1, 1u);  // ids outside every document
    const unsigned long long b0 = blk * DEC_SUB;
    if (b0 >= n) return;
    const unsigned long long i0 = b0 + (unsigned long long)lane * 8u;
    int32_t id[8];
    dec2_load_ids(ids, i0, n, id);
    uint32_t l[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) l[k] = (i0 + k < n && (uint32_t)id[k] < D.vocab_lines) ? __ldg(D.t_len8 + id[k]) : 256u;
    uint32_t sum = 0;
    unsigned int flag = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (i0 + k >= n) continue;
        if (l[k] < 255u) {
            sum += l[k];
        } else if (l[k] == 256u) {
            flag |= 1u;
        } else {
            const uint32_t m = __ldg(D.t_meta + id[k]);
            if (m & T_MISSING) flag |= 1u;
            if (m & T_UNSAFE) flag |= m & T_SOFT ? 4u : 2u;
            sum += m & T_LEN_MASK;
        }
    }
    if (PFX) {
        const unsigned long long b1 = b0 + DEC_SUB < n ? b0 + DEC_SUB : n;
        const uint32_t d_lo = __ldg(first_doc + blk);
        for (unsigned long long base = d_lo; d_lo != DEC2_NO_DOC; base += 32) {
            const unsigned long long d = base + lane;
            bool in = d < ndocs;
            unsigned long long t = 0;
            if (in) {
                t = __ldg(tok_off + d);
                in = t < b1;
            }
            if (in && __ldg(tok_off + d + 1) > t) {
                const int32_t f = ids[t];
                if ((uint32_t)f < D.vocab_lines) {
                    const uint32_t m = __ldg(D.t_meta + f);
                    if (m & T_UNSAFE_START) flag |= 2u;
                    sum -= (m >> T_PS_SHIFT) & T_PS_MASK;
                }
            }
            if (!__all_sync(0xFFFFFFFFu, in)) break;
        }
    }
    if (flag & 1u) atomicExch(bad, 1u);
    if (flag & 2u) atomicExch(bad + 1, 1u);
    if (flag & 4u) atomicExch(bad + 2, 1u);  // dec2_soft_kernel decides
#pragma unroll
    for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, d);
    if (lane == 0) sub_len[blk] = sum;
}

// Warp per sub-block (n / DEC_SUB + 1 of them: document boundaries at the very end belong to the last one).  Every
// lane takes 8 consecutive ids and loads their 16-byte slots at once, the warp scans the lengths, the bytes are put
// together in shared memory at the 16-byte phase of their destination and leave in 16-byte stores; text_off of the
// documents that start inside the sub-block comes from the same scan.
constexpr int G2_WARPS = 8;
#ifndef G2_CTAS
#define G2_CTAS 4
#endif
constexpr int G2_STAGE = 2048;  // bytes of one sub-block that fit the staging buffer; larger ones are written directly

struct G2Warp {
    uint4 stage[(G2_STAGE + 48) / 16];
    uint32_t off[DEC_SUB + 4];   // exclusive scan of the lengths inside the sub-block, [DEC_SUB] = total
    uint32_t skip[DEC_SUB / 4];  // PFX: bytes dropped from the front of each token (one byte per id)
};

__device__ __forceinline__ void dec2_put4(uint8_t* p, uint32_t w, uint32_t cnt) {
    p[0] = (uint8_t)w;
    if (cnt > 1) p[1] = (uint8_t)(w >> 8);
    if (cnt > 2) p[2] = (uint8_t)(w >> 16);
    if (cnt > 3) p[3] = (uint8_t)(w >> 24);
}

template <bool PFX>
__global__ void __launch_bounds__(G2_WARPS * 32, G2_CTAS) dec2_gather_kernel(DecModel D, const int32_t* ids, unsigned long long n,
                                                                    const unsigned long long* tok_off,
                                                                    unsigned long long ndocs,
                                                                    const uint32_t* first_doc,
                                                                    const unsigned long long* sub_off, uint8_t* text,
                                                                    unsigned long long* text_off) {
    __shared__ G2Warp sm[G2_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long blk = (unsigned long long)blockIdx.x * G2_WARPS + warp;
    if (blk > n / DEC_SUB) return;  // warps never synchronise with each other
    G2Warp& W = sm[warp];
    const unsigned long long b0 = blk * DEC_SUB, i0 = b0 + (unsigned long long)lane * 8u;
    int32_t id[8];
    dec2_load_ids(ids, i0, n, id);
    uint4 s[8];
    uint32_t l[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const bool ok = i0 + k < n && (uint32_t)id[k] < D.vocab_lines;
        s[k] = ok ? __ldg(D.t_slot + id[k]) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        l[k] = s[k].w >> 24;
        if (l[k] == 255u) l[k] = __ldg(D.t_meta + id[k]) & T_LEN_MASK;
    }
    const uint32_t d_lo = __ldg(first_doc + blk);
    uint32_t sk[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) sk[k] = 0;
    if (PFX) {
        W.skip[lane * 2] = 0;
        W.skip[lane * 2 + 1] = 0;
        __syncwarp();
        const unsigned long long b1 = b0 + DEC_SUB < n ? b0 + DEC_SUB : n;
        for (unsigned long long base = d_lo; d_lo != DEC2_NO_DOC; base += 32) {
            const unsigned long long d = base + lane;
            bool in = d < ndocs;
            unsigned long long t = 0;
            if (in) {
                t = __ldg(tok_off + d);
                in = t < b1;
            }
            if (in && __ldg(tok_off + d + 1) > t) {
                const int32_t f = ids[t];
                if ((uint32_t)f < D.vocab_lines)
                    reinterpret_cast<uint8_t*>(W.skip)[t - b0] = (uint8_t)((__ldg(D.t_meta + f) >> T_PS_SHIFT) & T_PS_MASK);
            }
            if (!__all_sync(0xFFFFFFFFu, in)) break;
        }
        __syncwarp();
        const uint32_t w0 = W.skip[lane * 2], w1 = W.skip[lane * 2 + 1];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            sk[k] = ((k < 4 ? w0 : w1) >> ((k & 3) * 8)) & 0xFFu;
            l[k] -= sk[k];
        }
    }
    uint32_t mine = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) mine += l[k];
    uint32_t inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += t;
    }
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, inc, 31), before = inc - mine;
    const unsigned long long g0 = __ldg(sub_off + blk);
    if (d_lo != DEC2_NO_DOC) {
        uint32_t run = before;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            W.off[lane * 8 + k] = run;
            run += l[k];
        }
        if (lane == 31) W.off[DEC_SUB] = run;
        __syncwarp();
    }
    for (unsigned long long base = d_lo; d_lo != DEC2_NO_DOC; base += 32) {  // documents (and the end mark) that start here
        const unsigned long long d = base + lane;
        bool in = d <= ndocs;
        unsigned long long t = 0;
        if (in) {
            t = __ldg(tok_off + d);
            in = t < b0 + DEC_SUB;
        }
        if (in) text_off[d] = g0 + W.off[t - b0];
        if (!__all_sync(0xFFFFFFFFu, in)) break;
    }
    uint8_t* const gdst = text + g0;
    const bool staged = total <= (uint32_t)G2_STAGE;
    const uint32_t ph = (uint32_t)(reinterpret_cast<uintptr_t>(gdst) & 15u);
    uint8_t* sp = staged ? reinterpret_cast<uint8_t*>(W.stage) + ph + before : gdst + before;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t L = l[k];
        if (!L) continue;
        if (L + sk[k] <= 15u) {
            uint4 v = s[k];
            if (PFX)
                for (uint32_t c = 0; c < sk[k]; ++c) {
                    v.x = __funnelshift_r(v.x, v.y, 8);
                    v.y = __funnelshift_r(v.y, v.z, 8);
                    v.z = __funnelshift_r(v.z, v.w, 8);
                    v.w >>= 8;
                }
            dec2_put4(sp, v.x, L);
            if (L > 4) dec2_put4(sp + 4, v.y, L - 4);
            if (L > 8) dec2_put4(sp + 8, v.z, L - 8);
            if (L > 12) dec2_put4(sp + 12, v.w, L - 12);
        } else {
            const uint8_t* src = D.t_pool + __ldg(D.t_off + id[k]) + sk[k];
            for (uint32_t c = 0; c < L; ++c) sp[c] = __ldg(src + c);
        }
        sp += L;
    }
    if (!staged) return;
    __syncwarp();
    const uint32_t end = ph + total;
    const uint8_t* const sb = reinterpret_cast<const uint8_t*>(W.stage);
    uint8_t* const gb = gdst - ph;  // 16-byte aligned
    for (uint32_t w = lane; w * 16u < end; w += 32) {
        const uint32_t lo = w * 16u, hi = lo + 16u;
        if (lo >= ph && hi <= end) {
            reinterpret_cast<uint4*>(gb)[w] = W.stage[w];
        } else {  // a 16-byte word shared with the neighbouring sub-block: whole 32-bit words, then single bytes
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) {
                const uint32_t l4 = lo + 4u * j, h4 = l4 + 4u;
                if (l4 >= ph && h4 <= end) {
                    reinterpret_cast<uint32_t*>(gb)[l4 >> 2] = reinterpret_cast<const uint32_t*>(sb)[l4 >> 2];
                } else {
                    const uint32_t a = l4 > ph ? l4 : ph, b = h4 < end ? h4 : end;
                    for (uint32_t q = a; q < b; ++q) gb[q] = sb[q];
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// One-launch path for a few ids (hutoken.decode of a sentence): one CTA does the whole token-wise pipeline on
// pinned host buffers the device reads and writes directly (ids, tok_off in; text, text_off, header out) and raises
// `done` last, after a system-wide fence; the host spins on it.  Anything it cannot finish -- an unsafe or soft
// token, more text than the staging buffer -- is reported in hdr->flags and the call goes through dec_run.
// ------------------------------------------------------------------------------------------------
constexpr int DS_NT = 256;
constexpr int DS_MAX_IDS = 2048;
constexpr int DS_MAX_DOCS = 256;
constexpr int DS_TEXT = 32768;  // decoded bytes staged in shared memory
constexpr unsigned int DS_BAD_ID = 1u, DS_FALLBACK = 2u;

struct DecSmallHdr {
    unsigned long long total;
    unsigned int flags;
    unsigned int pad;
};

template <bool PFX>
__global__ void __launch_bounds__(DS_NT, 1) dec_small_kernel(DecModel D, const int32_t* ids, uint32_t n,
                                                             const unsigned long long* tok_off, uint32_t ndocs,
                                                             uint8_t* text, unsigned long long* text_off,
                                                             DecSmallHdr* hdr, volatile unsigned int* done,
                                                             unsigned int seq) {
    __shared__ uint4 stage[DS_TEXT / 16];
    __shared__ uint32_t off[DS_MAX_IDS + 1];
    __shared__ uint8_t skip[DS_MAX_IDS];
    __shared__ unsigned long long scan_sm[32];
    __shared__ unsigned long long tot;
    const int tid = threadIdx.x;
    constexpr int PER = DS_MAX_IDS / DS_NT;
    unsigned int flags = 0;
    if (PFX) {
        for (int i = tid; i < DS_MAX_IDS; i += DS_NT) skip[i] = 0;
        __syncthreads();
        for (uint32_t d = tid; d < ndocs; d += DS_NT) {
            const unsigned long long t = tok_off[d];
            if (tok_off[d + 1] > t) {
                const int32_t f = ids[t];
                if ((uint32_t)f < D.vocab_lines) {
                    const uint32_t m = __ldg(D.t_meta + f);
                    if (m & T_UNSAFE_START) flags |= DS_FALLBACK;
                    skip[t] = (uint8_t)((m >> T_PS_SHIFT) & T_PS_MASK);
                }
            }
        }
        __syncthreads();
    }
    int32_t id[PER];
    uint32_t l[PER];
    uint32_t mine = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const uint32_t i = (uint32_t)tid * PER + k;
        id[k] = i < n ? ids[i] : 0;
        l[k] = 0;
        if (i < n) {
            if ((uint32_t)id[k] >= D.vocab_lines) {
                flags |= DS_BAD_ID;
            } else {
                const uint32_t m = __ldg(D.t_meta + id[k]);
                if (m & T_MISSING) flags |= DS_BAD_ID;
                if (m & T_UNSAFE) flags |= DS_FALLBACK;
                l[k] = (m & T_LEN_MASK) - (PFX ? (uint32_t)skip[i] : 0u);
            }
        }
        mine += l[k];
    }
    const unsigned long long ex = block_exclusive_u64(mine, &tot, scan_sm);
    const unsigned int any = (__syncthreads_or((int)(flags & DS_BAD_ID)) ? DS_BAD_ID : 0u) |
                             (__syncthreads_or((int)(flags & DS_FALLBACK)) ? DS_FALLBACK : 0u);
    const unsigned long long total = tot;
    const bool give_up = (any & (DS_BAD_ID | DS_FALLBACK)) || total > (unsigned long long)DS_TEXT;
    if (!give_up) {
        uint32_t run = (uint32_t)ex;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const uint32_t i = (uint32_t)tid * PER + k;
            if (i <= n) off[i] = run;
            if (i < n && l[k]) {
                const uint8_t* src = D.t_pool + __ldg(D.t_off + id[k]) + (PFX ? (uint32_t)skip[i] : 0u);
                uint8_t* dst = reinterpret_cast<uint8_t*>(stage) + run;
                for (uint32_t c = 0; c < l[k]; ++c) dst[c] = __ldg(src + c);
            }
            run += l[k];
        }
        if (tid == DS_NT - 1 && n == (uint32_t)DS_MAX_IDS) off[n] = run;
        __syncthreads();
        for (uint32_t d = tid; d <= ndocs; d += DS_NT) text_off[d] = off[tok_off[d]];
        const uint32_t words = ((uint32_t)total + 15u) >> 4;  // the host buffer is 16-byte aligned with room to spare
        for (uint32_t w = tid; w < words; w += DS_NT) reinterpret_cast<uint4*>(text)[w] = stage[w];
    }
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
        hdr->total = total;
        hdr->flags = (any & DS_BAD_ID) | (give_up && !(any & DS_BAD_ID) ? DS_FALLBACK : 0u);
        __threadfence_system();
        *done = seq;
    }
}

}  // namespace hut
