/* sm_100a device code of the BPE encode engine, part 2: the single-pass tile kernel.  Included only by hut_engine.cu.
 *
 *   tile2_kernel    persistent, WARP per 960-byte tile, tiles claimed in order from an atomic ticket:
 *                     TMA bulk copy of the tile (+ halos) into the warp's shared-memory slice
 *                     bit-sliced class masks + local start rule (hut_bitslice.h)   <- parser_next_token (parser.c:24-88)
 *                     rounds of 32 pretokens, one per lane: probe of the unique-pretoken cache (one 32-byte sector)
 *                     cache misses of up to 32 symbols: merged right there by the whole warp, one symbol per lane
 *                                                       <- bpe_encode_arena_string (core.c:66-209), MinPQ (queue.c)
 *                     ids of the tile are put together in the warp's shared-memory slice
 *                     decoupled look-back over the per-tile id counts  <- IntVector growth (vector.c:41-55)
 *                     ids leave once, coalesced, at their final position  <- list building (lib.c:699-715, 810-860)
 *
 * Nothing is staged in global memory and no second pass reads the ids again.  The only pretokens the kernel does not
 * finish itself are those of more than 32 bytes / symbols (or running past the tile's halo): they go to a side list
 * (DefEntry), the ids around them are written as if they produced nothing ("pre-patch" positions), and the host runs
 * the patch phase (long_kernel + patch kernels, hut_pipeline.cuh) only when the list is not empty.
 */
#pragma once

#include "hut_kernels.cuh"

#ifndef HUT_TILE_CTAS
#define HUT_TILE_CTAS 4  // resident CTAs per SM the tile kernel is compiled for (64 registers per thread)
#endif

namespace hut {

constexpr int IDCAP = 512;            // ids of a tile kept in shared memory before they leave (a fuller tile flushes early)
constexpr int LANE_IDS_MAX = 40;      // most ids one pretoken merged in the tile kernel can produce (32 symbols + prefix ids)
constexpr unsigned long long ST_MASK = (1ull << 62) - 1;
constexpr unsigned long long ST_AGG = 1ull << 62;   // status: the tile's own id count is known
constexpr unsigned long long ST_PFX = 2ull << 62;   // status: the inclusive prefix (ids of tiles 0..t) is known

// shared memory of ONE WARP of tile2_kernel
struct __align__(16) WarpSm2 {
    uint8_t text[TEXTB];        // span bytes: [tile_start - 32, tile_start + 992 + 16), then zero padding
    uint32_t dsm[CHUNK + 4];    // per chunk: document starts
    uint16_t pstart[SPAN + 8];  // span positions of the boundaries in chunks 1..31 (31 = halo), then the end sentinel
    int32_t ids[IDCAP];         // ids of the tile, in order
    uint32_t sc[64];            // symbols of the cache miss being merged
    unsigned long long mbar;
};

struct TileSm2 {
    uint4 kmask[16];            // kmask[len]: byte masks of a 16-byte key holding len text bytes
    uint4 kmask32[64];          // kmask32[2 * len], [2 * len + 1]: the same for the 32-byte keys of the second cache
    uint32_t byte1[256];
    WarpSm2 w[WPB];
};

// scalar restatement (hut_classify.h) of one 32-byte chunk: only used when an over-long UTF-8 form is in the tile
__device__ __noinline__ void generic_chunk(const uint8_t* text, const uint32_t* dsm, int lane, uint32_t* ST, uint32_t* NU) {
    auto dsbit = [&](int x) -> uint32_t { return (x >= 0 && x < SPAN) ? ((dsm[x >> 5] >> (x & 31)) & 1u) : 0u; };
    auto lead = [&](int x) -> uint32_t {
        uint32_t b0 = text[x], b1 = text[x + 1], b2 = text[x + 2], b3 = text[x + 3];
        if (dsbit(x + 1)) b1 = b2 = b3 = 0;
        else if (dsbit(x + 2)) b2 = b3 = 0;
        else if (dsbit(x + 3)) b3 = 0;
        return hut_lead_class(b0, b1, b2, b3);
    };
    auto code = [&](int x) -> uint32_t {
        if (x < 0) return HC_END;
        const uint32_t b0 = text[x];
        if ((b0 & 0xC0u) == 0x80u) {
            for (int k = 1; k <= 3; ++k) {
                const int q = x - k;
                if (q < 0) break;
                if ((text[q] & 0xC0u) == 0x80u) continue;
                if ((lead(q) >> 4) > (uint32_t)k) return HC_CONT;
                break;
            }
            return HC_BAD;
        }
        return lead(x) & 15u;
    };
    uint32_t st = 0, nu = 0;
    for (int j = 0; j < 32; ++j) {
        const int x = lane * 32 + j;
        const uint32_t c = code(x);
        if (c == HC_END) {
            nu |= 1u << j;
            continue;
        }
        if (c == HC_CONT) continue;
        const uint32_t p1 = code(x - 1);
        uint32_t pc = p1;
        for (int k = 2; pc == HC_CONT && k <= 4; ++k) pc = code(x - k);
        const int lone = dsbit(x - 1) || code(x - 2) != HC_SPACE;
        if (hut_is_start(c, (int)dsbit(x), p1, pc, lone)) st |= 1u << j;
    }
    *ST = st;
    *NU = nu;
}

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Decoupled look-back: ids produced by tiles 0 .. tile-1.  Lane i looks at tile (hi - i); the window moves back by 32
// until a tile with a known inclusive prefix is found.  Status words carry flag and value together, so one 64-bit load
// is a consistent snapshot.  Tiles are claimed in order from the ticket, so every tile waited for is being worked on.
__device__ __noinline__ unsigned long long lookback(const unsigned long long* status, uint32_t tile) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t FULL = 0xFFFFFFFFu;
    unsigned long long sum = 0;
    long long hi = (long long)tile - 1;
    for (;;) {
        const long long t = hi - (long long)lane;
        const unsigned long long v = t < 0 ? ST_PFX : ld_relaxed_u64(status + t);
        const uint32_t flag = (uint32_t)(v >> 62);
        const uint32_t pm = __ballot_sync(FULL, flag == 2u);
        const uint32_t zm = __ballot_sync(FULL, flag == 0u);
        const int jp = pm ? __ffs((int)pm) - 1 : 32;
        const uint32_t upto = jp >= 31 ? FULL : ((2u << jp) - 1u);  // lanes 0..jp
        if (zm & upto) {
            __nanosleep(40);
            continue;
        }
        unsigned long long x = ((upto >> lane) & 1u) ? (v & ST_MASK) : 0ull;
#pragma unroll
        for (int d = 16; d; d >>= 1) x += __shfl_xor_sync(FULL, x, d);
        sum += x;
        if (jp < 32) return sum;
        hi -= 32;
    }
}

// appends a deferred pretoken to the global list and links it to its tile's chain (one lane calls)
__device__ __forceinline__ void defer_push2(const Job& J, uint32_t tile, unsigned long long pos, uint32_t local_off,
                                            uint32_t len, uint32_t first, uint32_t& head, uint32_t& tail) {
    const uint32_t idx = atomicAdd(&J.ctr->n_def, 1u);
    if (idx >= J.def_cap) {
        set_flag(J, FLAG_DEF_CAP);
        return;
    }
    DefEntry e;
    e.pos = pos;
    e.tile = tile;
    e.local_off = local_off;
    e.len = len;
    e.flags = first | (head == NODOC ? 4u : 0u);  // bit 2: first entry of its tile's chain
    e.count = 0;
    e.pad = NODOC;  // next entry of the same tile
    e.ids_off = 0;
    J.def[idx] = e;
    if (head == NODOC) head = idx;
    else J.def[tail].pad = idx;
    tail = idx;
}

// ------------------------------------------------------------------------------------------------
// One cache miss merged by the whole warp: pretoken [s, s + len) of the span in shared memory, one symbol per lane.
// Per merge: redux.min over the pair ranks + ballot for the leftmost minimum -- the (rank, left index) order of the
// reference's heap (queue.c:162-164) -- one shuffle to close the gap, two pair-table probes for the new neighbours.
// Replaces bpe_encode_arena_string (core.c:66-209) for the pretokens the cache has not seen.
// The ids (prefix ids first, core.c:421-446) are written to dst[0 .. total); returns total, or 0xFFFFFFFF when the
// pretoken has more than 32 symbols (the caller defers it to long_kernel).  All lanes call with the same arguments.
// ------------------------------------------------------------------------------------------------
__device__ __noinline__ uint32_t wmerge_one(const DevModel& M, const uint32_t* byte1, const uint8_t* text, uint32_t* sc,
                                            int s, int len, uint32_t first, int32_t* dst) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t FULL = 0xFFFFFFFFu;
    auto at = [&](unsigned long long p) -> uint32_t { return (uint32_t)text[p]; };
    const uint32_t b0 = text[s];
    const bool with_prefix = first && b0 != ' ';
    const bool alone_first = first && b0 == ' ';
    const uint32_t npfx = with_prefix ? M.n_prefix_syms : 0u;
    const uint32_t nal = alone_first ? M.n_prefix_alone : 0u;
    // ---- initial symbols (A.3): lane p looks at text byte p ----
    uint32_t unit = M.inert, pb = 0, cnt = 0;
    if ((int)lane < len)
        cnt = char_symbols(M, byte1, at, (unsigned long long)(s + (int)lane), (unsigned long long)s,
                           (unsigned long long)(s + len), &unit, &pb);
    uint32_t inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, inc, d);
        if (lane >= (uint32_t)d) inc += t;
    }
    const uint32_t nsym = __shfl_sync(FULL, inc, 31);
    uint32_t m = npfx + nsym;
    if (m > 32u || nal > 8u) return 0xFFFFFFFFu;
    __syncwarp();
    if (lane < npfx) sc[lane] = M.prefix_syms[lane];
    {
        const uint32_t o = npfx + inc - cnt;
        if (cnt == 1) sc[o] = unit;
        else
            for (uint32_t k = 0; k < cnt; ++k) sc[o + k] = __ldg(M.sym_pool + pb + k);
    }
    __syncwarp();
    uint32_t u = lane < m ? sc[lane] : NONE;
    __syncwarp();
    // ---- merge loop (A.4) ----
    {
        const uint32_t un = __shfl_down_sync(FULL, u, 1);
        uint32_t mg = (lane + 1 < m) ? pair_probe(M, u, un) : NONE;
        uint32_t rk = rank_of_unit(M, mg);
        for (;;) {
            const uint32_t rmin = __reduce_min_sync(FULL, rk);
            if (rmin == 0xFFFFFFFFu) break;
            const uint32_t j = (uint32_t)__ffs((int)__ballot_sync(FULL, rk == rmin)) - 1u;
            const uint32_t mj = __shfl_sync(FULL, mg, (int)j);
            // lane j takes the merged unit; lanes > j take their right neighbour's state
            const uint32_t u_r = __shfl_down_sync(FULL, u, 1);
            const uint32_t mg_r = __shfl_down_sync(FULL, mg, 1);
            const uint32_t rk_r = __shfl_down_sync(FULL, rk, 1);
            if (lane == j) u = mj;
            if (lane > j) {
                u = u_r;
                mg = mg_r;
                rk = rk_r;
            }
            m--;
            if (lane >= m) {
                u = NONE;
                mg = NONE;
                rk = 0xFFFFFFFFu;
            }
            // only the pairs touching the new token change: (j-1, j) and (j, j+1)
            const uint32_t nb = __shfl_down_sync(FULL, u, 1);
            if (lane == j || lane + 1 == j) {
                mg = (lane + 1 < m) ? pair_probe(M, u, nb) : NONE;
                rk = rank_of_unit(M, mg);
            }
        }
    }
    if (lane < nal) dst[lane] = M.prefix_alone[lane];
    if (lane < m) dst[nal + lane] = id_of_unit(M, u);
    __syncwarp();
    return nal + m;
}

// PFX: the model has a prefix (the first pretoken of a document is keyed apart and merged with the prefix symbols).
template <bool PFX>
__global__ void __launch_bounds__(NT, HUT_TILE_CTAS) tile2_kernel(const __grid_constant__ DevModel M,
                                                                   const __grid_constant__ Job J) {
    __shared__ TileSm2 sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t FULL = 0xFFFFFFFFu;
    WarpSm2& ws = sm.w[warp];

    for (int i = tid; i < 256; i += NT) sm.byte1[i] = M.byte1[i];
    if (tid < 16) {
        auto wm = [&](int rem) -> uint32_t { return rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : ((1u << (8 * rem)) - 1u)); };
        sm.kmask[tid] = make_uint4(wm(tid), wm(tid - 4), wm(tid - 8), wm(tid - 12));
    }
    if (tid < 64) {
        const int len = tid >> 1, o = (tid & 1) * 16;
        auto wm = [&](int rem) -> uint32_t { return rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : ((1u << (8 * rem)) - 1u)); };
        sm.kmask32[tid] = make_uint4(wm(len - o), wm(len - o - 4), wm(len - o - 8), wm(len - o - 12));
    }
    if (lane == 0) {
        mbar_init(&ws.mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();  // the only block-wide barrier: the tables above are read-only afterwards

    uint32_t parity = 0;
    unsigned long long st_pre = 0, st_miss = 0, st_def = 0;  // lane 0 counts
    uint32_t tile = 0;
    if (lane == 0) tile = atomicAdd(&J.ctr->ticket, 1u);
    tile = __shfl_sync(FULL, tile, 0);

    while (tile < J.ntiles) {
        const long long g0 = (long long)tile * OWN - BACK;  // global position of span index 0
        const long long lo = g0 < 0 ? 0 : g0;
        const long long hi = (g0 + LOADB) < (long long)J.nbytes ? (g0 + LOADB) : (long long)J.nbytes;
        const int lo_li = (int)(lo - g0), hi_li = (int)(hi - g0);
        const uint32_t body = (uint32_t)(hi_li - lo_li) & ~15u;

        // ---- stage the span: one TMA bulk copy for the 16-byte-aligned body, plain loads for the ragged rest ----
        __syncwarp();
        if (lane == 0 && body) {
            fence_proxy_async();
            mbar_expect_tx(&ws.mbar, body);
            tma_load_1d(ws.text + lo_li, J.text + lo, body, &ws.mbar);
        }
        uint32_t next_tile = 0;  // claimed now, used at the end of the loop: the atomic's latency is hidden
        if (lane == 0) next_tile = atomicAdd(&J.ctr->ticket, 1u);
        for (int li = lane; li < lo_li; li += 32) ws.text[li] = 0;
        for (int li = lo_li + (int)body + lane; li < TEXTB; li += 32) ws.text[li] = li < hi_li ? J.text[g0 + li] : 0;
        ws.dsm[lane] = 0;
        if (lane < 4) ws.dsm[CHUNK + lane] = 0;
        // ---- document starts inside the span ----
        uint32_t fd = NODOC;
        if (lane < 3) {
            const long long tt = (long long)tile - 1 + lane;
            if (tt >= 0 && tt < (long long)J.ntiles) fd = J.first_doc[tt];
        }
        const uint32_t f0 = __shfl_sync(FULL, fd, 0), dfirst = __shfl_sync(FULL, fd, 1), f2 = __shfl_sync(FULL, fd, 2);
        const uint32_t dscan = f0 != NODOC ? f0 : (dfirst != NODOC ? dfirst : f2);
        __syncwarp();
        if (dscan != NODOC) {
            const long long lim = g0 + SPAN;
            for (unsigned long long d = (unsigned long long)dscan + lane;; d += 32) {
                const long long pos = d < J.ndocs ? (long long)J.doc_off[d] : lim;
                const bool in = pos < lim;
                if (in && pos >= g0) {
                    const int li = (int)(pos - g0);
                    atomicOr(&ws.dsm[li >> 5], 1u << (li & 31));
                }
                if (!__all_sync(FULL, in)) break;
            }
        }
        if (body) {
            mbar_wait(&ws.mbar, parity);
            parity ^= 1u;
        }
        __syncwarp();

        // ---- classify: bit planes -> class masks -> pretoken starts (hut_bitslice.h) ----
        uint32_t ST, NU;
        bool tile_docs;
        {
            const uint4* tp = reinterpret_cast<const uint4*>(ws.text + lane * CHUNK);
            const uint4 a = tp[0], b = tp[1];
            uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
            uint32_t planes[8];
            hut_bitslice32(w, planes);
            hut_bs_local loc;
            const uint32_t mydsm = ws.dsm[lane];
            tile_docs = __any_sync(FULL, mydsm != 0);
            hut_bs_classify_local(planes, mydsm, &loc);
            uint32_t nx = __shfl_down_sync(FULL, hut_bs_pack_next(&loc), 1);
            if (lane == 31) nx = 0;  // positions >= TRUST of the halo chunk are not used
            hut_bs_mid mid;
            hut_bs_lookahead(&loc, nx, &mid);
            uint32_t px = __shfl_up_sync(FULL, hut_bs_pack_prev(&loc, &mid), 1);
            if (lane == 0) px = 0;   // the look-back chunk's own starts are not used
            hut_bs_starts(&loc, &mid, px, &ST, &NU);
            if (__any_sync(FULL, mid.EX != 0)) generic_chunk(ws.text, ws.dsm, lane, &ST, &NU);
        }
        uint32_t bnd = ST | NU;
        if (lane == 31) bnd &= (1u << (TRUST - 31 * CHUNK)) - 1u;
        // the text ends at span position E = nbytes - g0 (> 32): the zero fill behind it reads as NUL boundaries, of
        // which only the one AT E is kept (it ends the last pretoken)
        const long long rem = (long long)J.nbytes - (g0 + lane * CHUNK);  // text bytes from this chunk's start on
        {
            const uint32_t valid = rem >= 32 ? FULL : (rem <= 0 ? 0u : ((1u << rem) - 1u));
            if (lane >= 1 && lane <= 30 && (NU & valid)) set_flag(J, FLAG_NUL);  // a NUL inside the text
            const uint32_t keep = rem >= 31 ? FULL : (rem < 0 ? 0u : ((2u << rem) - 1u));
            bnd &= keep;
        }
        // ---- enumerate the boundaries of chunks 1..31; those of chunks 1..30 start this tile's pretokens ----
        uint32_t n;
        {
            uint32_t m = lane >= 1 ? bnd : 0u;
            const uint32_t c = __popc(m);
            uint32_t inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, d);
                if (lane >= d) inc += t;
            }
            n = __shfl_sync(FULL, inc, 30);
            const uint32_t total = __shfl_sync(FULL, inc, 31);
            uint32_t k = inc - c;
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                ws.pstart[k++] = (uint16_t)(lane * CHUNK + b);
            }
            if (lane == 0) ws.pstart[total] = (uint16_t)BEYOND;
            // the boundary at the end of the text is not a pretoken
            const long long E = (long long)J.nbytes - g0;
            if (E < (long long)(31 * CHUNK)) n -= 1u;
        }
        __syncwarp();

        // ---- rounds of up to 32 pretokens, one per lane ----
        uint32_t run = 0;                        // ids in ws.ids
        uint32_t flushed = 0;                    // ids of this tile that left early (tile fuller than IDCAP)
        unsigned long long base = 0;             // ids of the tiles before this one
        bool have_base = tile == 0;
        uint32_t def_head = NODOC, def_tail = NODOC;
        uint32_t n_miss = 0, n_def = 0;
        // ids [0, run) of the slice leave for their final position; the look-back happens at the first flush
        auto flush = [&]() {
            if (!have_base) {
                base = lookback(J.status, tile);
                have_base = true;
            }
            const unsigned long long o = base + flushed;
            __syncwarp();
            if (o + run > J.out_cap) {
                if (lane == 0) set_flag(J, FLAG_OUT_CAP);
            } else {
                for (uint32_t i = lane; i < run; i += 32) st_stream_s32(J.out + o + i, ws.ids[i]);
            }
            __syncwarp();
            flushed += run;
            run = 0;
        };

        for (uint32_t r0 = 0; r0 < n; r0 += 32) {
            const uint32_t q = r0 + lane;
            const bool active = q < n;
            const int s = active ? ws.pstart[q] : BACK;
            const int e = active ? ws.pstart[q + 1] : BACK;
            const int len = e - s;                       // > MAXW when the end lies beyond the halo (BEYOND)
            const bool keyed = len <= CACHE_KEY_MAX;
            uint32_t dsb = 0;
            if (tile_docs) dsb = (ws.dsm[s >> 5] >> (s & 31)) & 1u;
            const uint32_t first = PFX ? dsb : 0u;
            uint32_t key[4], w[4] = {0u, 0u, 0u, 0u};
            const uint32_t b0 = build_key(ws.text, sm.kmask, s, keyed ? len : 0, first, key);
            bool hit = false;
            // pretokens of 16..31 bytes (rare): second cache, 32-byte keys
            if (__any_sync(FULL, !keyed)) {
                if (active && !keyed && len <= CACHE2_KEY_MAX && M.cache_ok) {
                    uint32_t k32[8];
                    build_key32(ws.text, sm.kmask32, s, len, first, k32);
                    hit = cache2_lookup(J, k32, w);
                }
            }
            if (active && keyed && M.cache_ok) hit = cache_lookup(J, key, w);
            // ---- counts of the hits -> offsets ----
            const uint32_t c = hit ? val_count(w[0]) : 0u;
            uint32_t inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, d);
                if (lane >= d) inc += t;
            }
            const uint32_t H = __shfl_sync(FULL, inc, 31);
            const uint32_t off = inc - c;                // ids of the round's hits in the lanes before this one
            const uint32_t need = __ballot_sync(FULL, active && !hit);
            if (need == 0) {
                // ---- every pretoken of the round was answered by the cache ----
                if (run + H > (uint32_t)IDCAP) flush();
                int32_t* const p = ws.ids + run + off;
                if (c > 0) p[0] = sext21(w[0]);
                if (c > 1) p[1] = sext21(w[1]);
                if (c > 2) p[2] = sext21(w[2]);
                if (c > 3) p[3] = sext21(w[3]);
                if (c > 4) p[4] = val_id4(w);
                if (tile_docs && active && dsb) {  // documents starting at this pretoken: ids of the tile before it
                    const unsigned long long gpos = (unsigned long long)(g0 + s);
                    unsigned long long d = dfirst;
                    while (d < J.ndocs && J.doc_off[d] < gpos) ++d;
                    while (d < J.ndocs && J.doc_off[d] == gpos) J.tok_off[d++] = flushed + run + off;
                }
                run += H;
                continue;
            }
            // ---- some pretokens were not answered by the cache (about one round in four on a cold cache): they are
            // merged by the whole warp, in lane order, so that the position of everybody's ids is known when it is
            // needed.  V(lane) = off + extra = ids of the round before this lane's.  UB = what the round can produce
            // at most; if that may not fit behind `run`, the earlier ids leave first, and if the round alone may not
            // fit it is handled in four parts of 8 lanes (8 * LANE_IDS_MAX <= IDCAP) with a flush in between. ----
            const uint32_t UB = H + (uint32_t)__popc(need) * (uint32_t)LANE_IDS_MAX;
            if (run + UB > (uint32_t)IDCAP) flush();
            const int nparts = UB > (uint32_t)IDCAP ? 4 : 1;
            const uint32_t tl0 = flushed + run;  // tile-local position of the round's first id
            int rbase = (int)run;                // slice index of the round's first id (negative after a part flush)
            uint32_t extra = 0;                  // ids of the misses in the lanes before this one
            uint32_t extra_all = 0;              // ids of all misses of the round so far
            for (int part = 0; part < nparts; ++part) {
                const uint32_t pmask = nparts == 1 ? FULL : (0xFFu << (8 * part));
                if (part) {  // the ids of the lanes before this part are complete: they leave, positions restart
                    const uint32_t v0 = __shfl_sync(FULL, off + extra, 8 * part);
                    run = (uint32_t)(rbase + (int)v0);
                    flush();
                    rbase = -(int)v0;
                }
                uint32_t mm = need & pmask;
                while (mm) {
                    const int j = __ffs((int)mm) - 1;
                    mm &= mm - 1;
                    const int sj = __shfl_sync(FULL, s, j), ej = __shfl_sync(FULL, e, j);
                    const uint32_t fj = __shfl_sync(FULL, first, j), bj = __shfl_sync(FULL, b0, j);
                    const uint32_t vj = __shfl_sync(FULL, off + extra, j);
                    const int lj = ej - sj;
                    uint32_t total = 0xFFFFFFFFu;
                    if (bj == 0) total = 0;  // a NUL in the text: the call fails (FLAG_NUL); emit nothing, cache nothing
                    else if (lj <= MAXW)
                        total = wmerge_one(M, sm.byte1, ws.text, ws.sc, sj, lj, fj, ws.ids + (rbase + (int)vj));
                    if (total == 0xFFFFFFFFu) {  // more than 32 bytes / symbols -> side list, patch phase
                        if (lane == 0)
                            defer_push2(J, tile, (unsigned long long)(g0 + sj), tl0 + vj,
                                        ej == (int)BEYOND ? 0u : (uint32_t)lj, fj, def_head, def_tail);
                        def_head = __shfl_sync(FULL, def_head, 0);
                        def_tail = __shfl_sync(FULL, def_tail, 0);
                        n_def++;
                        total = 0;
                    } else if (bj != 0) {
                        n_miss++;
                        if (lane == j && M.cache_ok && total >= 1u && total <= (uint32_t)CACHE_IDS &&
                            len <= CACHE2_KEY_MAX) {  // the owner still holds the key
                            const int32_t* src = ws.ids + (rbase + (int)vj);
                            int32_t t5[CACHE_IDS];
                            bool fits = true;  // ids are packed as signed 21-bit fields
#pragma unroll
                            for (int z = 0; z < CACHE_IDS; ++z) {
                                t5[z] = (uint32_t)z < total ? src[z] : 0;
                                fits = fits && t5[z] >= -(1 << 20) && t5[z] < (1 << 20);
                            }
                            if (fits) {
                                uint32_t pw[4];
                                val_pack(t5, total, pw);
                                if (keyed) {
                                    cache_insert(J, key, pw);
                                } else {
                                    uint32_t k32[8];
                                    build_key32(ws.text, sm.kmask32, s, len, first, k32);
                                    cache2_insert(J, k32, pw);
                                }
                            }
                        }
                    }
                    if (lane > j) extra += total;
                    extra_all += total;
                }
                // ---- ids of the part's hits (their positions are final now) ----
                if ((pmask >> lane) & 1u) {
                    int32_t* const p = ws.ids + (rbase + (int)(off + extra));
                    if (c > 0) p[0] = sext21(w[0]);
                    if (c > 1) p[1] = sext21(w[1]);
                    if (c > 2) p[2] = sext21(w[2]);
                    if (c > 3) p[3] = sext21(w[3]);
                    if (c > 4) p[4] = val_id4(w);
                }
                __syncwarp();
            }
            if (tile_docs && active && dsb) {
                const unsigned long long gpos = (unsigned long long)(g0 + s);
                unsigned long long d = dfirst;
                while (d < J.ndocs && J.doc_off[d] < gpos) ++d;
                while (d < J.ndocs && J.doc_off[d] == gpos) J.tok_off[d++] = tl0 + off + extra;
            }
            run = (uint32_t)(rbase + (int)(H + extra_all));
        }

        // ---- the tile's ids are complete: publish the count, learn the base, let the ids go ----
        const unsigned long long cnt = (unsigned long long)flushed + run;
        if (!have_base) {
            if (lane == 0) st_relaxed_u64(J.status + tile, ST_AGG | cnt);
            base = lookback(J.status, tile);
            have_base = true;
        }
        if (lane == 0) st_relaxed_u64(J.status + tile, ST_PFX | (base + cnt));
        if (run) flush();
        // documents starting in this tile hold tile-local positions: make them global (pre-patch) ones
        __syncwarp();
        if (base && tile_docs && dfirst != NODOC) {
            const unsigned long long tile_end = ((unsigned long long)tile + 1) * OWN;
            for (unsigned long long d = (unsigned long long)dfirst + lane;; d += 32) {
                const bool in = d < J.ndocs && J.doc_off[d] < tile_end && J.doc_off[d] < J.nbytes;
                if (in) J.tok_off[d] += base;
                if (!__all_sync(FULL, in)) break;
            }
        }
        if (lane == 0) {
            st_pre += n;
            st_miss += n_miss;
            st_def += n_def;
        }
        if (tile == J.ntiles - 1) {
            // the last tile closes the offsets: empty documents at the very end of the text and entry n_docs
            const unsigned long long total = base + cnt;
            if (lane == 0) J.ctr->total_ids = total;
            unsigned long long a = 0, b = J.ndocs;  // first document starting at / behind the end of the text
            while (a < b) {
                const unsigned long long mid = (a + b) >> 1;
                if (J.doc_off[mid] < J.nbytes) a = mid + 1;
                else b = mid;
            }
            for (unsigned long long d = a + lane; d <= J.ndocs; d += 32) J.tok_off[d] = total;
        }
        tile = __shfl_sync(FULL, next_tile, 0);
    }
    // ---- statistics ----
    if (lane == 0) {
        Counters* c = J.ctr;
        if (st_pre) atomicAdd(&c->st_pretokens, st_pre);
        if (st_pre - st_miss - st_def) atomicAdd(&c->st_hits, st_pre - st_miss - st_def);
        if (st_def) atomicAdd(&c->st_block, st_def);
        if (st_miss) atomicAdd(&c->st_warp, st_miss);
    }
}

}  // namespace hut
