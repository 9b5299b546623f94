/* sm_100a device code of the BPE encode engine, part 2: the single-pass tile kernel.  Included only by hut_engine.cu.
 *
 *   tile2_kernel    persistent, WARP per 960-byte tile, tiles claimed in order from an atomic ticket:
 *                     TMA bulk copy of the tile (+ halos) into the warp's shared-memory slice
 *                     bit-sliced class masks + local start rule (hut_bitslice.h)   <- parser_next_token (parser.c:24-88)
 *                     rounds of 32 pretokens, one per lane: probe of the unique-pretoken cache (one 32-byte sector)
 *                     cache misses of up to 32 symbols: merged right there by the whole warp, one symbol per lane
 *                                                       <- bpe_encode_arena_string (core.c:66-209), MinPQ (queue.c)
 *                     ids of the tile are put together in the warp's shared-memory slice
 *                     device-wide exclusive scan of the per-tile id counts, two levels: counts per tile, decoupled
 *                     look-back over groups of 128 tiles                <- IntVector growth (vector.c:41-55)
 *                     ids leave once, coalesced, at their final position; a warp already works on its next tile
 *                     while the previous one waits for its base         <- list building (lib.c:699-715, 810-860)
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

#ifdef HUT_TIMING
#define HUT_T(i) do { const long long t_ = clock64(); tm_[i] += t_ - t_last; t_last = t_; } while (0)
#else
#define HUT_T(i) do { } while (0)
#endif

namespace hut {

constexpr int RING = 1024;            // ids per warp kept in shared memory: the tile in work + up to 4 finished tiles that wait
                                      // for their base (a ring, so the depth adapts to how many ids the tiles hold)
constexpr int RMASK = RING - 1;
constexpr int PENDQ = 4;              // finished tiles a warp can have waiting
constexpr int LP_MIN = 6;             // cache misses in one round from which on they are merged lane-parallel
constexpr int WPB2 = 7;               // warps per CTA of tile2_kernel (they never synchronise with each other)
constexpr int NT2 = WPB2 * 32;
constexpr int GRP = 128;              // tiles per group of the two-level scan
constexpr int LANE_IDS_MAX = 40;      // most ids one pretoken merged in the tile kernel can produce (32 symbols + prefix ids)
constexpr unsigned long long ST_MASK = (1ull << 62) - 1;
constexpr unsigned long long ST_AGG = 1ull << 62;   // status: the tile's own id count is known
constexpr unsigned long long ST_PFX = 2ull << 62;   // status: the inclusive prefix (ids of tiles 0..t) is known

// shared memory of ONE WARP of tile2_kernel
struct __align__(16) WarpSm2 {
    uint8_t text[TEXTB];        // span bytes: [tile_start - 32, tile_start + 992 + 16), then zero padding
    uint32_t dsm[CHUNK + 4];    // per chunk: document starts
    uint16_t pstart[SPAN + 8];  // span positions of the boundaries in chunks 1..31 (31 = halo), then the end sentinel
    int32_t ring[RING];         // ids of the finished tiles that wait for their base, then of the tile in work
    uint32_t pq_tile[PENDQ], pq_start[PENDQ], pq_cnt[PENDQ], pq_flushed[PENDQ], pq_dfirst[PENDQ];
    unsigned long long pq_base[PENDQ];  // ~0 = not known yet
    uint32_t sc[64];            // symbols of the cache miss being merged
    unsigned long long mbar;
};

struct TileSm2 {
    uint4 kmask[16];            // kmask[len]: byte masks of a 16-byte key holding len text bytes
    uint4 kmask32[64];          // kmask32[2 * len], [2 * len + 1]: the same for the 32-byte keys of the second cache
    uint32_t byte1[256];
    WarpSm2 w[WPB2];
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
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Device-wide scan of the id counts, two levels (a look-back chain over single tiles advances 32 tiles per L2 round
// trip, which is slower than the tiles are produced):
//   level 0  cnt0[tile] = id count + 1 (0 = not known yet), written when the tile's rounds are done
//   level 1  groups of GRP tiles: every tile adds its count to grp_acc[g] (tiles done << 40 | ids); the warp that
//            completes the group publishes its total in grp_st[g] and turns it into an inclusive prefix by decoupled
//            look-back over the GROUPS (lookback() below, 32 groups = 4096 tiles per L2 round trip)
//   base(tile) = prefix of the group before + counts of the tiles before it in its own group (wait_base)
// Tiles are claimed in order from the ticket and every wait is for lower tiles / groups only, so the lowest
// unfinished tile never waits.
//
// lookback(): ids produced by groups 0 .. g-1.  Lane i looks at group (hi - i); the window moves back by 32 until a
// group with a known inclusive prefix is found.  Status words carry flag and value together, so one 64-bit load is a
// consistent snapshot.
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

// ids produced by the tiles before `tile`; block = wait until they are all known, else ~0 if they are not yet
__device__ __noinline__ unsigned long long tile_base(const Job& J, uint32_t tile, bool block) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t g = tile / GRP, i0 = g * GRP;
    for (;;) {
        bool ok = true;
        unsigned long long sum = 0;
        uint32_t v[GRP / 32];
#pragma unroll
        for (int k = 0; k < GRP / 32; ++k) {
            const uint32_t idx = i0 + lane + 32u * k;
            v[k] = idx < tile ? ld_relaxed_u32(J.cnt0 + idx) : 1u;
        }
#pragma unroll
        for (int k = 0; k < GRP / 32; ++k) {
            ok = ok && v[k] != 0u;
            sum += v[k] - 1u;
        }
        if (g > 0 && lane == 0) {
            const unsigned long long pf = ld_relaxed_u64(J.grp_st + (g - 1));
            ok = ok && (pf >> 62) == 2ull;
            sum += pf & ST_MASK;
        }
        if (__all_sync(FULL, ok)) {
#pragma unroll
            for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(FULL, sum, d);
            return sum;
        }
        if (!block) return ~0ull;
#ifdef HUT_TIMING
        {
            const long long t0 = clock64();
            __nanosleep(64);
            if (lane == 0) atomicAdd(&J.ctr->tw[0], (unsigned long long)(clock64() - t0) + 1500ull);
        }
#else
        __nanosleep(64);
#endif
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
                                            int s, int len, uint32_t first, int32_t* ring, uint32_t dst) {
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
    if (lane < nal) ring[(dst + lane) & RMASK] = M.prefix_alone[lane];
    if (lane < m) ring[(dst + nal + lane) & RMASK] = id_of_unit(M, u);
    __syncwarp();
    return nal + m;
}

// ------------------------------------------------------------------------------------------------
// The same merge by ONE LANE (arrays in local memory, which the hardware interleaves by lane): used when a round has
// many cache misses (a cold cache), so that up to 32 of them run at once and their table probes overlap.  Per merge one
// scan for the leftmost minimal rank -- the (rank, left index) order of the reference's heap (queue.c:162-164) -- and
// two pair-table probes (issued together) for the new neighbours; symbols never move, a merged-away symbol is cleared
// in `alive`.  out[0 .. total) = ids; returns total or 0xFFFFFFFF (more than 32 symbols: deferred).
// ------------------------------------------------------------------------------------------------
__device__ __noinline__ uint32_t merge_lane(const DevModel& M, const uint32_t* byte1, const uint8_t* text, int s, int len,
                                            uint32_t first, int32_t* out) {
    uint32_t U[32], MG[32];
    auto at = [&](unsigned long long p) -> uint32_t { return (uint32_t)text[p]; };
    const uint32_t b0 = text[s];
    const bool with_prefix = first && b0 != ' ';
    const bool alone_first = first && b0 == ' ';
    const uint32_t npfx = with_prefix ? M.n_prefix_syms : 0u;
    const uint32_t nal = alone_first ? M.n_prefix_alone : 0u;
    if (npfx > 32u || nal > 8u) return 0xFFFFFFFFu;
    uint32_t m = 0;
    for (uint32_t k = 0; k < npfx; ++k) U[m++] = M.prefix_syms[k];
    for (int p = 0; p < len; ++p) {
        uint32_t unit = M.inert, pb = 0;
        const uint32_t cnt = char_symbols(M, byte1, at, (unsigned long long)(s + p), (unsigned long long)s,
                                          (unsigned long long)(s + len), &unit, &pb);
        if (m + cnt > 32u) return 0xFFFFFFFFu;
        if (cnt == 1) U[m++] = unit;
        else
            for (uint32_t k = 0; k < cnt; ++k) U[m++] = __ldg(M.sym_pool + pb + k);
    }
    const uint32_t m0 = m;
    for (uint32_t k = 0; k < m0; ++k) MG[k] = k + 1 < m0 ? pair_probe(M, U[k], U[k + 1]) : NONE;
    uint32_t alive = m0 >= 32 ? 0xFFFFFFFFu : ((1u << m0) - 1u);
    for (;;) {
        uint32_t best = 0xFFFFFFFFu, j = 0, mu = NONE;
        for (uint32_t k = 0; k + 1 < m0; ++k) {
            const uint32_t mg = MG[k];
            const uint32_t rk = rank_of_unit(M, mg);
            if (rk < best) {
                best = rk;
                j = k;
                mu = mg;
            }
        }
        if (best == 0xFFFFFFFFu) break;
        const uint32_t r = j + (uint32_t)__ffs((int)(alive >> (j + 1)));  // right partner of j
        alive &= ~(1u << r);
        U[j] = mu;
        MG[r] = NONE;
        const uint32_t below = alive & ((1u << j) - 1u), above = r >= 31 ? 0u : (alive >> (r + 1));
        const int l = below ? 31 - __clz((int)below) : -1;
        const int r2 = above ? (int)(r + 1) + (__ffs((int)above) - 1) : -1;
        const uint32_t ul = l >= 0 ? U[l] : NONE, ur = r2 >= 0 ? U[r2] : NONE;
        uint32_t ml, mr;
        pair_probe2(M, ul, mu, ur, &ml, &mr);
        if (l >= 0) MG[l] = ml;
        MG[j] = mr;  // NONE when j is the last symbol now
    }
    uint32_t k = 0;
    for (; k < nal; ++k) out[k] = M.prefix_alone[k];
    for (uint32_t i = 0; i < m0; ++i)
        if ((alive >> i) & 1u) out[k++] = id_of_unit(M, U[i]);
    return k;
}


// ------------------------------------------------------------------------------------------------
// The warp's ring of ids and its queue of finished tiles.  These helpers are deliberately NOT inlined and take the
// state through a struct that only lives around the call: the hot loop of tile2_kernel keeps the same values in
// registers (an address-taken struct would sit in local memory for the whole kernel).
// ------------------------------------------------------------------------------------------------
struct WState {
    uint32_t ph, pn;          // pending queue: index of the oldest entry, entries
    uint32_t pend_ids;        // ids held by the pending tiles
    uint32_t rpos;            // ring index where the tile in work starts
    uint32_t run, flushed;    // tile in work: ids in the ring, ids that left early
    unsigned long long base;  // tile in work: ids of the tiles before it, ~0 = not known yet
};

__device__ __noinline__ void emit_ids(const Job& J, const int32_t* ring, uint32_t start, unsigned long long o, uint32_t cnt) {
    const uint32_t lane = threadIdx.x & 31;
    __syncwarp();
    if (o + cnt > J.out_cap) {
        if (lane == 0) set_flag(J, FLAG_OUT_CAP);
    } else {
        for (uint32_t i = lane; i < cnt; i += 32) st_stream_s32(J.out + o + i, ring[(start + i) & RMASK]);
    }
    __syncwarp();
}

// the oldest waiting tile gets its base (block: waits for it), its ids leave, its documents' offsets become global
// (pre-patch) ones.  Returns false if the base is not known yet.
__device__ __noinline__ bool finalize_oldest(const Job& J, WarpSm2& ws, WState& st, bool block) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t k = st.ph & (PENDQ - 1);
    const uint32_t p_tile = ws.pq_tile[k];
    unsigned long long p_base = ws.pq_base[k];
    if (p_base == ~0ull) {
        p_base = tile_base(J, p_tile, block);
        if (p_base == ~0ull) return false;
    }
    const uint32_t p_cnt = ws.pq_cnt[k], p_flushed = ws.pq_flushed[k], p_dfirst = ws.pq_dfirst[k];
    emit_ids(J, ws.ring, ws.pq_start[k], p_base + p_flushed, p_cnt);
    if (lane == 0) J.status[p_tile] = p_base + p_flushed + p_cnt;  // inclusive prefix, read by the patch phase
    if (p_base && p_dfirst != NODOC) {
        const unsigned long long tile_end = ((unsigned long long)p_tile + 1) * OWN;
        for (unsigned long long d = (unsigned long long)p_dfirst + lane;; d += 32) {
            const bool in = d < J.ndocs && J.doc_off[d] < tile_end && J.doc_off[d] < J.nbytes;
            if (in) J.tok_off[d] += p_base;
            if (!__all_sync(0xFFFFFFFFu, in)) break;
        }
    }
    st.pend_ids -= p_cnt;
    st.ph++;
    st.pn--;
    return true;
}

// makes room for k more ids of the tile in work: waiting tiles leave first (each waits for its base), and if the ring
// holds the tile in work alone its ids so far leave early, which needs that tile's base now
__device__ __noinline__ void make_room(const Job& J, WarpSm2& ws, WState& st, uint32_t tile, uint32_t k) {
    while (st.pend_ids + st.run + k > (uint32_t)RING) {
        if (st.pn) {
            finalize_oldest(J, ws, st, true);
        } else {
            if (st.base == ~0ull) st.base = tile_base(J, tile, true);
            emit_ids(J, ws.ring, st.rpos, st.base + st.flushed, st.run);
            st.flushed += st.run;
            st.rpos = (st.rpos + st.run) & RMASK;
            st.run = 0;
        }
    }
}

// end of a tile: it joins the queue of waiting tiles; those whose base has arrived leave now (the newest one had no
// time yet: it is looked at after the next tile)
__device__ __noinline__ void tile_done(const Job& J, WarpSm2& ws, WState& st, uint32_t tile, uint32_t dfirst) {
    const uint32_t lane = threadIdx.x & 31;
    if (st.pn == (uint32_t)PENDQ) finalize_oldest(J, ws, st, true);
    const uint32_t k = (st.ph + st.pn) & (PENDQ - 1);
    __syncwarp();
    if (lane == 0) {
        ws.pq_tile[k] = tile;
        ws.pq_start[k] = st.rpos;
        ws.pq_cnt[k] = st.run;
        ws.pq_flushed[k] = st.flushed;
        ws.pq_dfirst[k] = dfirst;
        ws.pq_base[k] = st.base;
    }
    __syncwarp();
    st.pn++;
    st.pend_ids += st.run;
    st.rpos = (st.rpos + st.run) & RMASK;
    while (st.pn > 1 && finalize_oldest(J, ws, st, false)) {}
}

// the warp that completes a group turns the group's total into an inclusive prefix (and the last group closes the
// document offsets: empty documents at the very end of the text and entry n_docs)
__device__ __noinline__ void close_group(const Job& J, uint32_t g, uint32_t ngroups, unsigned long long gtotal) {
    const uint32_t lane = threadIdx.x & 31;
    unsigned long long before = 0;
    if (g > 0) {
        if (lane == 0) st_relaxed_u64(J.grp_st + g, ST_AGG | gtotal);
        before = lookback(J.grp_st, g);
    }
    if (lane == 0) st_relaxed_u64(J.grp_st + g, ST_PFX | (before + gtotal));
    if (g + 1 == ngroups) {
        const unsigned long long total = before + gtotal;
        if (lane == 0) J.ctr->total_ids = total;
        unsigned long long a = 0, b = J.ndocs;  // first document starting at / behind the end of the text
        while (a < b) {
            const unsigned long long mid = (a + b) >> 1;
            if (J.doc_off[mid] < J.nbytes) a = mid + 1;
            else b = mid;
        }
        for (unsigned long long d = a + lane; d <= J.ndocs; d += 32) J.tok_off[d] = total;
    }
}

__device__ __forceinline__ void put_hit_ids(int32_t* ring, uint32_t p, uint32_t c, const uint32_t w[4]) {
    if (c > 0) ring[p & RMASK] = sext21(w[0]);
    if (c > 1) ring[(p + 1) & RMASK] = sext21(w[1]);
    if (c > 2) ring[(p + 2) & RMASK] = sext21(w[2]);
    if (c > 3) ring[(p + 3) & RMASK] = sext21(w[3]);
    if (c > 4) ring[(p + 4) & RMASK] = val_id4(w);
}

// documents starting at a pretoken of this tile: their id offset = ids of the tile before the pretoken (made global
// when the tile's base is known)
__device__ __noinline__ void doc_offsets_at(const Job& J, unsigned long long gpos, unsigned long long dfirst, uint32_t v) {
    unsigned long long d = dfirst;
    while (d < J.ndocs && J.doc_off[d] < gpos) ++d;
    while (d < J.ndocs && J.doc_off[d] == gpos) J.tok_off[d++] = v;
}

// what a round of the tile kernel hands to rare_round (lane values + warp-uniform state)
struct RareArgs {
    // per lane
    int s, e;
    uint32_t first, b0, c, off, dsb;
    uint32_t key[4], w[4];
    bool active, hit;
    // uniform
    uint32_t H, need, tile, dfirst;
    long long g0;
    bool tile_docs;
    uint32_t def_head, def_tail, n_miss, n_def;
    WState st;
};

// A round in which some pretokens were not answered by the cache.  Few misses: each is merged by the whole warp
// (wmerge_one), in lane order, so that the position of its ids is known when they are written.  Many misses (a cold
// cache): every lane merges its own (merge_lane), all at once.  Pretokens of more than 32 bytes / symbols go to the
// side list.  UB = what the round can produce at most (hits: their count; misses: one id per symbol + prefix ids;
// deferred: nothing); if the round alone may not fit the ring it is handled in four parts of 8 lanes
// (8 * LANE_IDS_MAX <= RING) with the ring emptied in between.
template <bool PFX>
__device__ __noinline__ void rare_round(const DevModel& M, const Job& J, TileSm2& sm, WarpSm2& ws, RareArgs& A) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t FULL = 0xFFFFFFFFu;
    const int s = A.s, e = A.e, len = e - s;
    const uint32_t first = A.first, b0 = A.b0, off = A.off, H = A.H, need = A.need, tile = A.tile;
    const bool keyed = len <= CACHE_KEY_MAX;
    const bool mine = A.active && !A.hit;
    uint32_t c = A.c;
    WState& st = A.st;
    uint32_t UB;
    {
        uint32_t ub = c;
        if (mine && len <= MAXW) {
            ub = (uint32_t)len * M.kmax + (first ? M.first_extra : 0u);
            if (ub > (uint32_t)LANE_IDS_MAX) ub = LANE_IDS_MAX;
        }
#pragma unroll
        for (int d = 16; d; d >>= 1) ub += __shfl_xor_sync(FULL, ub, d);
        UB = ub;
    }
    const int nparts = UB > (uint32_t)RING / 2 ? 4 : 1;
    make_room(J, ws, st, tile, nparts == 4 ? (uint32_t)RING : UB);  // 4 parts: they start from an empty ring
    const uint32_t tl0 = st.flushed + st.run;  // tile-local position of the round's first id
    uint32_t V = off;                          // ids of the round before this lane's
    uint32_t round_total = H;
    auto insert = [&](const int32_t* t5, uint32_t total) {
        uint32_t pw[4];
        val_pack(t5, total, pw);
        if (keyed) {
            cache_insert(J, A.key, pw);
        } else {
            uint32_t k32[8];
            build_key32(ws.text, sm.kmask32, s, len, first, k32);
            cache2_insert(J, k32, pw);
        }
    };
    auto defer = [&](int sj, int ej, uint32_t fj, uint32_t vj) {
        if (lane == 0)
            defer_push2(J, tile, (unsigned long long)(A.g0 + sj), tl0 + vj, ej == (int)BEYOND ? 0u : (uint32_t)(ej - sj), fj,
                        A.def_head, A.def_tail);
        A.def_head = __shfl_sync(FULL, A.def_head, 0);
        A.def_tail = __shfl_sync(FULL, A.def_tail, 0);
    };

    if (nparts == 1 && __popc(need) >= LP_MIN) {
        // ---- many misses: lane-parallel ----
        int32_t mine_ids[LANE_IDS_MAX];
        uint32_t total = 0;
        bool dfr = false;
        if (mine) {
            if (b0 == 0) total = 0;  // a NUL in the text: the call fails (FLAG_NUL); emit nothing, cache nothing
            else if (len > MAXW) dfr = true;
            else {
#ifdef HUT_TIMING
                const long long t0 = clock64();
#endif
                total = merge_lane(M, sm.byte1, ws.text, s, len, first, mine_ids);
#ifdef HUT_TIMING
                atomicAdd(&J.ctr->tw[3], (unsigned long long)(clock64() - t0));
#endif
                if (total == 0xFFFFFFFFu) {
                    dfr = true;
                    total = 0;
                }
            }
            c = total;
        }
        uint32_t inc2 = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(FULL, inc2, d);
            if (lane >= (uint32_t)d) inc2 += t;
        }
        round_total = __shfl_sync(FULL, inc2, 31);
        V = inc2 - c;
        uint32_t dm = __ballot_sync(FULL, dfr);
        A.n_def += __popc(dm);
        A.n_miss += __popc(need & ~dm);
        while (dm) {  // in lane order
            const int j = __ffs((int)dm) - 1;
            dm &= dm - 1;
            defer(__shfl_sync(FULL, s, j), __shfl_sync(FULL, e, j), __shfl_sync(FULL, first, j), __shfl_sync(FULL, V, j));
        }
        const uint32_t p = st.rpos + st.run + V;
        if (mine) {
            for (uint32_t z = 0; z < total; ++z) ws.ring[(p + z) & RMASK] = mine_ids[z];
            if (!dfr && b0 != 0 && M.cache_ok && total >= 1u && total <= (uint32_t)CACHE_IDS && len <= CACHE2_KEY_MAX) {
                int32_t t5[CACHE_IDS];
#pragma unroll
                for (int z = 0; z < CACHE_IDS; ++z) t5[z] = (uint32_t)z < total ? mine_ids[z] : 0;
                insert(t5, total);
            }
        } else {
            put_hit_ids(ws.ring, p, c, A.w);
        }
        __syncwarp();
    } else {
        // ---- few misses: one after the other, each by the whole warp ----
        uint32_t extra = 0;       // ids of the misses in the lanes before this one
        uint32_t extra_all = 0;   // ids of all misses of the round so far
        uint32_t gone = 0;        // ids of the round that left with an earlier part
        for (int part = 0; part < nparts; ++part) {
            const uint32_t pmask = nparts == 1 ? FULL : (0xFFu << (8 * part));
            if (part) {  // the ids of the lanes before this part are complete: they leave, the ring is empty again
                const uint32_t v0 = __shfl_sync(FULL, off + extra, 8 * part);
                st.run += v0 - gone;
                gone = v0;
                make_room(J, ws, st, tile, RING);
            }
            uint32_t mm = need & pmask;
            while (mm) {
                const int j = __ffs((int)mm) - 1;
                mm &= mm - 1;
                const int sj = __shfl_sync(FULL, s, j), ej = __shfl_sync(FULL, e, j);
                const uint32_t fj = __shfl_sync(FULL, first, j), bj = __shfl_sync(FULL, b0, j);
                const uint32_t vj = __shfl_sync(FULL, off + extra, j);
                const int lj = ej - sj;
                const uint32_t dst = st.rpos + st.run + vj - gone;
                uint32_t total = 0xFFFFFFFFu;
                if (bj == 0) total = 0;  // a NUL in the text: the call fails (FLAG_NUL); emit nothing, cache nothing
                else if (lj <= MAXW) {
#ifdef HUT_TIMING
                    const long long t0 = clock64();
#endif
                    total = wmerge_one(M, sm.byte1, ws.text, ws.sc, sj, lj, fj, ws.ring, dst);
#ifdef HUT_TIMING
                    if (lane == 0) atomicAdd(&J.ctr->tw[2], (unsigned long long)(clock64() - t0));
#endif
                }
                if (total == 0xFFFFFFFFu) {
                    defer(sj, ej, fj, vj);
                    A.n_def++;
                    total = 0;
                } else if (bj != 0) {
                    A.n_miss++;
                    if (lane == (uint32_t)j && M.cache_ok && total >= 1u && total <= (uint32_t)CACHE_IDS &&
                        len <= CACHE2_KEY_MAX) {  // the owner still holds the key
                        int32_t t5[CACHE_IDS];
#pragma unroll
                        for (int z = 0; z < CACHE_IDS; ++z) t5[z] = (uint32_t)z < total ? ws.ring[(dst + z) & RMASK] : 0;
                        insert(t5, total);
                    }
                }
                if (lane > (uint32_t)j) extra += total;
                extra_all += total;
            }
            // ---- ids of the part's hits (their positions are final now) ----
            if ((pmask >> lane) & 1u) put_hit_ids(ws.ring, st.rpos + st.run + off + extra - gone, c, A.w);
            __syncwarp();
        }
        V = off + extra;
        round_total = H + extra_all - gone;
    }
    if (A.tile_docs && A.active && A.dsb) doc_offsets_at(J, (unsigned long long)(A.g0 + s), A.dfirst, tl0 + V);
    st.run += round_total;
}

// PFX: the model has a prefix (the first pretoken of a document is keyed apart and merged with the prefix symbols).
template <bool PFX>
__global__ void __launch_bounds__(NT2, HUT_TILE_CTAS) tile2_kernel(const __grid_constant__ DevModel M,
                                                                    const __grid_constant__ Job J) {
    extern __shared__ __align__(16) unsigned char tile2_smem[];
    TileSm2& sm = *reinterpret_cast<TileSm2*>(tile2_smem);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t FULL = 0xFFFFFFFFu;
    WarpSm2& ws = sm.w[warp];

    for (int i = tid; i < 256; i += NT2) sm.byte1[i] = M.byte1[i];
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

    // ---- the warp's ring of ids: finished tiles that wait for their base (oldest first), then the tile in work.  A
    // warp goes on with its next tiles while earlier ones wait, so a tile that takes long (many cache misses) does not
    // stop the tiles behind it until the ring is full.  (WState, kept in registers here.) ----
    uint32_t ph = 0, pn = 0, pend_ids = 0, rpos = 0;
    const uint32_t ngroups = (J.ntiles + GRP - 1) / GRP;

#ifdef HUT_TIMING
    long long tm_[8] = {0, 0, 0, 0, 0, 0, 0, 0}, t_last = clock64();
#endif
    while (tile < J.ntiles) {
        HUT_T(7);
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
        HUT_T(0);

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
        {
            const long long rem = (long long)J.nbytes - (g0 + lane * CHUNK);  // text bytes from this chunk's start on
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
        HUT_T(1);

#ifdef HUT_TIMING
        const long long t_tile = clock64();
#endif
        // ---- rounds of up to 32 pretokens, one per lane ----
        uint32_t run = 0;                        // ids of this tile in the ring, at rpos ..
        uint32_t flushed = 0;                    // ids of this tile that left early (the ring was full of this tile alone)
        unsigned long long base = ~0ull;         // ids of the tiles before this one, ~0 = not known yet
        uint32_t def_head = NODOC, def_tail = NODOC;
        uint32_t n_miss = 0, n_def = 0;

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
                if (pend_ids + run + H > (uint32_t)RING) {
                    WState st = {ph, pn, pend_ids, rpos, run, flushed, base};
                    make_room(J, ws, st, tile, H);
                    ph = st.ph, pn = st.pn, pend_ids = st.pend_ids, rpos = st.rpos, run = st.run, flushed = st.flushed, base = st.base;
                }
                put_hit_ids(ws.ring, rpos + run + off, c, w);
                if (tile_docs && active && dsb) doc_offsets_at(J, (unsigned long long)(g0 + s), dfirst, flushed + run + off);
                run += H;
                HUT_T(2);
                continue;
            }
            {
                RareArgs A;
                A.s = s, A.e = e, A.first = first, A.b0 = b0, A.c = c, A.off = off, A.dsb = dsb;
#pragma unroll
                for (int z = 0; z < 4; ++z) A.key[z] = key[z], A.w[z] = w[z];
                A.active = active, A.hit = hit;
                A.H = H, A.need = need, A.tile = tile, A.dfirst = dfirst, A.g0 = g0, A.tile_docs = tile_docs;
                A.def_head = def_head, A.def_tail = def_tail, A.n_miss = n_miss, A.n_def = n_def;
                A.st = WState{ph, pn, pend_ids, rpos, run, flushed, base};
                rare_round<PFX>(M, J, sm, ws, A);
                def_head = A.def_head, def_tail = A.def_tail, n_miss = A.n_miss, n_def = A.n_def;
                ph = A.st.ph, pn = A.st.pn, pend_ids = A.st.pend_ids, rpos = A.st.rpos, run = A.st.run, flushed = A.st.flushed,
                base = A.st.base;
            }
            HUT_T(3);
        }

#ifdef HUT_TIMING
        if (lane == 0) {
            const long long dt = clock64() - t_tile;
            atomicAdd(&J.ctr->hist[63 - __clzll(dt | 1)], 1u);
            atomicAdd(&J.ctr->hmiss[n_miss > 31 ? 31 : n_miss], 1u);
        }
#endif
        // ---- the tile's ids are complete: publish the count; the ids wait in the ring for their base ----
        const uint32_t cnt = flushed + run;
        const uint32_t g = tile / GRP;
        unsigned long long acc_old = 0;
        if (lane == 0) {
            st_relaxed_u32(J.cnt0 + tile, cnt + 1u);
            acc_old = atomicAdd(J.grp_acc + g, (1ull << 40) | (unsigned long long)cnt);
        }
        HUT_T(4);
        {
            WState st = {ph, pn, pend_ids, rpos, run, flushed, base};
            tile_done(J, ws, st, tile, tile_docs ? dfirst : NODOC);
            ph = st.ph, pn = st.pn, pend_ids = st.pend_ids, rpos = st.rpos;
        }
        HUT_T(5);
        if (lane == 0) {
            st_pre += n;
            st_miss += n_miss;
            st_def += n_def;
        }
        acc_old = __shfl_sync(FULL, acc_old, 0);
        {
            const uint32_t gsize = g + 1 < ngroups ? (uint32_t)GRP : J.ntiles - g * GRP;
            if ((uint32_t)(acc_old >> 40) + 1u == gsize) close_group(J, g, ngroups, (acc_old & ((1ull << 40) - 1)) + cnt);
        }
        HUT_T(6);
        tile = __shfl_sync(FULL, next_tile, 0);
    }
    {
        WState st = {ph, pn, pend_ids, rpos, 0u, 0u, ~0ull};
        while (st.pn) finalize_oldest(J, ws, st, true);
    }
#ifdef HUT_TIMING
    if (lane == 0)
        for (int i = 0; i < 8; ++i) atomicAdd(&J.ctr->tm[i], (unsigned long long)tm_[i]);
#endif
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
