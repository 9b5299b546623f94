/* sm_100a device code of the BPE encode engine, part 2: the kernels of the pipeline (see hut_kernels.cuh for
 * the overview).  Included only by hut_engine.cu. */
#pragma once

#include "hut_kernels.cuh"

#ifndef HUT_TILE_CTAS
#define HUT_TILE_CTAS 4  // resident CTAs per SM tile_kernel is compiled for (64 registers per thread)
#endif

namespace hut {

// shared memory of ONE WARP of tile_kernel
struct __align__(16) WarpSm {
    uint8_t text[TEXTB];        // span bytes: [tile_start - 32, tile_start + 992 + 16), then zero padding
    uint32_t dsm[CHUNK + 4];    // per chunk: document starts
    uint16_t pstart[OWN + 8];   // span positions of the boundaries inside the owned chunks, then the end sentinel
    uint4 q[64];                // cache misses waiting for a full batch: pos_off (x, y), meta (z), tile (w)
    unsigned long long mbar;
};

struct TileSm {
    uint4 kmask[16];            // kmask[len]: byte masks of a 16-byte key holding len text bytes
    uint4 kmask32[64];          // kmask32[2 * len], [2 * len + 1]: the same for the 32-byte keys of the second cache
    uint32_t byte1[256];
    WarpSm w[WPB];
};

__global__ void init_kernel(Job J) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        Counters* c = J.ctr;
        c->flags = 0;
        c->n_events = 0;
        c->n_def = 0;
        c->n_miss = 0;
        c->total_ids = 0;
        c->long_used = 0;
        c->st_pretokens = c->st_hits = c->st_warp = c->st_block = c->st_big = c->st_cut = 0;
        for (int i = 0; i < 32; ++i) c->pool_lock[i] = 0;
    }
}

// first_doc[t] = first document starting inside tile t (pre-set to NODOC)
__global__ void doc_kernel(Job J) {
    const unsigned long long d = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= J.ndocs) return;
    const unsigned long long pos = J.doc_off[d];
    if (pos >= J.nbytes) return;
    const unsigned long long t = pos / OWN;
    if (d == 0 || J.doc_off[d - 1] / OWN != t) J.first_doc[t] = (uint32_t)d;
}

// scalar restatement (hut_classify.h) of one 32-byte chunk: only used when an over-long UTF-8 form is in the tile
// start of every tile's staging slot: tile * stride, plus slot_extra ids for each document starting before the tile
// (lower bound of the tile's first byte in doc_off) for a prefix model; see stage_slot()
__global__ void slot_docs_kernel(Job J, unsigned long long* out) {
    const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t > J.ntiles) return;
    unsigned long long a = 0;
    if (J.slot_extra) {
        const unsigned long long pos = t * OWN;
        unsigned long long b = J.ndocs;
        while (a < b) {
            const unsigned long long mid = (a + b) >> 1;
            if (J.doc_off[mid] < pos) a = mid + 1;
            else b = mid;
        }
    }
    out[t] = t * J.stride + a * J.slot_extra;
}

__device__ __noinline__ void generic_chunk(const WarpSm& ws, int lane, uint32_t* ST, uint32_t* NU) {
    auto dsbit = [&](int x) -> uint32_t { return (x >= 0 && x < SPAN) ? ((ws.dsm[x >> 5] >> (x & 31)) & 1u) : 0u; };
    auto lead = [&](int x) -> uint32_t {
        uint32_t b0 = ws.text[x], b1 = ws.text[x + 1], b2 = ws.text[x + 2], b3 = ws.text[x + 3];
        if (dsbit(x + 1)) b1 = b2 = b3 = 0;
        else if (dsbit(x + 2)) b2 = b3 = 0;
        else if (dsbit(x + 3)) b3 = 0;
        return hut_lead_class(b0, b1, b2, b3);
    };
    auto code = [&](int x) -> uint32_t {
        if (x < 0) return HC_END;
        const uint32_t b0 = ws.text[x];
        if ((b0 & 0xC0u) == 0x80u) {
            for (int k = 1; k <= 3; ++k) {
                const int q = x - k;
                if (q < 0) break;
                if ((ws.text[q] & 0xC0u) == 0x80u) continue;
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

// appends a deferred pretoken to the global list and links it to its tile's chain (one lane calls)
__device__ __forceinline__ void defer_push(const Job& J, uint32_t tile, unsigned long long pos, uint32_t local_off,
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
    e.flags = first;
    e.count = 0;
    e.pad = NODOC;  // next entry of the same tile
    e.ids_off = 0;
    J.def[idx] = e;
    if (head == NODOC) head = idx;
    else J.def[tail].pad = idx;
    tail = idx;
}

// merges the first n (<= 32) queued cache misses, one per lane (merge_entry).  The room a pretoken of another tile
// did not use is subtracted from that tile's count here; for the tile in progress it is returned (warp sum).
__device__ __forceinline__ uint32_t run_batch(const DevModel& M, const Job& J, const uint32_t* byte1, const uint4* q,
                                           uint32_t n, uint32_t cur_tile) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t slack = 0;
    if (lane < n) {
        const uint4 E = q[lane];
        const unsigned long long pos_off = E.x | ((unsigned long long)E.y << 32);
        const unsigned long long pos = pos_off & ((1ull << 40) - 1);
        const uint32_t loff = (uint32_t)(pos_off >> 40);
        const uint32_t len = E.z & 0xFFu, first = (E.z >> 8) & 1u, room = E.z >> 16;
        slack = merge_entry<MISS_SMALL>(M, J, byte1, pos, len, first, room,
                                        stage_slot(J, E.w) + loff);
        if (slack && E.w != cur_tile) {
            atomicSub(&J.tot[E.w], slack);
            slack = 0;
        }
    }
    __syncwarp();
#pragma unroll
    for (int d = 16; d; d >>= 1) slack += __shfl_xor_sync(0xFFFFFFFFu, slack, d);
    return slack;
}

// PFX: the model has a prefix (first pretoken of a document keyed apart, staging slots of uneven size read from
// J.slot_off); without one the slot of a tile is tile * stride, a warp-uniform product that needs no load.
template <bool PFX>
__global__ void __launch_bounds__(NT, HUT_TILE_CTAS) tile_kernel(const __grid_constant__ DevModel M, const __grid_constant__ Job J) {
    __shared__ TileSm sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lt_mask = (1u << lane) - 1u;
    WarpSm& ws = sm.w[warp];

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
    __syncthreads();  // the only block-wide barrier: kmask is read-only afterwards

    uint32_t parity = 0;
    uint32_t st_pre = 0, st_hit = 0;  // per lane
    unsigned long long st_def = 0, st_wm = 0;
    uint32_t qn = 0;                  // cache misses queued in ws.q
    const uint32_t gw = blockIdx.x * WPB + warp, nw = gridDim.x * WPB;

    for (uint32_t tile = J.tile_lo + gw; tile < J.tile_hi; tile += nw) {
        unsigned long long slot_o = 0;
        if (PFX) slot_o = __ldg(J.slot_off + tile);  // asked for early: used only after the classifier
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
            if (__any_sync(FULL, mid.EX != 0)) generic_chunk(ws, lane, &ST, &NU);
        }
        uint32_t bnd = ST | NU;
        if (lane == 31) bnd &= (1u << (TRUST - 31 * CHUNK)) - 1u;
        const bool own = lane >= 1 && lane <= 30;
        if (own && NU) {  // a NUL inside the text (not the zero fill past its end)?
            const long long rem = (long long)J.nbytes - (g0 + lane * CHUNK);
            const uint32_t valid = rem >= 32 ? FULL : (rem <= 0 ? 0u : ((1u << rem) - 1u));
            if (NU & valid) set_flag(J, FLAG_NUL);
        }
        // ---- enumerate the boundaries of the owned chunks ----
        uint32_t n;
        {
            const uint32_t c = own ? __popc(bnd) : 0u;
            uint32_t inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, d);
                if (lane >= d) inc += t;
            }
            n = __shfl_sync(FULL, inc, 31);
            uint32_t k = inc - c, m = own ? bnd : 0u;
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                ws.pstart[k++] = (uint16_t)(lane * CHUNK + b);
            }
            if (lane == 31) ws.pstart[n] = bnd ? (uint16_t)(31 * CHUNK + __ffs(bnd) - 1) : (uint16_t)BEYOND;
        }
        __syncwarp();
#ifdef HUT_EXP_NOROUND
        n = 0;  // timing experiment only: staging + classify + enumerate without the rounds
#endif

        // ---- rounds of up to 32 pretokens, one per lane ----
        uint32_t run = 0;                       // entries of this tile's staging slot used so far
        uint32_t def_head = NODOC, def_tail = NODOC;
        bool holes = false;
        uint32_t slack = 0;                     // room reserved for misses of this tile that stayed unused
        int32_t* const slot = J.stage + (PFX ? slot_o : (unsigned long long)tile * J.stride);
#ifdef HUT_PIPE_PROBE
        // software pipeline of the cache probe: the first slot of round r + 1 is requested right after round r's
        // compare, so that the scan and the stores of round r run under its latency (one request per lane and round,
        // as before)
        int n_s = BACK, n_e = BACK;
        uint32_t n_dsb = 0, n_b0 = 0, n_i = 0, n_key[4] = {0u, 0u, 0u, 0u};
        uint4 n_a = make_uint4(0, 0, 0, 0), n_b = make_uint4(0, 0, 0, 0);
        auto prep = [&](uint32_t r) {
            const uint32_t q = r + lane;
            const bool active = q < n;
            n_s = active ? ws.pstart[q] : BACK;
            n_e = active ? ws.pstart[q + 1] : BACK;
            const int len = n_e - n_s;
            const bool keyed = len <= CACHE_KEY_MAX;
            n_dsb = 0;
            if (tile_docs) n_dsb = (ws.dsm[n_s >> 5] >> (n_s & 31)) & 1u;
            n_b0 = build_key(ws.text, sm.kmask, n_s, keyed ? len : 0, PFX ? n_dsb : 0u, n_key);
            n_a = make_uint4(0, 0, 0, 0);
            if (active && n_b0 != 0 && keyed && M.cache_ok) {
                n_i = cache_hash(n_key, J.cache_bits);
                ld_slot256(J.cache + n_i, n_a, n_b);
            }
        };
        if (n) prep(0);
#endif
        for (uint32_t r0 = 0; r0 < n; r0 += 32) {
            const uint32_t q = r0 + lane;
            const bool active = q < n;
#ifdef HUT_PIPE_PROBE
            const int s = n_s, e = n_e;
            const int len = e - s;
            const bool keyed = len <= CACHE_KEY_MAX;
            const uint32_t dsb = n_dsb;
            const uint32_t first = PFX ? dsb : 0u;
            uint32_t key[4] = {n_key[0], n_key[1], n_key[2], n_key[3]}, w[4] = {0u, 0u, 0u, 0u};
            const uint32_t b0 = n_b0;
            const bool dummy = !active || b0 == 0;
            const uint4 pa = n_a, pb = n_b;
            const uint32_t pi = n_i;
#else
            const int s = active ? ws.pstart[q] : BACK;
            const int e = active ? ws.pstart[q + 1] : BACK;
            const int len = e - s;                       // > MAXW when the end lies beyond the halo (BEYOND)
            const bool keyed = len <= CACHE_KEY_MAX;
            uint32_t dsb = 0;
            if (tile_docs) dsb = (ws.dsm[s >> 5] >> (s & 31)) & 1u;
            const uint32_t first = PFX ? dsb : 0u;
            uint32_t key[4], w[4] = {0u, 0u, 0u, 0u};
            const uint32_t b0 = build_key(ws.text, sm.kmask, s, keyed ? len : 0, first, key);
            const bool dummy = !active || b0 == 0;       // NUL boundary (zero fill past the end of the text)
#endif
            // state: 0 nothing, 1 ids ready, 2 cache miss merged by this warp (batches of 32), 3 long pretoken ->
            // long_kernel, 4 miss too long for the cache / the small per-lane arrays -> miss_kernel
            uint32_t state = 0;
            // pretokens of 16..31 bytes (rare): second cache, 32-byte keys
            bool hit2 = false;
            if (__any_sync(FULL, !keyed)) {
                const bool keyed2 = !dummy && !keyed && len <= CACHE2_KEY_MAX && M.cache_ok;
                if (keyed2) {
                    uint32_t k32[8];
                    build_key32(ws.text, sm.kmask32, s, len, first, k32);
                    hit2 = cache2_lookup(J, k32, w);
                }
            }
#ifdef HUT_PIPE_PROBE
            bool hit1 = false;
            if (!dummy && !hit2 && keyed && M.cache_ok) {
                if ((pa.x | pa.y) == 0) {
                    hit1 = false;
                } else if (pa.x == key[0] && pa.y == key[1] && pa.z == key[2] && pa.w == key[3]) {
                    w[0] = pb.x;
                    w[1] = pb.y;
                    w[2] = pb.z;
                    w[3] = pb.w;
                    hit1 = val_count(pb.x) != 0;
                } else {
                    hit1 = cache_lookup_from(J, key, w, (pi + 1) & ((1u << J.cache_bits) - 1), CACHE_PROBES - 1);
                }
            }
            if (r0 + 32 < n) prep(r0 + 32);
            if (!dummy) {
                if (hit2 || hit1) state = 1;
#else
            if (!dummy) {
                if (hit2 || (keyed && M.cache_ok && cache_lookup(J, key, w))) state = 1;
#endif
                else if ((uint32_t)len > M.miss_len_max) state = 3;
                else state = (len <= CACHE2_KEY_MAX &&
                              (uint32_t)len * M.kmax + (first ? M.n_prefix_syms : 0u) <= (uint32_t)MISS_SMALL) ? 2u : 4u;
            }
            // ---- counts -> offsets inside the tile's slot.  A miss reserves room for its worst case. ----
            uint32_t c = 0;
            if (state == 1) c = val_count(w[0]);
            else if (state == 2 || state == 4) c = (uint32_t)len * M.kmax + (first ? M.first_extra : 0u);
            uint32_t inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, d);
                if (lane >= d) inc += t;
            }
            const uint32_t round_total = __shfl_sync(FULL, inc, 31);
            const uint32_t off = run + inc - c;
            // ---- ids -> staging slot ----
            {
                int32_t* const p = slot + off;
                const uint32_t cs = state == 1 ? c : 0u;
                if (cs > 0) p[0] = sext21(w[0]);
                if (cs > 1) p[1] = sext21(w[1]);
                if (cs > 2) p[2] = sext21(w[2]);
                if (cs > 3) p[3] = sext21(w[3]);
                if (cs > 4) p[4] = val_id4(w);
            }
            // everything below up to the per-document offsets concerns pretokens the cache did not answer: one vote
            // decides for the whole round (99 % of the rounds skip it)
            if (__any_sync(FULL, state >= 2)) {
            // ---- cache misses -> this warp's queue; every 32 of them are merged at once, one per lane ----
            const uint32_t m2 = __ballot_sync(FULL, state == 2);
            if (m2) {
                holes = true;
                const unsigned long long po = (unsigned long long)(g0 + s) | ((unsigned long long)off << 40);
                if (state == 2)
                    ws.q[qn + __popc(m2 & lt_mask)] = make_uint4((uint32_t)po, (uint32_t)(po >> 32),
                                                                 (uint32_t)len | (first << 8) | (c << 16), tile);
                qn += __popc(m2);
                __syncwarp();
                while (qn >= 32) {
                    slack += run_batch(M, J, sm.byte1, ws.q, 32, tile);
                    st_wm += 32;
                    uint4 t = make_uint4(0, 0, 0, 0);
                    if (lane + 32 < qn) t = ws.q[lane + 32];
                    __syncwarp();
                    if (lane + 32 < qn) ws.q[lane] = t;
                    qn -= 32;
                    __syncwarp();
                }
            }
            // ---- the other misses -> the global miss list (one atomic per round) ----
            const uint32_t mm = __ballot_sync(FULL, state == 4);
            if (mm) {
                holes = true;
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&J.ctr->n_miss, (unsigned)__popc(mm));
                base = __shfl_sync(FULL, base, 0);
                if (state == 4) {
                    const uint32_t idx = base + __popc(mm & lt_mask);
                    if (idx < J.miss_cap) {
                        MissEntry E;
                        E.pos_off = (unsigned long long)(g0 + s) | ((unsigned long long)off << 40);
                        E.meta = (uint32_t)len | (first << 8) | (c << 16);
                        E.pad = 0;
                        J.miss[idx] = E;
                    } else {
                        set_flag(J, FLAG_MISS_CAP);
                    }
                }
            }
            // ---- long pretokens, per-document offsets (rare) ----
            uint32_t dm = __ballot_sync(FULL, state == 3);
            while (dm) {
                const int src = __ffs(dm) - 1;
                dm &= dm - 1;
                const uint32_t o = __shfl_sync(FULL, off, src);
                const int ss = __shfl_sync(FULL, s, src), ee = __shfl_sync(FULL, e, src);
                const uint32_t ff = __shfl_sync(FULL, first, src);
                if (lane == 0) {
                    defer_push(J, tile, (unsigned long long)(g0 + ss), o, ee == (int)BEYOND ? 0u : (uint32_t)(ee - ss),
                               ff, def_head, def_tail);
                    st_def++;
                }
                def_head = __shfl_sync(FULL, def_head, 0);
                def_tail = __shfl_sync(FULL, def_tail, 0);
            }
            }
            if (tile_docs && !dummy && dsb) {  // documents starting at this pretoken: slot entries before it
                const unsigned long long gpos = (unsigned long long)(g0 + s);
                unsigned long long d = dfirst;
                while (d < J.ndocs && J.doc_off[d] < gpos) ++d;
                while (d < J.ndocs && J.doc_off[d] == gpos) J.tok_off[d++] = off;
            }
            st_pre += dummy ? 0u : 1u;
            st_hit += state == 1 ? 1u : 0u;
            run += round_total;
        }
        if (lane == 0) {
            J.cnt[tile] = run | (def_head != NODOC ? CNT_CHAIN : 0u) | (holes ? CNT_HOLES : 0u);
            J.tot[tile] = run - slack;
            J.def_first[tile] = def_head;
        }
    }
    // ---- the misses still queued ----
    __syncwarp();
    if (qn) {
        run_batch(M, J, sm.byte1, ws.q, qn, NODOC);
        st_wm += qn;
    }
    // ---- statistics ----
    unsigned long long sum_pre = st_pre, sum_hit = st_hit;
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        sum_pre += __shfl_xor_sync(FULL, sum_pre, d);
        sum_hit += __shfl_xor_sync(FULL, sum_hit, d);
    }
    if (lane == 0) {
        Counters* c = J.ctr;
        if (sum_pre) atomicAdd(&c->st_pretokens, sum_pre);
        if (sum_hit) atomicAdd(&c->st_hits, sum_hit);
        if (st_def) atomicAdd(&c->st_block, st_def);
        if (st_wm) atomicAdd(&c->st_warp, st_wm);
    }
}


// ------------------------------------------------------------------------------------------------
// miss_kernel: LANE per entry of the global miss list = the cache misses tile_kernel did not merge itself
// (pretokens too long for the cache key or for its small per-lane arrays).  See merge_entry.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(MNT) miss_kernel(const DevModel M, const Job J) {
    __shared__ uint32_t byte1[256];
    const uint32_t tid = threadIdx.x;
    for (uint32_t i = tid; i < 256; i += MNT) byte1[i] = M.byte1[i];
    __syncthreads();
    uint32_t hi = J.ctr->n_miss;
    if (hi > J.miss_cap) hi = J.miss_cap;
    uint32_t n_done = 0;
    for (uint32_t ei = blockIdx.x * MNT + tid; ei < hi; ei += gridDim.x * MNT) {
        const MissEntry E = J.miss[ei];
        const unsigned long long pos = E.pos_off & ((1ull << 40) - 1);
        const uint32_t loff = (uint32_t)(E.pos_off >> 40);
        const uint32_t len = E.meta & 0xFFu, first = (E.meta >> 8) & 1u, room = E.meta >> 16;
        const unsigned long long tile = pos / OWN;
        const uint32_t slack = merge_entry<MISS_SYMS>(M, J, byte1, pos, len, first, room,
                                                      stage_slot(J, tile) + loff);
        if (slack) atomicSub(&J.tot[tile], slack);
        n_done++;
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) n_done += __shfl_xor_sync(0xFFFFFFFFu, n_done, d);
    if ((tid & 31) == 0 && n_done) atomicAdd(&J.ctr->st_warp, (unsigned long long)n_done);
}

// ------------------------------------------------------------------------------------------------
// long_kernel: one CTA per deferred pretoken (longer than 32 bytes / symbols, or running past the tile's halo)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(LNT) long_kernel(const DevModel M, const Job J) {
    __shared__ LongSm sm;
    const uint32_t tid = threadIdx.x;
    for (uint32_t i = tid; i < 256; i += LNT) sm.byte1[i] = M.byte1[i];
    __syncthreads();
    uint32_t n = J.ctr->n_def;
    if (n > J.def_cap) n = J.def_cap;
    uint32_t* const scratch = J.cta_scratch + (unsigned long long)blockIdx.x * J.cta_scratch_words;
    unsigned long long n_big = 0;
    for (uint32_t e = blockIdx.x; e < n; e += gridDim.x) {
        DefEntry* E = J.def + e;
        const unsigned long long gs = E->pos;
        const bool first = (E->flags & 1u) != 0;
        __syncthreads();
        if (tid == 0) {  // end of the document containing gs: first doc_off > gs
            unsigned long long a = 0, b = J.ndocs + 1;
            while (a < b) {
                unsigned long long mid = (a + b) >> 1;
                if (J.doc_off[mid] <= gs) a = mid + 1;
                else b = mid;
            }
            unsigned long long de = a <= J.ndocs ? J.doc_off[a] : J.nbytes;
            sm.doc_end = de < J.nbytes ? de : J.nbytes;
        }
        __syncthreads();
        const unsigned long long dend = sm.doc_end;
        unsigned long long ge;
        if (E->len) {
            ge = gs + E->len;
        } else {
            uint32_t cls = gclass(J, gs, gs, dend);
            if (cls == HC_SPACE) {
                const uint32_t c1 = gclass(J, gs + 1, gs, dend);
                if (c1 >= HC_ALPHA && c1 <= HC_OTHER) cls = c1;
            }
            ge = find_run_end(J, sm, gs + 1, gs, dend, cls, gs + M.limit + 1);
        }
        const uint32_t b0 = gbyte(J, gs, dend);
        const bool with_prefix = first && b0 != ' ';
        const bool alone_first = first && b0 == ' ';
        if (ge - gs > M.limit) {
            if (M.compat_cut) {  // core.c:402-407,503: the document ends here
                if (tid == 0) {
                    E->flags |= 2u;
                    E->count = 0;
                }
            } else {
                set_flag(J, FLAG_BIG_CAP);
            }
            continue;
        }
        const unsigned long long bound = (ge - gs) * M.kmax + M.n_prefix_syms;
        const bool big = bound > J.med_cap;
        uint32_t *U, *MG, *NX, *PV, cap;
        if (!big) {
            U = scratch;
            MG = U + J.med_cap;
            NX = MG + J.med_cap;
            PV = NX + J.med_cap;
            cap = J.med_cap;
        } else {
            if (tid == 0) {
                uint32_t slot = blockIdx.x % J.big_slots;
                for (;;) {
                    if (atomicCAS(&J.ctr->pool_lock[slot], 0u, 1u) == 0u) break;
                    slot = (slot + 1) % J.big_slots;
                    __nanosleep(200);
                }
                sm.slot = slot;
            }
            __syncthreads();
            U = J.big_pool + (unsigned long long)sm.slot * J.big_slot_words;
            MG = U + J.big_cap;
            NX = MG + J.big_cap;
            PV = NX + J.big_cap;
            cap = J.big_cap;
            n_big++;
        }
        // tournament-tree levels behind the four arrays (8-byte aligned; the areas hold 5 * cap + 8 words)
        unsigned long long* T = reinterpret_cast<unsigned long long*>(
            (reinterpret_cast<uintptr_t>(PV + cap) + 7u) & ~(uintptr_t)7u);
        uint32_t m = 0;
        uint32_t cnt = cta_merge_run(M, J, sm, gs, ge, with_prefix, alone_first, U, MG, NX, PV, T, cap, &m);
        if (cnt == 0xFFFFFFFFu) {
            set_flag(J, FLAG_BIG_CAP);
            cnt = 0;
            m = 0;
        }
        __syncthreads();
        if (tid == 0) sm.ids_off = atomicAdd(&J.ctr->long_used, (unsigned long long)cnt);
        __syncthreads();
        const unsigned long long off = sm.ids_off;
        if (off + cnt > J.long_cap) {
            set_flag(J, FLAG_LONG_CAP);
        } else if (cnt) {
            cta_merge_emit(M, sm, U, m, alone_first, J.long_ids + off);
        }
        __syncthreads();
        if (tid == 0) {
            E->count = cnt;
            E->ids_off = off;
            atomicAdd(&J.tot[E->tile], cnt);
            if (big) {
                __threadfence();
                atomicExch(&J.ctr->pool_lock[sm.slot], 0u);
            }
        }
    }
    if (tid == 0 && n_big) atomicAdd(&J.ctr->st_big, n_big);
}

// ------------------------------------------------------------------------------------------------
// exclusive scan of the per-tile id counts: scan1 per block of SCAN_BLK tiles, scan2 over the block sums
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) scan1_kernel(Job J) {
    __shared__ uint32_t wsum[32];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long base = (unsigned long long)blockIdx.x * SCAN_BLK + tid * 4ull;
    uint32_t v[4], sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[i] = base + i < J.ntiles ? J.tot[base + i] : 0u;
        sum += v[i];
    }
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = wsum[lane];
        uint32_t winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, winc, d);
            if (lane >= d) winc += t;
        }
        wsum[lane] = winc - w;
        if (lane == 31) J.bsum[blockIdx.x] = winc;
    }
    __syncthreads();
    uint32_t run = wsum[warp] + inc - sum;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (base + i < J.ntiles) J.tot[base + i] = run;
        run += v[i];
    }
}

__global__ void __launch_bounds__(1024) scan2_kernel(Job J) {
    __shared__ unsigned long long wsum[32];
    __shared__ unsigned long long carry_s;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t b0 = 0; b0 < J.nblocks; b0 += 1024) {
        const uint32_t i = b0 + tid;
        const unsigned long long v = i < J.nblocks ? J.bsum[i] : 0ull;
        unsigned long long inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const unsigned long long w = wsum[lane];
            unsigned long long winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, winc, d);
                if (lane >= d) winc += t;
            }
            wsum[lane] = winc - w;
        }
        __syncthreads();
        const unsigned long long carry = carry_s;
        const unsigned long long excl = carry + wsum[warp] + inc - v;
        if (i < J.nblocks) J.bsum[i] = excl;
        __syncthreads();
        if (tid == 1023) carry_s = excl + v;
        __syncthreads();
    }
    if (tid == 0) J.ctr->total_ids = carry_s;
}

__device__ __forceinline__ unsigned long long tile_base(const Job& J, unsigned long long t) {
    return J.bsum[t / SCAN_BLK] + J.tot[t];
}

// warp copy of n ids into the dense output at position o
__device__ __forceinline__ void copy_ids(const Job& J, const int32_t* src, unsigned long long o, uint32_t n) {
    const uint32_t lane = threadIdx.x & 31;
    if (o + n > J.out_cap) {
        if (lane == 0) set_flag(J, FLAG_OUT_CAP);
        return;
    }
    for (uint32_t base = 0; base < n; base += 128) {
        int32_t v[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t i = base + 32 * g + lane;
            v[g] = i < n ? ld_stream_s32(src + i) : 0;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t i = base + 32 * g + lane;
            if (i < n) st_stream_s32(J.out + o + i, v[g]);
        }
    }
}

// warp copy of the non-HOLE entries of src[0 .. n) to the dense output at position o; returns how many there were.
// Four 128-byte loads are in flight per step.
__device__ __forceinline__ uint32_t copy_ids_filtered(const Job& J, const int32_t* src, unsigned long long o, uint32_t n) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t kept = 0;
    for (uint32_t base = 0; base < n; base += 128) {
        int32_t v[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t i = base + 32 * g + lane;
            v[g] = i < n ? ld_stream_s32(src + i) : HOLE;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const bool keep = v[g] != HOLE;
            const uint32_t mask = __ballot_sync(0xFFFFFFFFu, keep);
            if (keep) {
                const unsigned long long at = o + kept + __popc(mask & lt);
                if (at < J.out_cap) st_stream_s32(J.out + at, v[g]);
                else set_flag(J, FLAG_OUT_CAP);
            }
            kept += __popc(mask);
        }
    }
    return kept;
}

// The same for a whole tile whose slot holds HOLEs, and the documents starting in the tile get their final id offset
// on the way: tile_kernel left the slot position of their first pretoken in tok_off, which is translated to the
// number of ids kept before it.
__device__ __forceinline__ void compact_tile_holes(const Job& J, uint32_t tile, const int32_t* src, unsigned long long o,
                                                   uint32_t n) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const unsigned long long tile_end = ((unsigned long long)tile + 1) * OWN;
    unsigned long long d = J.first_doc[tile];
    uint32_t so = 0xFFFFFFFFu;  // slot position of document d's first pretoken; none left: 0xFFFFFFFF
    auto load_doc = [&]() {
        so = 0xFFFFFFFFu;
        if (d != NODOC && d < J.ndocs) {
            const unsigned long long pos = J.doc_off[d];
            if (pos < tile_end && pos < J.nbytes) so = (uint32_t)J.tok_off[d];
        }
    };
    load_doc();
    uint32_t kept = 0;
    for (uint32_t base = 0; base < n; base += 128) {
        int32_t v[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t i = base + 32 * g + lane;
            v[g] = i < n ? ld_stream_s32(src + i) : HOLE;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const uint32_t gbase = base + 32 * g;
            const bool keep = v[g] != HOLE;
            const uint32_t mask = __ballot_sync(0xFFFFFFFFu, keep);
            while (so < gbase + 32) {  // warp-uniform
                if (lane == 0) J.tok_off[d] = o + kept + __popc(mask & ((1u << (so - gbase)) - 1u));
                ++d;
                load_doc();
            }
            if (keep) {
                const unsigned long long at = o + kept + __popc(mask & lt);
                if (at < J.out_cap) st_stream_s32(J.out + at, v[g]);
                else set_flag(J, FLAG_OUT_CAP);
            }
            kept += __popc(mask);
        }
    }
    while (so != 0xFFFFFFFFu) {  // documents starting behind the last entry of the slot
        if (lane == 0) J.tok_off[d] = o + kept;
        ++d;
        load_doc();
    }
}

// staging slots -> dense ids (warp per tile); drops the HOLEs left by deferred cache misses and splices in the ids
// of the pretokens long_kernel merged
__global__ void __launch_bounds__(256, 8) compact_kernel(const Job J) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), nw = gridDim.x * (blockDim.x >> 5);
    for (uint32_t tile = gw; tile < J.ntiles; tile += nw) {
        const uint32_t c = J.cnt[tile];
        const uint32_t ns = c & CNT_MASK;
        unsigned long long o = tile_base(J, tile);
        const int32_t* src = stage_slot(J, tile);
        if (!(c & (CNT_CHAIN | CNT_HOLES))) {
            copy_ids(J, src, o, ns);
            continue;
        }
        if (!(c & CNT_CHAIN)) {
            compact_tile_holes(J, tile, src, o, ns);
            continue;
        }
        uint32_t prev = 0;
        for (uint32_t idx = J.def_first[tile]; idx != NODOC;) {
            const DefEntry E = J.def[idx];
            o += copy_ids_filtered(J, src + prev, o, E.local_off - prev);
            prev = E.local_off;
            if (E.flags & 2u) {  // reference-compatible cut: everything from here to the end of the document goes
                if (lane == 0) {
                    const uint32_t ev = atomicAdd(&J.ctr->n_events, 1u);
                    if (ev < (uint32_t)MAX_EVENTS) {
                        J.ctr->events[ev].tok_index = o;
                        J.ctr->events[ev].pos = E.pos;
                    } else {
                        set_flag(J, FLAG_EVENTS);
                    }
                }
            } else if (E.count && E.ids_off + E.count <= J.long_cap) {
                copy_ids(J, J.long_ids + E.ids_off, o, E.count);
                o += E.count;
            }
            idx = E.pad;
        }
        copy_ids_filtered(J, src + prev, o, ns - prev);
    }
}

// tile-local document offsets -> global ones
__global__ void fixdoc_kernel(const Job J) {
    const unsigned long long d = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (d > J.ndocs) return;
    const unsigned long long total = J.ctr->total_ids;
    const unsigned long long pos = d < J.ndocs ? J.doc_off[d] : J.nbytes;
    if (d == J.ndocs || pos >= J.nbytes) {
        J.tok_off[d] = total;
        return;
    }
    const unsigned long long t = pos / OWN;
    const uint32_t c = J.cnt[t];
    if ((c & (CNT_HOLES | CNT_CHAIN)) == CNT_HOLES) return;  // compact_kernel has translated these already
    unsigned long long v = J.tok_off[d];  // entries of the tile's slot before the document
    if (c & CNT_HOLES) {
        const int32_t* src = stage_slot(J, t);
        uint32_t h = 0;
        for (uint32_t i = 0; i < (uint32_t)v; ++i) h += src[i] == HOLE;
        v -= h;
    }
    v += tile_base(J, t);
    if (c & CNT_CHAIN)
        for (uint32_t idx = J.def_first[t]; idx != NODOC; idx = J.def[idx].pad)
            if (J.def[idx].pos < pos) v += J.def[idx].count;
    J.tok_off[d] = v;
}

// ------------------------------------------------------------------------------------------------
// reference-compatible document cuts (rare): drop the ids of [tok_index, end of document) per event
// ------------------------------------------------------------------------------------------------
struct CutRanges {
    unsigned int n;
    unsigned int pad;
    unsigned long long new_total;
    unsigned long long x[MAX_EVENTS], e[MAX_EVENTS], del_before[MAX_EVENTS + 1];
};

__global__ void cut_prepare_kernel(Job J, CutRanges* R) {
    if (threadIdx.x || blockIdx.x) return;
    Counters* c = J.ctr;
    unsigned int n = c->n_events < (unsigned)MAX_EVENTS ? c->n_events : (unsigned)MAX_EVENTS;
    // end of each event's document in id space
    for (unsigned int i = 0; i < n; ++i) {
        unsigned long long pos = c->events[i].pos, a = 0, b = J.ndocs + 1;
        while (a < b) {
            unsigned long long mid = (a + b) >> 1;
            if (J.doc_off[mid] <= pos) a = mid + 1;
            else b = mid;
        }
        R->x[i] = c->events[i].tok_index;
        R->e[i] = J.tok_off[a <= J.ndocs ? a : J.ndocs];
    }
    for (unsigned int i = 1; i < n; ++i) {  // insertion sort by x
        unsigned long long x = R->x[i], e = R->e[i];
        unsigned int j = i;
        while (j && R->x[j - 1] > x) {
            R->x[j] = R->x[j - 1];
            R->e[j] = R->e[j - 1];
            --j;
        }
        R->x[j] = x;
        R->e[j] = e;
    }
    unsigned int m = 0;  // merge overlapping
    for (unsigned int i = 0; i < n; ++i) {
        if (R->e[i] <= R->x[i]) continue;
        if (m && R->x[i] <= R->e[m - 1]) {
            if (R->e[i] > R->e[m - 1]) R->e[m - 1] = R->e[i];
        } else {
            R->x[m] = R->x[i];
            R->e[m] = R->e[i];
            ++m;
        }
    }
    unsigned long long del = 0;
    for (unsigned int i = 0; i < m; ++i) {
        R->del_before[i] = del;
        del += R->e[i] - R->x[i];
    }
    R->del_before[m] = del;
    R->n = m;
    R->new_total = c->total_ids - del;
    c->st_cut = m;
}

__device__ __forceinline__ unsigned long long cut_map(const CutRanges* R, unsigned long long i, bool* deleted) {
    unsigned int a = 0, b = R->n;  // number of ranges with x <= i
    while (a < b) {
        unsigned int mid = (a + b) >> 1;
        if (R->x[mid] <= i) a = mid + 1;
        else b = mid;
    }
    *deleted = false;
    if (a == 0) return i;
    if (i < R->e[a - 1]) {
        *deleted = true;
        return R->x[a - 1] - R->del_before[a - 1];
    }
    return i - R->del_before[a];
}

__global__ void cut_apply_ids_kernel(const int32_t* src, int32_t* dst, unsigned long long n, const CutRanges* R) {
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        bool del;
        unsigned long long j = cut_map(R, i, &del);
        if (!del) dst[j] = src[i];
    }
}

__global__ void cut_apply_offs_kernel(unsigned long long* tok_off, unsigned long long n, const CutRanges* R) {
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        bool del;
        tok_off[i] = cut_map(R, tok_off[i], &del);
    }
}

}  // namespace hut
