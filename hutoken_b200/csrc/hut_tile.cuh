/* sm_100a device code of the BPE encode engine, part 2: the fused tile kernel (see hut_kernels.cuh for the pipeline).
 * Included only by hut_engine.cu (through hut_pipeline.cuh). */
#pragma once

#include "hut_kernels.cuh"

#ifndef HUT_TILE_CTAS
#define HUT_TILE_CTAS 4  // resident CTAs per SM the tile kernel is compiled for (64 registers per thread)
#endif

namespace hut {

#define HUT_RARE(x) __builtin_expect(!!(x), 0)    // keeps the code behind it out of the hot path's cache lines
#define HUT_USUAL(x) __builtin_expect(!!(x), 1)

// Cache misses are merged 32 at a time, one per lane, and a batch takes as long as its longest pretoken (the merge
// loop is quadratic in the symbols).  They therefore wait in one queue per length class, so that the many short
// ones (numbers, short rare words) are not held up by a long word in every batch.
#ifndef HUT_MISS_CLASSES
#define HUT_MISS_CLASSES 1
#endif
#ifndef HUT_MISS_SHORT
#define HUT_MISS_SHORT 7
#endif
constexpr int MISS_CLASSES = HUT_MISS_CLASSES;
constexpr int MISS_SHORT = HUT_MISS_SHORT;  // longest pretoken (bytes) of class 0

// shared memory of ONE WARP of tile_kernel
struct __align__(16) WarpSm {
    uint8_t text[TEXTB];        // span bytes: [tile_start - 32, tile_start + 992 + 16), then zero padding
    uint32_t dsm[CHUNK + 4];    // per chunk: document starts
    uint16_t pstart[SPAN + 8];  // span positions of the boundaries in chunks 1..31 (31 = halo), then the end sentinel
    unsigned long long q[MISS_CLASSES][64];  // cache misses waiting for a full batch (by length class), one word each: see q_pack
    unsigned long long mbar;
    // what only the rounds with cache misses / long pretokens touch lives here, not in registers (warp-uniform values,
    // written by lane 0 or by all lanes alike)
    uint32_t qn[MISS_CLASSES];  // entries of q[]
    uint32_t slack;             // room reserved for misses of the tile in work that stayed unused
    uint32_t holes;             // the tile in work has reserved room
    uint32_t def_head, def_tail;  // chain of the tile's pretokens handed to long_kernel
    uint32_t st_pre, st_miss, st_def, st_pad;  // statistics of this warp
};


struct TileSm {
    uint4 kmask[16];            // kmask[len]: byte masks of a 16-byte key holding len text bytes
    uint32_t byte1[256];
    WarpSm w[WPB];
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

// A queued miss in one word: byte position (38 bits: 256 GiB per shard), position of its room in the tile's staging
// slot (18 bits), length in bytes (6 bits), "first pretoken of its document" (1 bit).  Its tile is position / OWN, its
// room the worst case tile_kernel reserved: one id per symbol (+ the prefix ids of a first pretoken).
__device__ __forceinline__ unsigned long long q_pack(unsigned long long pos, uint32_t loff, uint32_t len, uint32_t first) {
    return pos | ((unsigned long long)loff << 38) | ((unsigned long long)len << 56) | ((unsigned long long)first << 62);
}

// merges the first n (<= 32) queued cache misses, one per lane (merge_entry).  The room a pretoken of another tile
// did not use is subtracted from that tile's count here; for the tile in progress it is returned (warp sum).
__device__ __forceinline__ uint32_t run_batch(const DevModel& M, const Job& J, const uint32_t* byte1, const unsigned long long* q,
                                           uint32_t n, uint32_t cur_tile) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t slack = 0;
    if (lane < n) {
        const unsigned long long E = q[lane];
        const unsigned long long pos = E & ((1ull << 38) - 1);
        const uint32_t loff = (uint32_t)(E >> 38) & 0x3FFFFu;
        const uint32_t len = (uint32_t)(E >> 56) & 63u, first = (uint32_t)(E >> 62) & 1u;
        const uint32_t room = len * M.kmax + (first ? M.first_extra : 0u);
        const uint32_t tile = (uint32_t)(pos / (unsigned long long)OWN);
        slack = merge_entry<MISS_SMALL>(M, J, byte1, pos, len, first, room, stage_slot(J, tile) + loff);
        if (slack && tile != cur_tile) {
            atomicSub(&J.tot[tile], slack);
            slack = 0;
        }
    }
    __syncwarp();
#pragma unroll
    for (int d = 16; d; d >>= 1) slack += __shfl_xor_sync(0xFFFFFFFFu, slack, d);
    return slack;
}


__device__ __forceinline__ void put_hit_ids(int32_t* p, uint32_t c, const uint32_t w[4]) {
    if (c > 0) p[0] = sext21(w[0]);
    if (c > 1) p[1] = sext21(w[1]);
    if (c > 2) p[2] = sext21(w[2]);
    if (c > 3) p[3] = sext21(w[3]);
    if (c > 4) p[4] = val_id4(w);
}

// documents starting at a pretoken of this tile: their id offset = entries of the tile's slot before the pretoken
// (fixdoc_kernel / compact_kernel turn that into the global offset)
__device__ __forceinline__ void doc_offsets_at(const Job& J, unsigned long long gpos, unsigned long long dfirst, uint32_t v) {
    unsigned long long d = dfirst;
    while (d < J.ndocs && J.doc_off[d] < gpos) ++d;
    while (d < J.ndocs && J.doc_off[d] == gpos) J.tok_off[d++] = v;
}

// The part of a round that concerns pretokens the cache did not answer (state 2: cache miss, 3: long pretoken).  All it
// shares with the hot loop beyond the arguments lives in the warp's shared-memory slice, so it costs the loop no
// registers.  (Inlined on purpose: a call inside the loop makes the compiler keep the loop's variables in local
// memory -- 140 bytes of spills against none.)
__device__ __forceinline__ void rare_round(const DevModel& M, const Job& J, TileSm& sm, WarpSm& ws, uint32_t tile, int s, int e,
                                        uint32_t first, uint32_t c, uint32_t off, uint32_t state) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const long long g0 = (long long)tile * OWN - BACK;
    const int len = e - s;
    {
        // ---- cache misses -> this warp's queue; every 32 of them -- of this tile and of earlier ones -- are
        // merged at once, one per lane (merge_entry), so that their table probes overlap; what a miss leaves
        // unused of its room stays behind as HOLEs, which compact_kernel drops ----
        if (__any_sync(FULL, state == 2)) {
            uint32_t slack = 0, nb = 0;
            const unsigned long long po = q_pack((unsigned long long)(g0 + s), off, (uint32_t)len, first);
            const uint32_t mycls = (MISS_CLASSES > 1 && len > MISS_SHORT) ? 1u : 0u;
#pragma unroll 1
            for (uint32_t cls = 0; cls < (uint32_t)MISS_CLASSES; ++cls) {  // one copy of the merge code for all classes
                const uint32_t m2 = __ballot_sync(FULL, state == 2 && mycls == cls);
                if (!m2) continue;
                unsigned long long* const q = ws.q[cls];
                uint32_t qn = ws.qn[cls];
                if (state == 2 && mycls == cls)
                    q[qn + __popc(m2 & lt_mask)] = po;
                qn += __popc(m2);
                __syncwarp();
                while (qn >= 32) {
                    slack += run_batch(M, J, sm.byte1, q, 32, tile);
                    nb += 32;
                    unsigned long long t = 0;
                    if (lane + 32 < qn) t = q[lane + 32];
                    __syncwarp();
                    if (lane + 32 < qn) q[lane] = t;
                    qn -= 32;
                    __syncwarp();
                }
                if (lane == 0) ws.qn[cls] = qn;
            }
            if (lane == 0) {
                ws.holes = 1;
                ws.slack += slack;
                ws.st_miss += nb;
            }
            __syncwarp();
        }
        // ---- long pretokens -> side list, in lane order ----
        uint32_t dm = __ballot_sync(FULL, state == 3);
        if (dm) {
            uint32_t def_head = ws.def_head, def_tail = ws.def_tail;
            const uint32_t nd = __popc(dm);
            while (dm) {
                const int src = __ffs((int)dm) - 1;
                dm &= dm - 1;
                const uint32_t o = __shfl_sync(FULL, off, src);
                const int ss = __shfl_sync(FULL, s, src), ee = __shfl_sync(FULL, e, src);
                const uint32_t ff = __shfl_sync(FULL, first, src);
                if (lane == 0)
                    defer_push(J, tile, (unsigned long long)(g0 + ss), o,
                               ee == (int)BEYOND ? 0u : (uint32_t)(ee - ss), ff, def_head, def_tail);
            }
            if (lane == 0) {
                ws.def_head = def_head;
                ws.def_tail = def_tail;
                ws.st_def += nd;
            }
            __syncwarp();
        }
    }
}

// PFX: the model has a prefix (first pretoken of a document keyed apart and merged with the prefix symbols; staging
// slots of uneven size read from J.slot_off); without one the slot of a tile is tile * stride.
// DYN: tiles are handed out by a ticket counter (the ticket of the next tile is asked for while the current one is in
// work), so that warps on faster SMs take more of them; otherwise warp gw takes tiles gw, gw + nw, ...
template <bool PFX, bool DYN>
__device__ __forceinline__ void tile_body(const DevModel& M, const Job& J, TileSm& sm, const uint32_t gw, const uint32_t nw) {
    const int tid = threadIdx.x, warp = tid >> 5;
    int lane;  // read once from the special register (kept from being recomputed from %tid inside the loops)
    asm volatile("mov.u32 %0, %%laneid;" : "=r"(lane));
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t le_mask = 0xFFFFFFFFu >> (31 - lane);  // lanes up to and including this one
    WarpSm& ws = sm.w[warp];

    for (int i = tid; i < 256; i += NT) sm.byte1[i] = M.byte1[i];
    if (tid < 16) {
        auto wm = [&](int rem) -> uint32_t { return rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : ((1u << (8 * rem)) - 1u)); };
        sm.kmask[tid] = make_uint4(wm(tid), wm(tid - 4), wm(tid - 8), wm(tid - 12));
    }
    if (lane == 0) {
        mbar_init(&ws.mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();  // the only block-wide barrier: the tables above are read-only afterwards

    const uint32_t hshift = 32u - J.cache_bits;
    const uint32_t probe_len = M.cache_ok ? (uint32_t)CACHE_KEY_MAX : 0u;  // longest pretoken looked up in the cache
    uint32_t parity = 0;
    if (lane == 0) {
        ws.st_pre = ws.st_miss = ws.st_def = 0;
        for (int cls = 0; cls < MISS_CLASSES; ++cls) ws.qn[cls] = 0;
    }
    __syncwarp();

    uint32_t tile = gw;
    if (DYN) {
        if (lane == 0) tile = atomicAdd(&J.ctr->tile_ticket, 1u);
        tile = __shfl_sync(FULL, tile, 0);
    }
    while (tile < J.ntiles) {
        uint32_t next_tile = tile + nw;
        if (DYN && lane == 0) next_tile = atomicAdd(&J.ctr->tile_ticket, 1u);  // answered long before the tile is done
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
#pragma unroll 1
        for (int li = lane; li < lo_li; li += 32) ws.text[li] = 0;
#pragma unroll 1
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
            if (HUT_RARE(__any_sync(FULL, mid.EX != 0))) generic_chunk(ws.text, ws.dsm, lane, &ST, &NU);
        }
        uint32_t bnd = ST | NU;
        if (lane == 31) bnd &= (1u << (TRUST - 31 * CHUNK)) - 1u;
        // the text ends at span position E = nbytes - g0 (> 32): the zero fill behind it reads as NUL boundaries, of
        // which only the one AT E is kept (it ends the last pretoken)
        {
            const long long rem = (long long)J.nbytes - (g0 + lane * CHUNK);  // text bytes from this chunk's start on
            const uint32_t valid = rem >= 32 ? FULL : (rem <= 0 ? 0u : ((1u << rem) - 1u));
            if (HUT_RARE(lane >= 1 && lane <= 30 && (NU & valid))) set_flag(J, FLAG_NUL);  // a NUL inside the text
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
        uint32_t run = 0;                       // ids of this tile written to its staging slot so far
        if (lane == 0) {
            ws.slack = 0;
            ws.holes = 0;
            ws.def_head = ws.def_tail = NODOC;
        }
        __syncwarp();
        int32_t* const slot = J.stage + (PFX ? slot_o : (unsigned long long)tile * J.stride);
        const uint32_t hmask = (1u << J.cache_bits) - 1u;
        for (uint32_t r0 = 0; r0 < n; r0 += 32) {
            // ---- one pretoken per lane: key -> one 32-byte probe of the pretoken cache.  The code from here to the
            // `continue` of the all-answered case is the hot path of the engine and is kept short on purpose: the SM's
            // instruction cache holds little more than this loop, the classifier and one copy of the merge code ----
            const uint32_t q = r0 + lane;
            const bool active = q < n;
            int s = BACK, e = BACK;
            if (active) {
                s = ws.pstart[q];
                e = ws.pstart[q + 1];
            }
            const int len = e - s;                   // > MAXW when the end lies beyond the halo (BEYOND); 0 on idle lanes
            const bool keyed = len <= CACHE_KEY_MAX;
            uint32_t dsb = 0;
            if (tile_docs) dsb = (ws.dsm[s >> 5] >> (s & 31)) & 1u;
            uint32_t key[4];
            const uint32_t b0 = build_key(ws.text, sm.kmask, s, keyed ? len : 0, PFX ? dsb : 0u, key);
            const uint32_t slot_i = cache_hash_raw(key) >> hshift;
            // an idle lane, a long pretoken or a disabled cache ask for nothing: the slot then reads as all zero, which no
            // key equals (the meta byte of a key is never zero)
            uint4 sa = make_uint4(0u, 0u, 0u, 0u), sb = make_uint4(0u, 0u, 0u, 0u);
            if (active && (uint32_t)len <= probe_len) ld_slot256(J.cache + slot_i, sa, sb);
            bool hit = sa.x == key[0] && sa.y == key[1] && sa.z == key[2] && sa.w == key[3];
            if (HUT_RARE(!hit && (sa.x | sa.y) != 0u)) {       // the slot holds another pretoken: walk on (few lanes, few steps)
                uint32_t i = slot_i;
#pragma unroll 1
                for (int p = 1; p < CACHE_PROBES; ++p) {
                    i = (i + 1u) & hmask;
                    ld_slot256(J.cache + i, sa, sb);
                    if ((sa.x | sa.y) == 0u) break;
                    hit = sa.x == key[0] && sa.y == key[1] && sa.z == key[2] && sa.w == key[3];
                    if (hit) break;
                }
            }
            uint32_t c = hit ? val_count(sb.x) : 0u;  // count 0: the value is still being written -> a miss
            if (HUT_USUAL(!__any_sync(FULL, active && c == 0u))) {
                // ---- every pretoken of the round was answered by the cache: counts are 1..5, three ballots scan them ----
                const uint32_t m0 = __ballot_sync(FULL, (c & 1u) != 0u), m1 = __ballot_sync(FULL, (c & 2u) != 0u),
                               m2 = __ballot_sync(FULL, (c & 4u) != 0u);
                const uint32_t incl = __popc(m0 & le_mask) + 2u * __popc(m1 & le_mask) + 4u * __popc(m2 & le_mask);
                const uint32_t off = run + incl - c;
                const uint32_t w[4] = {sb.x, sb.y, sb.z, sb.w};
                if (c) put_hit_ids(slot + off, c, w);
                if (HUT_RARE(tile_docs && active && dsb)) doc_offsets_at(J, (unsigned long long)(g0 + s), dfirst, off);
                run += __shfl_sync(FULL, incl, 31);
                continue;
            }
            // ---- some pretokens were not answered.  Pretokens of 16..31 bytes ask the second cache (32-byte keys) first ----
            uint32_t w[4] = {sb.x, sb.y, sb.z, sb.w};
            if (HUT_RARE(active && !keyed && len <= CACHE2_KEY_MAX && M.cache_ok)) {
                uint32_t k32[8];
                build_key32(ws.text, s, len, PFX ? dsb : 0u, k32);
                hit = cache2_lookup(J, k32, w);
                c = hit ? val_count(w[0]) : 0u;
            }
            // state of a lane's pretoken: 0 nothing / answered, 2 cache miss merged by this warp (batches of 32), 3 long
            // pretoken -> long_kernel.  A miss reserves room for its worst case (one id per symbol + prefix ids).
            {
                const uint32_t first = PFX ? dsb : 0u;
                uint32_t state = 0, cc = c;
                if (active && c == 0u && b0 != 0) {  // b0 == 0: a NUL in the text, the call fails (FLAG_NUL)
                    const bool lng = (uint32_t)len > M.miss_len_max ||
                                     (uint32_t)len * M.kmax + (first ? M.n_prefix_syms : 0u) > (uint32_t)MISS_SMALL;
                    state = lng ? 3u : 2u;
                    if (!lng) cc = (uint32_t)len * M.kmax + (first ? M.first_extra : 0u);
                }
                uint32_t inc = cc;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t t = __shfl_up_sync(FULL, inc, d);
                    if (lane >= d) inc += t;
                }
                const uint32_t round_total = __shfl_sync(FULL, inc, 31);
                const uint32_t off = run + inc - cc;     // position of this lane's ids (or room) in the tile's slot
                if (c) put_hit_ids(slot + off, c, w);
                if (__any_sync(FULL, state != 0)) rare_round(M, J, sm, ws, tile, s, e, first, cc, off, state);
                if (HUT_RARE(tile_docs && active && dsb)) doc_offsets_at(J, (unsigned long long)(g0 + s), dfirst, off);
                run += round_total;
            }
        }
        if (lane == 0) {
            const uint32_t def_head = ws.def_head;
            J.cnt[tile] = run | (def_head != NODOC ? CNT_CHAIN : 0u) | (ws.holes ? CNT_HOLES : 0u);
            J.tot[tile] = run - ws.slack;
            if (def_head != NODOC) J.def_first[tile] = def_head;
            ws.st_pre += n;
        }
        tile = DYN ? __shfl_sync(FULL, next_tile, 0) : next_tile;
    }
    // ---- the misses still queued ----
    __syncwarp();
#pragma unroll 1
    for (uint32_t cls = 0; cls < (uint32_t)MISS_CLASSES; ++cls) {
        const uint32_t qn = ws.qn[cls];
        if (qn) {
            run_batch(M, J, sm.byte1, ws.q[cls], qn, NODOC);
            if (lane == 0) ws.st_miss += qn;
        }
        __syncwarp();
    }
    // ---- statistics (a warp sees fewer than 2^32 pretokens: 2 TB of text over >= 8 warps) ----
    if (lane == 0) {
        Counters* c = J.ctr;
        const unsigned long long st_pre = ws.st_pre, st_miss = ws.st_miss, st_def = ws.st_def;
        if (st_pre) atomicAdd(&c->st_pretokens, st_pre);
        if (st_pre - st_miss - st_def) atomicAdd(&c->st_hits, st_pre - st_miss - st_def);
        if (st_def) atomicAdd(&c->st_block, st_def);
        if (st_miss) atomicAdd(&c->st_warp, st_miss);
    }
}

template <bool PFX>
__global__ void __launch_bounds__(NT, HUT_TILE_CTAS) tile_kernel(const __grid_constant__ DevModel M,
                                                                 const __grid_constant__ Job J) {
    __shared__ TileSm sm;
#ifdef HUT_STATIC_TILES
    tile_body<PFX, false>(M, J, sm, blockIdx.x * WPB + (threadIdx.x >> 5), gridDim.x * WPB);
#else
    tile_body<PFX, true>(M, J, sm, blockIdx.x * WPB + (threadIdx.x >> 5), gridDim.x * WPB);
#endif
}

}  // namespace hut
