/* sm_100a device code of the BPE encode engine, part 3: the small kernels around tile_kernel (see hut_kernels.cuh for
 * the overview).  Included only by hut_engine.cu. */
#pragma once

#include "hut_tile.cuh"

namespace hut {

__global__ void init_kernel(Job J) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        Counters* c = J.ctr;
        c->flags = 0;
        c->n_events = 0;
        c->n_def = 0;
        c->ticket = 0;
        c->ticket2 = 0;
        c->tile_ticket = 0;
        c->total_ids = 0;
        c->long_used = 0;
        c->st_pretokens = c->st_hits = c->st_warp = c->st_block = c->st_big = c->st_cut = 0;
#ifdef HUT_LONGSTAT
        c->lp[0] = c->lp[1] = c->lp[2] = c->lp[3] = 0;
#endif
    }
}

// first_doc[t] = first document starting inside tile t (pre-set to NODOC)
__device__ __forceinline__ void doc_first(const Job& J, unsigned long long d) {
    if (d >= J.ndocs) return;
    const unsigned long long pos = J.doc_off[d];
    if (pos >= J.nbytes) return;
    const unsigned long long t = pos / OWN;
    if (d == 0 || J.doc_off[d - 1] / OWN != t) J.first_doc[t] = (uint32_t)d;
}
__global__ void doc_kernel(Job J) { doc_first(J, (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x); }

// start of every tile's staging slot: tile * stride, plus slot_extra ids for each document starting before the tile
// (lower bound of the tile's first byte in doc_off) for a prefix model; see stage_slot()
__device__ __forceinline__ void slot_docs(const Job& J, unsigned long long* out, unsigned long long t) {
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
__global__ void slot_docs_kernel(Job J, unsigned long long* out) {
    slot_docs(J, out, (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x);
}

// ------------------------------------------------------------------------------------------------
// long_kernel / huge_kernel: one CTA per deferred pretoken (longer than 32 bytes / symbols, or running past the
// tile's halo).  long_kernel (128 threads) merges those of up to med_cap symbols in its CTA's own scratch area and
// leaves the longer ones (DefEntry flag 8) to huge_kernel (1024 threads, one working area of the big pool per CTA),
// which runs cooperative rounds when the model's ranks are monotone (cta_merge_run).
// ------------------------------------------------------------------------------------------------
template <bool HUGE>
__device__ __forceinline__ void long_entries(const DevModel& M, const Job& J, LongSm& sm) {
    const uint32_t tid = threadIdx.x;
    for (uint32_t i = tid; i < 256; i += blockDim.x) sm.byte1[i] = M.byte1[i];
    __syncthreads();
    uint32_t n = J.ctr->n_def;
    if (n > J.def_cap) n = J.def_cap;
    uint32_t* const scratch = HUGE ? J.big_pool + (unsigned long long)blockIdx.x * J.big_slot_words
                                   : J.cta_scratch + (unsigned long long)blockIdx.x * J.cta_scratch_words;
    const uint32_t cap_g = HUGE ? J.big_cap : J.med_cap;
    extern __shared__ __align__(16) uint32_t long_dyn[];  // long_kernel: 4 * LSM_CAP words
    unsigned long long n_big = 0;
    for (uint32_t e = blockIdx.x; e < n; e += gridDim.x) {
        DefEntry* E = J.def + e;
        __syncthreads();
        if (HUGE && !(E->flags & 8u)) continue;  // uniform: every thread reads the same word
        if (!HUGE && (E->flags & 4u)) continue;  // done by mid_kernel
        const unsigned long long gs = E->pos;
        const bool first = (E->flags & 1u) != 0;
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
        if (!HUGE) {
            const unsigned long long bound = (ge - gs) * M.kmax + M.n_prefix_syms;
            if (bound > J.med_cap) {  // huge_kernel's; it is told the length found here (< 2^32: limit * kmax fits)
                __syncthreads();
                if (tid == 0) {
                    E->len = (uint32_t)(ge - gs);
                    E->flags |= 8u;
                    if (J.big_slots == 0) set_flag(J, FLAG_BIG_NEED);  // the host allocates the pool and runs the shard again
                }
                continue;
            }
        }
        // a pretoken of up to LSM_CAP symbols (most of them) has its arrays and its whole tree in shared memory: every
        // merge is a chain of a dozen dependent accesses, so their latency is what it costs
        const bool in_sm = !HUGE && (ge - gs) * M.kmax + M.n_prefix_syms <= (unsigned long long)LSM_CAP;
        const uint32_t cap = in_sm ? (uint32_t)LSM_CAP : cap_g;
        uint32_t* U = in_sm ? long_dyn : scratch;
        uint32_t* MG = U + cap;
        uint32_t* NX = MG + cap;
        uint32_t* PV = NX + cap;
        // tournament-tree levels behind the four arrays (8-byte aligned; cap / 8 words are plenty), then -- huge_kernel
        // only -- the lists of the cooperative rounds
        unsigned long long* T = reinterpret_cast<unsigned long long*>(
            (reinterpret_cast<uintptr_t>(PV + cap) + 7u) & ~(uintptr_t)7u);
        uint32_t* X = HUGE ? reinterpret_cast<uint32_t*>(T) + (cap / 8u + 64u) : nullptr;
        uint32_t m = 0;
        uint32_t cnt = cta_merge_run(M, J, sm, gs, ge, with_prefix, alone_first, U, MG, NX, PV, T, cap, &m, X);
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
        }
        if (HUGE) n_big++;
    }
    if (tid == 0 && n_big) atomicAdd(&J.ctr->st_big, n_big);
}

// mid_kernel<CAP>: one WARP per deferred pretoken of at most CAP symbols (most of them: a long word, a number, a
// blob of a few hundred bytes) -- warp_merge_count, arrays in shared memory (16 bytes per symbol), no tree.  A merge
// is one dependent chain (reduction, links, one table probe, stores), so what counts is how many chains the GPU has
// in flight and how long a step of one is: CAP 1024 runs 12 one-warp CTAs per SM, CAP 4096 three (long_kernel: three
// four-warp CTAs of which one warp works, and a tree to refresh), each probing the pair table for 32 candidate merges
// at once (warp_merge_count).  Entries are handed out by a ticket counter.  Where tile_kernel did not see the end of the
// pretoken the warp finds it (and leaves the length in the entry for the kernels after it).  Entries finished here
// carry flag 4; long_kernel takes the rest: more symbols than 4096, pretokens to be cut, ranks of 2^20 and above.
template <uint32_t CAP>
__global__ void __launch_bounds__(32) mid_kernel(const DevModel M, const Job J) {
    __shared__ uint32_t byte1[256];
    extern __shared__ __align__(16) uint32_t mid_dyn[];  // 4 * CAP words
    uint32_t* const W = mid_dyn;
    const uint32_t lane = threadIdx.x;
    uint32_t n = J.ctr->n_def;
    if (n > J.def_cap) n = J.def_cap;
    if (blockIdx.x >= n) return;
    for (uint32_t i = lane; i < 256; i += 32) byte1[i] = M.byte1[i];
    __syncwarp();
    unsigned int* const ticket = CAP == 1024u ? &J.ctr->ticket : &J.ctr->ticket2;
    for (;;) {  // entries handed out one at a time: their lengths differ by two orders of magnitude
        uint32_t e = 0;
        if (lane == 0) e = atomicAdd(ticket, 1u);
        e = __shfl_sync(0xFFFFFFFFu, e, 0);
        if (e >= n) break;
        DefEntry* E = J.def + e;
        const uint32_t flags = E->flags;
        if (flags & (4u | 16u)) continue;  // done by the smaller variant / known to be beyond this path
        uint32_t len = E->len;
        const unsigned long long gs = E->pos;
        if (len == 0) {  // the end lies beyond what the tile had loaded: look for it, up to 4096 bytes on
            unsigned long long a = 0, b = J.ndocs + 1;  // end of the document containing gs: first doc_off > gs
            while (a < b) {
                const unsigned long long mid = (a + b) >> 1;
                if (J.doc_off[mid] <= gs) a = mid + 1;
                else b = mid;
            }
            unsigned long long dend = a <= J.ndocs ? J.doc_off[a] : J.nbytes;
            if (dend > J.nbytes) dend = J.nbytes;
            uint32_t cls = gclass(J, gs, gs, dend);
            if (cls == HC_SPACE) {
                const uint32_t c1 = gclass(J, gs + 1, gs, dend);
                if (c1 >= HC_ALPHA && c1 <= HC_OTHER) cls = c1;
            }
            const unsigned long long ge = warp_find_run_end(J, gs + 1, gs, dend, cls, gs + 4096u);
            if (ge - gs > 4096u) {
                if (lane == 0) E->flags = flags | 16u;
                continue;
            }
            len = (uint32_t)(ge - gs);
            if (lane == 0) E->len = len;
        }
        if (len > CAP || len > M.limit) continue;
        const unsigned long long ge = gs + len;
        const bool first = (flags & 1u) != 0;
        const uint32_t b0 = __ldg(J.text + gs);
        const bool with_prefix = first && b0 != ' ', alone_first = first && b0 == ' ';
        uint32_t m = 0;
        const uint32_t cnt = warp_merge_count<CAP>(M, J, byte1, gs, ge, with_prefix, alone_first, W, &m);
        if (cnt == 0xFFFFFFFFu) continue;  // more than CAP symbols
        unsigned long long off = 0;
        if (lane == 0) off = atomicAdd(&J.ctr->long_used, (unsigned long long)cnt);
        off = __shfl_sync(0xFFFFFFFFu, off, 0);
        if (off + cnt > J.long_cap) {
            set_flag(J, FLAG_LONG_CAP);  // the host grows long_ids and runs the shard again
        } else if (cnt) {
            warp_merge_emit(M, W, m, alone_first, J.long_ids + off);
        }
        __syncwarp();
        if (lane == 0) {
            E->count = cnt;
            E->ids_off = off;
            E->flags = flags | 4u;
            atomicAdd(&J.tot[E->tile], cnt);
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(LNT) long_kernel(const DevModel M, const Job J) {
    __shared__ LongSm sm;
    long_entries<false>(M, J, sm);
}

__global__ void __launch_bounds__(HNT, 1) huge_kernel(const DevModel M, const Job J) {
    __shared__ LongSm sm;
    if (J.ctr->flags & FLAG_BIG_NEED) return;  // no pool yet
    long_entries<true>(M, J, sm);
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
    if (tid == 0) {
        J.ctr->total_ids = carry_s;
        J.ctr->tile_ticket = 0;  // tile_kernel is done with it: compact_kernel hands its tiles out the same way
    }
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
// staging slot of one tile -> dense ids (warp); drops the HOLEs left by deferred cache misses and splices in the ids of
// the pretokens the long-path kernels merged
__device__ __forceinline__ void compact_one_tile(const Job& J, const uint32_t tile, const uint32_t lane) {
    {
        const uint32_t c = J.cnt[tile];
        const uint32_t ns = c & CNT_MASK;
        unsigned long long o = tile_base(J, tile);
        const int32_t* src = stage_slot(J, tile);
        if (!(c & (CNT_CHAIN | CNT_HOLES))) {
            copy_ids(J, src, o, ns);
            return;
        }
        if (!(c & CNT_CHAIN)) {
            compact_tile_holes(J, tile, src, o, ns);
            return;
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


// DYN: groups of COMPACT_GROUP consecutive tiles are handed out by the ticket counter (one atomic per group: a ticket per
// tile, 1.1 M atomics on one address inside half a millisecond, serialises in L2 -- measured)
#ifndef HUT_COMPACT_GROUP
#define HUT_COMPACT_GROUP 8
#endif
constexpr uint32_t COMPACT_GROUP = HUT_COMPACT_GROUP;
template <bool DYN>
__device__ __forceinline__ void compact_tiles(const Job& J, const uint32_t gw, const uint32_t nw) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t tile = gw, next_base = 0;
    if (DYN) {
        if (lane == 0) tile = atomicAdd(&J.ctr->tile_ticket, COMPACT_GROUP);
        tile = __shfl_sync(0xFFFFFFFFu, tile, 0);
        if (lane == 0) next_base = atomicAdd(&J.ctr->tile_ticket, COMPACT_GROUP);
    }
    for (; tile < J.ntiles;) {
        const uint32_t this_tile = tile;
        if (DYN) {  // next tile: the following one of the group, or the first of the group asked for a group ago
            if (((tile + 1u) % COMPACT_GROUP) != 0u) {
                tile = tile + 1u;
            } else {
                tile = __shfl_sync(0xFFFFFFFFu, next_base, 0);
                if (lane == 0) next_base = atomicAdd(&J.ctr->tile_ticket, COMPACT_GROUP);
            }
        } else {
            tile += nw;
        }
        compact_one_tile(J, this_tile, lane);
    }
}

__global__ void __launch_bounds__(256, 8) compact_kernel(const Job J) {
#ifdef HUT_STATIC_TILES
    compact_tiles<false>(J, blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), gridDim.x * (blockDim.x >> 5));
#else
    compact_tiles<true>(J, blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), gridDim.x * (blockDim.x >> 5));
#endif
}

// tile-local document offsets -> global ones
__device__ __forceinline__ void fixdoc_doc(const Job& J, unsigned long long d) {
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
__global__ void fixdoc_kernel(const Job J) { fixdoc_doc(J, (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x); }

// ------------------------------------------------------------------------------------------------
// small_kernel: the whole pipeline for a SMALL batch (a sentence, a handful of short documents) in ONE launch of one
// CTA -- what hutoken.encode(text) costs is launch latency, not throughput (the reference's own check of this path is a
// per-call bound, tests/test_tokenizer.py:107-130).  Text, offsets, ids and counters live in pinned host memory the
// device reads and writes directly, so there is no copy either; the host waits on `done`.  Anything the one CTA does
// not do itself (pretokens handed to long_kernel) makes it give up: the host then takes the general path.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t SMALL_MAX_TILES = 72;  // 69 120 bytes
constexpr uint32_t SMALL_FALLBACK = 0x80000000u;

template <bool PFX>
__global__ void __launch_bounds__(NT, 1) small_kernel(const __grid_constant__ DevModel M, const __grid_constant__ Job J,
                                                      Counters* h_ctr, volatile unsigned int* done, unsigned int seq) {
    __shared__ TileSm sm;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) {
        Counters* c = J.ctr;
        c->flags = 0;
        c->n_events = 0;
        c->n_def = 0;
        c->total_ids = 0;
        c->long_used = 0;
        c->st_pretokens = c->st_hits = c->st_warp = c->st_block = c->st_big = c->st_cut = 0;
    }
    for (uint32_t t = tid; t <= J.ntiles; t += NT) J.first_doc[t] = NODOC;
    __syncthreads();
    for (unsigned long long d = tid; d < J.ndocs; d += NT) doc_first(J, d);
    if (PFX)
        for (uint32_t t = tid; t <= J.ntiles; t += NT) slot_docs(J, const_cast<unsigned long long*>(J.slot_off), t);
    __syncthreads();
    tile_body<PFX, false>(M, J, sm, tid >> 5, WPB);
    __syncthreads();
    const bool give_up = J.ctr->n_def != 0 || (J.ctr->flags & ~(unsigned)FLAG_NUL) != 0;
    if (!give_up) {
        if (tid == 0) {  // exclusive scan of the (few) per-tile id counts, in place like scan1 / scan2
            unsigned long long run = 0;
            for (uint32_t t = 0; t < J.ntiles; ++t) {
                const uint32_t v = J.tot[t];
                J.tot[t] = (uint32_t)run;
                run += v;
            }
            J.bsum[0] = 0;
            J.ctr->total_ids = run;
        }
        __syncthreads();
        compact_tiles<false>(J, tid >> 5, WPB);
        for (unsigned long long d = tid; d <= J.ndocs; d += NT) fixdoc_doc(J, d);
    }
    __syncthreads();
    if (tid == 0) {
        const Counters* c = J.ctr;
        h_ctr->flags = c->flags | (give_up ? SMALL_FALLBACK : 0u);
        h_ctr->total_ids = c->total_ids;
        h_ctr->st_pretokens = c->st_pretokens;
        h_ctr->st_hits = c->st_hits;
        h_ctr->st_warp = c->st_warp;
        __threadfence_system();
        *done = seq;
    }
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
