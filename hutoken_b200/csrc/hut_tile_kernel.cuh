/* The fused persistent tile kernel + its small companions.  Included only by hut_engine.cu. */
#pragma once

#include "hut_kernels.cuh"

namespace hut {

// bits of 32-byte segment `seg` that lie in the tile proper, i.e. span index in [BACK, BACK + TILE)
__device__ __forceinline__ uint32_t proper_mask(int seg) {
    int lo = seg * 32, hi = lo + 32;
    int a = lo > BACK ? lo : BACK, b = hi < BACK + TILE ? hi : BACK + TILE;
    if (a >= b) return 0u;
    uint32_t m = 0xFFFFFFFFu << (a - lo);
    if (b < hi) m &= (1u << (b - lo)) - 1u;
    return m;
}

__device__ __forceinline__ int next_boundary(const Smem& sm, int li) {
    int w = li >> 5, b = li & 31;
    uint32_t m = (b == 31) ? 0u : (sm.bndmask[w] & (0xFFFFFFFEu << b));
    while (m == 0) {
        if (++w >= NSEG) return SPAN;
        m = sm.bndmask[w];
    }
    return w * 32 + __ffs(m) - 1;
}

__device__ __forceinline__ bool doc_bit(const Smem& sm, int li) { return (sm.docmask[li >> 5] >> (li & 31)) & 1u; }

// cache key of the pretoken text[s .. s+len), len <= 23: k0 = meta | bytes 0..6, k1 = bytes 7..14, k2 = bytes 15..22
__device__ __forceinline__ void build_key(const Smem& sm, int s, int len, uint32_t flag, unsigned long long& k0,
                                          unsigned long long& k1, unsigned long long& k2) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(sm.text + (s & ~3));
    const uint32_t sh = (uint32_t)(s & 3) * 8u;
    uint32_t r[7], x[6];
#pragma unroll
    for (int i = 0; i < 7; ++i) r[i] = w[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        x[i] = __funnelshift_r(r[i], r[i + 1], sh);
        int rem = len - 4 * i;
        x[i] = rem >= 4 ? x[i] : (rem <= 0 ? 0u : (x[i] & ((1u << (8 * rem)) - 1u)));
    }
    unsigned long long a = x[0] | ((unsigned long long)x[1] << 32);
    unsigned long long b = x[2] | ((unsigned long long)x[3] << 32);
    unsigned long long c = x[4] | ((unsigned long long)x[5] << 32);
    unsigned long long meta = 0x80u | (flag << 5) | (uint32_t)len;
    k0 = (a << 8) | meta;
    k1 = (a >> 56) | (b << 8);
    k2 = (b >> 56) | (c << 8);
}

// ids before pretoken q of the current pass (q in [0, n]); c = pretokens per thread in the blocked scan
__device__ __forceinline__ uint32_t tokens_before(const Smem& sm, uint32_t q, uint32_t n, uint32_t c, uint32_t total) {
    if (q >= n) return total;
    uint32_t t = q / c;
    uint32_t v = sm.tbase[t];
    for (uint32_t x = t * c; x < q; ++x) v += val_count(sm.pval[x]);
    return v;
}

__global__ void prep_kernel(Job J) {
    unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < J.ntiles) {
        J.tile_state[t] = 0;
        unsigned long long g = t * TILE;
        g = g >= BACK ? g - BACK : 0;
        unsigned long long lo = 0, hi = J.ndocs + 1;  // first d in [0, ndocs] with doc_off[d] >= g
        while (lo < hi) {
            unsigned long long mid = (lo + hi) >> 1;
            if (J.doc_off[mid] < g) lo = mid + 1;
            else hi = mid;
        }
        J.first_doc[t] = (uint32_t)lo;
    }
    if (t == 0) {
        Counters* c = J.ctr;
        c->ticket = 0;
        c->flags = 0;
        c->n_events = 0;
        c->total_ids = 0;
        c->st_pretokens = c->st_hits = c->st_warp = c->st_block = c->st_big = c->st_cut = 0;
        for (int i = 0; i < 32; ++i) c->pool_lock[i] = 0;
    }
}

__device__ __forceinline__ void set_flag(const Job& J, uint32_t f) { atomicOr(&J.ctr->flags, f); }

__device__ __forceinline__ void put_id(const Job& J, unsigned long long idx, int32_t v) {
    if (idx < J.out_cap) st_stream_s32(J.out + idx, v);
    else set_flag(J, FLAG_OUT_CAP);
}

__global__ void __launch_bounds__(NT) tile_kernel(const DevModel M, const Job J) {
    __shared__ Smem sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int i = tid; i < 256; i += NT) sm.byte1[i] = M.byte1[i];
    if (tid == 0) {
        mbar_init(&sm.mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 6) sm.st[tid] = 0;
    __syncthreads();

    uint32_t parity = 0;
    uint32_t* const scratch = J.cta_scratch + (unsigned long long)blockIdx.x * J.cta_scratch_words;
    uint32_t* const medU = scratch;
    uint32_t* const medMG = medU + J.med_cap;
    uint32_t* const medNX = medMG + J.med_cap;
    uint32_t* const medPV = medNX + J.med_cap;
    int32_t* const ovf = reinterpret_cast<int32_t*>(medPV + J.med_cap);
    uint16_t* const missq = reinterpret_cast<uint16_t*>(sm.stage);

    for (;;) {
        __syncthreads();
        if (tid == 0) sm.tile = atomicAdd(&J.ctr->ticket, 1u);
        __syncthreads();
        const uint32_t tile = sm.tile;
        if (tile >= J.ntiles) break;
        const long long tile_start = (long long)tile * TILE;
        const long long g0 = tile_start - BACK;  // global position of span index 0
        const long long lo = g0 < 0 ? 0 : g0;
        const long long hi = (g0 + SPAN) < (long long)J.nbytes ? (g0 + SPAN) : (long long)J.nbytes;
        const int lo_li = (int)(lo - g0), hi_li = (int)(hi - g0);
        const uint32_t body = (uint32_t)(hi_li - lo_li) & ~15u;
        const bool last_tile = (tile + 1 == J.ntiles);

        // ---- stage the span: one TMA bulk copy for the aligned body, plain loads for the ragged rest ----
        if (tid == 0 && body) {
            fence_proxy_async();
            mbar_expect_tx(&sm.mbar, body);
            tma_load_1d(sm.text + lo_li, J.text + lo, body, &sm.mbar);
        }
        for (int li = tid; li < lo_li; li += NT) sm.text[li] = 0;
        for (int li = lo_li + (int)body + tid; li < SPAN; li += NT) sm.text[li] = li < hi_li ? J.text[g0 + li] : 0;
        for (int i = tid; i < NSEG + 2; i += NT) {
            sm.docmask[i] = 0;
            sm.startmask[i] = 0;
            sm.bndmask[i] = 0;
            sm.wcount[i] = 0;
        }
        if (tid == 0) sm.dfirst = 0xFFFFFFFFu;
        __syncthreads();
        {  // document starts inside the span (doc_off[ndocs] = nbytes acts as the final boundary)
            const unsigned long long lim = (unsigned long long)(g0 + SPAN);
            for (unsigned long long d = (unsigned long long)J.first_doc[tile] + tid; d <= J.ndocs; d += NT) {
                unsigned long long pos = J.doc_off[d];
                if (pos >= lim) break;
                int li = (int)((long long)pos - g0);
                atomicOr(&sm.docmask[li >> 5], 1u << (li & 31));
                if ((long long)pos >= tile_start) atomicMin(&sm.dfirst, (uint32_t)d);
            }
        }
        if (body) mbar_wait(&sm.mbar, parity);
        if (body) parity ^= 1u;
        __syncthreads();

        // ---- pass 1: class code of every byte as a potential char start (len-1 in bits 4-5) ----
        uint32_t multi = 0;
        for (int seg = tid, it = 0; seg < NSEG; seg += NT, ++it) {
            const uint32_t* tw = reinterpret_cast<const uint32_t*>(sm.text + seg * 32);
            uint32_t w[9];
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = tw[i];
            w[8] = (seg + 1 < NSEG) ? tw[8] : 0u;
            const unsigned long long dm =
                (unsigned long long)sm.docmask[seg] | ((unsigned long long)sm.docmask[seg + 1] << 32);
            uint32_t hib = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) hib |= w[i];
            uint32_t outw[8];
            if ((hib & 0x80808080u) == 0) {  // all ASCII
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    uint32_t o = 0;
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        uint32_t c0 = (w[i] >> (8 * b)) & 0xFFu;
                        o |= (hut_lead_class(c0, 0, 0, 0) & 15u) << (8 * b);
                    }
                    outw[i] = o;
                }
            } else {
                multi |= 1u << it;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    uint32_t o = 0;
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        const int j = i * 4 + b;
                        uint32_t c0 = (w[i] >> (8 * b)) & 0xFFu;
                        uint32_t code;
                        if (c0 < 0x80u) {
                            code = hut_lead_class(c0, 0, 0, 0) & 15u;
                        } else {
                            // following bytes, cut at the next document start
                            unsigned long long nx =
                                ((unsigned long long)w[(j + 1) >> 2] >> (8 * ((j + 1) & 3))) & 0xFFull;
                            uint32_t b1 = (uint32_t)nx;
                            uint32_t b2 = (w[(j + 2) >> 2] >> (8 * ((j + 2) & 3))) & 0xFFu;
                            uint32_t b3 = (w[(j + 3) >> 2] >> (8 * ((j + 3) & 3))) & 0xFFu;
                            uint32_t nb = (uint32_t)(dm >> (j + 1)) & 7u;
                            if (nb & 1u) b1 = b2 = b3 = 0;
                            else if (nb & 2u) b2 = b3 = 0;
                            else if (nb & 4u) b3 = 0;
                            uint32_t r = hut_lead_class(c0, b1, b2, b3);
                            code = (r & 15u) | (((r >> 4) - 1u) << 4);
                        }
                        o |= code << (8 * b);
                    }
                    outw[i] = o;
                }
            }
            uint32_t* cw = reinterpret_cast<uint32_t*>(sm.code + seg * 32);
#pragma unroll
            for (int i = 0; i < 8; ++i) cw[i] = outw[i];
        }
        __syncthreads();
        // ---- pass 2: bytes covered by a valid multi-byte lead become HC_CONT ----
        for (int seg = tid, it = 0; seg < NSEG; seg += NT, ++it) {
            if (!((multi >> it) & 1u)) continue;
            for (int j = 0; j < 32; ++j) {
                int li = seg * 32 + j;
                uint32_t c = sm.code[li];
                uint32_t extra = (c >> 4) & 3u;
                for (uint32_t k = 1; k <= extra; ++k)
                    if (li + (int)k < SPAN) sm.code[li + k] = HC_CONT;
            }
        }
        __syncthreads();
        // ---- pass 3: pretoken starts (local rule) ----
        for (int seg = tid; seg < NSEG; seg += NT) {
            const uint32_t* cw = reinterpret_cast<const uint32_t*>(sm.code + seg * 32);
            uint32_t w[9];  // w[0] = the 4 codes before the segment
            w[0] = seg ? cw[-1] : 0u;
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i + 1] = cw[i];
            const unsigned long long dmw = ((unsigned long long)sm.docmask[seg] << 32) |
                                           (seg ? (unsigned long long)sm.docmask[seg - 1] : 0ull);
            uint32_t smask = 0, emask = 0;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                auto cod = [&](int rel) -> uint32_t {  // class code at segment-relative index rel in [-4, 31]
                    int x = rel + 4;
                    return (w[x >> 2] >> (8 * (x & 3))) & 7u;
                };
                uint32_t c = cod(j);
                if (c == HC_END) {
                    emask |= 1u << j;
                    continue;
                }
                if (c == HC_CONT) continue;
                int ds = (int)((dmw >> (32 + j)) & 1ull);
                uint32_t p1 = cod(j - 1);
                uint32_t pc = p1;
                if (pc == HC_CONT) {
                    pc = cod(j - 2);
                    if (pc == HC_CONT) {
                        pc = cod(j - 3);
                        if (pc == HC_CONT) pc = cod(j - 4);
                    }
                }
                int lone = (int)((dmw >> (32 + j - 1)) & 1ull) || cod(j - 2) != HC_SPACE;
                if (hut_is_start(c, ds, p1, pc, lone)) smask |= 1u << j;
            }
            sm.startmask[seg] = smask;
            sm.bndmask[seg] = smask | emask;
            const uint32_t pm = proper_mask(seg);
            sm.wcount[seg] = (uint16_t)__popc(smask & pm);
            if (emask & pm) {  // a NUL inside the text (not past its end)?
                uint32_t e = emask & pm;
                while (e) {
                    int b = __ffs(e) - 1;
                    e &= e - 1;
                    if (g0 + seg * 32 + b < (long long)J.nbytes) set_flag(J, FLAG_NUL);
                }
            }
        }
        __syncthreads();
        if (warp == 0) {  // exclusive scan of the per-word start counts
            constexpr int PER = (NSEG + 31) / 32;
            uint32_t loc[PER], sum = 0;
#pragma unroll
            for (int i = 0; i < PER; ++i) {
                int s = lane * PER + i;
                loc[i] = s < NSEG ? sm.wcount[s] : 0u;
                sum += loc[i];
            }
            uint32_t inc = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t n = __shfl_up_sync(0xFFFFFFFFu, inc, d);
                if (lane >= d) inc += n;
            }
            uint32_t run = inc - sum;
#pragma unroll
            for (int i = 0; i < PER; ++i) {
                int s = lane * PER + i;
                if (s < NSEG) sm.wprefix[s] = (uint16_t)run;
                run += loc[i];
            }
            if (lane == 31) sm.np = inc;
        }
        __syncthreads();
        for (int seg = tid; seg < NSEG; seg += NT) {
            uint32_t m = sm.startmask[seg] & proper_mask(seg);
            uint32_t k = sm.wprefix[seg];
            while (m) {
                int b = __ffs(m) - 1;
                m &= m - 1;
                sm.pstart[k++] = (uint16_t)(seg * 32 + b);
            }
        }
        __syncthreads();
        const uint32_t np = sm.np;
        if (tid == 0) sm.st[0] += np;

        // the last pretoken of the tile may outgrow the span: it takes the block-cooperative "big" path
        bool big_last = false;
        if (np) big_last = next_boundary(sm, sm.pstart[np - 1]) > LIM;
        const bool slow = big_last || np > PCAP;
        if (slow) {
            if (warp == 0) {
                unsigned long long b = lookback(J, tile, 0, false);
                if (lane == 0) sm.base = b;
            }
            __syncthreads();
        }
        unsigned long long running = 0;
        const uint32_t npass = np ? (np + PCAP - 1) / PCAP : 1;
        for (uint32_t pass = 0; pass < npass; ++pass) {
            const uint32_t seg0 = pass * PCAP;
            const uint32_t nseg = np ? ((np - seg0) < (uint32_t)PCAP ? (np - seg0) : (uint32_t)PCAP) : 0u;
            const bool has_big = big_last && (pass + 1 == npass);
            const uint32_t nreal = nseg - (has_big ? 1u : 0u);
            __syncthreads();
            if (tid == 0) {
                sm.nmiss = 0;
                sm.nmed = 0;
                sm.ovf_used = 0;
                sm.nbigout = 0;
            }
            __syncthreads();
            // ---- unique-pretoken cache probe, one thread per pretoken ----
            uint32_t hits = 0;
            for (uint32_t q = tid; q < nreal; q += NT) {
                const int s = sm.pstart[seg0 + q];
                const int e = next_boundary(sm, s);
                const int len = e - s;
                unsigned long long v = 0;
                if (M.cache_ok && len <= CACHE_KEY_MAX) {
                    unsigned long long k0, k1, k2;
                    build_key(sm, s, len, (M.has_prefix && doc_bit(sm, s)) ? 1u : 0u, k0, k1, k2);
                    v = cache_lookup(J, k0, k1, k2);
                }
                sm.pval[q] = v;
                if (v == 0) missq[atomicAdd(&sm.nmiss, 1u)] = (uint16_t)q;
                else hits++;
            }
            if (hits) atomicAdd(&sm.st[1], (unsigned long long)hits);
            __syncthreads();
            // ---- misses: warp-per-pretoken merge ----
            const uint32_t nmiss = sm.nmiss;
            const bool all_block = M.multi_sym != 0;  // a byte may expand to several symbols: no register path
            for (uint32_t x = warp; x < nmiss && !all_block; x += NW) {
                const uint32_t q = missq[x];
                const int s = sm.pstart[seg0 + q];
                const int e = next_boundary(sm, s);
                const int len = e - s;
                const bool first = M.has_prefix && doc_bit(sm, s);
                const bool with_prefix = first && sm.text[s] != ' ';
                const bool alone_first = first && sm.text[s] == ' ';
                const uint32_t npfx = with_prefix ? M.n_prefix_syms : 0u;
                bool to_block = len > 128 || npfx > 8;
                uint32_t m = npfx;
                if (!to_block) {
                    if ((uint32_t)lane < npfx) sm.wscratch[warp][lane] = M.prefix_syms[lane];
                    auto at = [&](unsigned long long p) -> uint32_t { return sm.text[p]; };
                    for (int c0 = 0; c0 < len; c0 += 32) {
                        const int j = c0 + lane;
                        uint32_t cnt = 0, unit = M.inert, pb = 0;
                        if (j < len)
                            cnt = char_symbols(M, sm.byte1, at, (unsigned long long)(s + j), (unsigned long long)s,
                                               (unsigned long long)e, &unit, &pb);
                        const uint32_t mask = __ballot_sync(0xFFFFFFFFu, cnt != 0);
                        const uint32_t idx = m + __popc(mask & ((1u << lane) - 1u));
                        const uint32_t add = __popc(mask);
                        if (m + add > 32) {
                            to_block = true;
                            break;
                        }
                        if (cnt) sm.wscratch[warp][idx] = unit;
                        m += add;
                    }
                }
                if (to_block) {
                    if (lane == 0) {
                        uint32_t slot = atomicAdd(&sm.nmed, 1u);
                        if (slot < 160) sm.medq[slot] = q;
                        else set_flag(J, FLAG_BIG_CAP);  // cannot happen: > 32 symbols each
                    }
                    continue;
                }
                __syncwarp();
                uint32_t u = (uint32_t)lane < m ? sm.wscratch[warp][lane] : NONE;
                __syncwarp();
                const uint32_t cnt = warp_merge(M, u, m);
                const int32_t id = (uint32_t)lane < cnt ? id_of_unit(M, u) : -1;
                const uint32_t nal = alone_first ? M.n_prefix_alone : 0u;
                const uint32_t total = cnt + nal;
                if (M.cache_ok && total <= 3) {
                    unsigned long long v = 0;
#pragma unroll
                    for (uint32_t i = 0; i < 3; ++i) {
                        int32_t src = __shfl_sync(0xFFFFFFFFu, id, i >= nal ? (int)(i - nal) : 0);
                        int32_t t = i < nal ? M.prefix_alone[i] : src;
                        if (i < total) v |= (unsigned long long)(uint32_t)(t + 2) << (21 * i);
                    }
                    if (lane == 0) {
                        sm.pval[q] = v;
                        if (len <= CACHE_KEY_MAX) {
                            unsigned long long k0, k1, k2;
                            build_key(sm, s, len, first ? 1u : 0u, k0, k1, k2);
                            cache_insert(J, k0, k1, k2, v);
                        }
                    }
                } else {
                    uint32_t off = 0;
                    if (lane == 0) off = atomicAdd(&sm.ovf_used, total);
                    off = __shfl_sync(0xFFFFFFFFu, off, 0);
                    if (off + total <= J.ovf_cap) {
                        if ((uint32_t)lane < nal) ovf[off + lane] = M.prefix_alone[lane];
                        if ((uint32_t)lane < cnt) ovf[off + nal + lane] = id;
                    } else {
                        set_flag(J, FLAG_BIG_CAP);
                    }
                    if (lane == 0) sm.pval[q] = esc_make(total, off);
                }
                if (lane == 0) atomicAdd(&sm.st[2], 1ull);
            }
            __syncthreads();
            // ---- block-cooperative merge: pretokens > 32 symbols (and everything when a byte maps to several symbols) ----
            {
                const uint32_t nmed = all_block ? nmiss : (sm.nmed < 160u ? sm.nmed : 160u);
                for (uint32_t x = 0; x < nmed; ++x) {
                    const uint32_t q = all_block ? (uint32_t)missq[x] : sm.medq[x];
                    const int s = sm.pstart[seg0 + q];
                    const int e = next_boundary(sm, s);
                    const bool first = M.has_prefix && doc_bit(sm, s);
                    const bool with_prefix = first && sm.text[s] != ' ';
                    const bool alone_first = first && sm.text[s] == ' ';
                    __syncthreads();
                    const uint32_t off = sm.ovf_used;
                    uint32_t cnt = cta_merge(M, J, sm, (unsigned long long)(g0 + s), (unsigned long long)(g0 + e),
                                             with_prefix, alone_first, medU, medMG, medNX, medPV, J.med_cap, ovf + off,
                                             off < J.ovf_cap ? J.ovf_cap - off : 0u);
                    if (cnt == 0xFFFFFFFFu || off + cnt > J.ovf_cap) {
                        set_flag(J, FLAG_BIG_CAP);
                        cnt = 0;
                    }
                    if (tid == 0) {
                        sm.pval[q] = esc_make(cnt, off);
                        sm.ovf_used = off + cnt;
                        sm.st[3] += 1;
                    }
                    __syncthreads();
                }
            }
            // ---- counts -> blocked scan ----
            const uint32_t c = nreal ? (nreal + NT - 1) / NT : 1u;
            uint32_t k0 = tid * c, k1 = k0 + c;
            if (k0 > nreal) k0 = nreal;
            if (k1 > nreal) k1 = nreal;
            uint32_t mine = 0;
            for (uint32_t q = k0; q < k1; ++q) mine += val_count(sm.pval[q]);
            uint32_t total;
            const uint32_t excl = block_scan_excl(sm, mine, &total);
            sm.tbase[tid] = excl;
            __syncthreads();
            if (!slow) {  // normal tile: publish the aggregate, then look back for the exclusive prefix
                if (warp == 0) {
                    unsigned long long b = lookback(J, tile, total, true);
                    if (lane == 0) sm.base = b;
                }
                __syncthreads();
            }
            const unsigned long long gbase = sm.base + running;
            // ---- emit: scatter ids into the staging window, flush coalesced ----
            for (uint32_t cb = 0; cb < total; cb += STAGE) {
                uint32_t off = excl;
                for (uint32_t q = k0; q < k1; ++q) {
                    const unsigned long long v = sm.pval[q];
                    const uint32_t n = val_count(v);
                    if (off + n > cb && off < cb + STAGE) {
                        if (!(v & ESC)) {
#pragma unroll
                            for (uint32_t i = 0; i < 3; ++i) {
                                uint32_t p = off + i - cb;
                                if (i < n && p < (uint32_t)STAGE) sm.stage[p] = val_id(v, i);
                            }
                        } else if (n <= 64 || sm.nbigout >= 96) {
                            const int32_t* src = ovf + (uint32_t)v;
                            for (uint32_t i = 0; i < n; ++i) {
                                uint32_t p = off + i - cb;
                                if (p < (uint32_t)STAGE) sm.stage[p] = src[i];
                            }
                        } else if (cb <= off) {  // long run: copied cooperatively below (listed once)
                            uint32_t slot = atomicAdd(&sm.nbigout, 1u);
                            if (slot < 96) sm.bigout[slot] = q;
                        }
                    }
                    off += n;
                }
                __syncthreads();
                const uint32_t nflush = (total - cb) < (uint32_t)STAGE ? (total - cb) : (uint32_t)STAGE;
                for (uint32_t i = tid; i < nflush; i += NT) put_id(J, gbase + cb + i, sm.stage[i]);
                __syncthreads();
            }
            {
                const uint32_t nb = sm.nbigout < 96u ? sm.nbigout : 96u;
                for (uint32_t x = 0; x < nb; ++x) {
                    const uint32_t q = sm.bigout[x];
                    const unsigned long long v = sm.pval[q];
                    const uint32_t n = val_count(v);
                    const uint32_t off = tokens_before(sm, q, nreal, c, total);
                    const int32_t* src = ovf + (uint32_t)v;
                    for (uint32_t i = tid; i < n; i += NT) put_id(J, gbase + off + i, src[i]);
                }
            }
            // ---- per-document id offsets for documents starting at a pretoken of this pass ----
            if (sm.dfirst != 0xFFFFFFFFu) {
                const long long tile_end = tile_start + TILE;
                for (unsigned long long d = (unsigned long long)sm.dfirst + tid; d < J.ndocs; d += NT) {
                    const long long pos = (long long)J.doc_off[d];
                    if (pos >= tile_end || pos >= (long long)J.nbytes) break;
                    const int li = (int)(pos - g0);
                    const int w = li >> 5, b = li & 31;
                    const uint32_t kk = sm.wprefix[w] + __popc(sm.startmask[w] & proper_mask(w) & ((1u << b) - 1u));
                    if (kk >= seg0 && kk < seg0 + nseg)
                        J.tok_off[d] = gbase + tokens_before(sm, kk - seg0, nreal, c, total);
                }
            }
            running += total;
        }
        // ---- the pretoken that outgrows the span ----
        if (big_last) {
            const int s = sm.pstart[np - 1];
            const unsigned long long gs = (unsigned long long)(g0 + s);
            uint32_t cls = sm.code[s] & 7u;
            if (cls == HC_SPACE) {
                uint32_t c1 = sm.code[s + 1] & 7u;
                if (c1 >= HC_ALPHA && c1 <= HC_OTHER) cls = c1;
            }
            const bool first = M.has_prefix && doc_bit(sm, s);
            const bool with_prefix = first && sm.text[s] != ' ';
            const bool alone_first = first && sm.text[s] == ' ';
            if (tid == 0) {  // end of the document containing gs: first doc_off > gs
                unsigned long long a = 0, b = J.ndocs + 1;
                while (a < b) {
                    unsigned long long mid = (a + b) >> 1;
                    if (J.doc_off[mid] <= gs) a = mid + 1;
                    else b = mid;
                }
                sm.doc_end = a <= J.ndocs ? J.doc_off[a] : J.nbytes;
            }
            __syncthreads();
            const unsigned long long dend = sm.doc_end;
            const unsigned long long ge =
                find_run_end(J, sm, (unsigned long long)(g0 + LIM - 3), gs, dend, cls, gs + M.limit + 1);
            uint32_t cnt = 0;
            if (ge - gs > M.limit) {
                if (M.compat_cut) {  // core.c:402-407,503: the document ends here
                    if (tid == 0) {
                        uint32_t ev = atomicAdd(&J.ctr->n_events, 1u);
                        if (ev < (uint32_t)MAX_EVENTS) {
                            J.ctr->events[ev].tok_index = sm.base + running;
                            J.ctr->events[ev].pos = gs;
                        } else {
                            set_flag(J, FLAG_EVENTS);
                        }
                    }
                } else {
                    set_flag(J, FLAG_BIG_CAP);
                }
            } else {
                if (tid == 0) {
                    uint32_t slot = 0;
                    for (;;) {
                        if (atomicCAS(&J.ctr->pool_lock[slot], 0u, 1u) == 0u) break;
                        slot = (slot + 1) % J.big_slots;
                        __nanosleep(200);
                    }
                    sm.slot = slot;
                }
                __syncthreads();
                uint32_t* U = J.big_pool + (unsigned long long)sm.slot * J.big_slot_words;
                const unsigned long long at0 = sm.base + running;
                cnt = cta_merge(M, J, sm, gs, ge, with_prefix, alone_first, U, U + J.big_cap,
                                U + 2ull * J.big_cap, U + 3ull * J.big_cap, J.big_cap, J.out + at0,
                                at0 < J.out_cap ? J.out_cap - at0 : 0ull);
                if (cnt == 0xFFFFFFFFu) {
                    set_flag(J, FLAG_BIG_CAP);
                    cnt = 0;
                }
                if (at0 + cnt > J.out_cap) set_flag(J, FLAG_OUT_CAP);
                __syncthreads();
                if (tid == 0) {
                    __threadfence();
                    atomicExch(&J.ctr->pool_lock[sm.slot], 0u);
                    sm.st[4] += 1;
                }
            }
            running += cnt;
        }
        if (slow && tid == 0) {
            __threadfence();
            st_volatile_u64(J.tile_state + tile, ST_INC | (sm.base + running));
        }
        if (last_tile) {
            // documents that start at the very end of the text (empty trailing documents) and the grand total
            const unsigned long long tot = sm.base + running;
            for (unsigned long long d = J.ndocs - tid; d <= J.ndocs; d -= NT) {
                if (d < J.ndocs && J.doc_off[d] < J.nbytes) break;
                J.tok_off[d] = tot;
                if (d < (unsigned long long)NT) break;
            }
            if (tid == 0) J.ctr->total_ids = tot;
        }
    }
    if (tid == 0) {
        Counters* c = J.ctr;
        if (sm.st[0]) atomicAdd(&c->st_pretokens, sm.st[0]);
        if (sm.st[1]) atomicAdd(&c->st_hits, sm.st[1]);
        if (sm.st[2]) atomicAdd(&c->st_warp, sm.st[2]);
        if (sm.st[3]) atomicAdd(&c->st_block, sm.st[3]);
        if (sm.st[4]) atomicAdd(&c->st_big, sm.st[4]);
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
