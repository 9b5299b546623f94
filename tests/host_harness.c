/* TEST-ONLY scalar driver for the engine's HOST-side pieces: hut_model.c (loaders + derived unit / pair /
 * symbol tables) and hut_classify.h (local pretoken-start rule).  It walks one document with the same tables the
 * CUDA kernels read, so `pytest -m "not gpu"` can check table construction and boundary
 * classification against the oracle without a GPU.  It is NOT part of libhutoken_b200.so and is never a
 * fallback: the shipped library has no CPU encode path.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../hutoken_b200/csrc/hut_classify.h"
#include "../hutoken_b200/csrc/hut_model.h"

#define NONE HUT_UNIT_MASK

/* class codes + start flags of one document; returns number of pretokens, starts[] filled */
size_t harness_pretokens(const uint8_t* t, size_t n, uint32_t* starts, size_t cap) {
    uint8_t* code = calloc(n + 8, 1);
    for (size_t i = 0; i < n; ++i) {
        uint32_t b1 = i + 1 < n ? t[i + 1] : 0, b2 = i + 2 < n ? t[i + 2] : 0, b3 = i + 3 < n ? t[i + 3] : 0;
        uint32_t r = hut_lead_class(t[i], b1, b2, b3);
        code[i] = (uint8_t)((r & 15u) | (((r >> 4) - 1u) << 4));
    }
    for (size_t i = 0; i < n; ++i) {
        uint32_t extra = (code[i] >> 4) & 3u;
        if ((code[i] & 7u) == HC_CONT) continue;
        for (uint32_t k = 1; k <= extra && i + k < n; ++k) code[i + k] = HC_CONT;
    }
    size_t cnt = 0;
    for (size_t i = 0; i < n; ++i) {
        uint32_t c = code[i] & 7u;
        if (c == HC_END) break; /* NUL ends the document */
        if (c == HC_CONT) continue;
        uint32_t p1 = i ? (code[i - 1] & 7u) : HC_END;
        uint32_t pc = p1;
        for (int k = 2; pc == HC_CONT && k <= 4 && (size_t)k <= i; ++k) pc = code[i - k] & 7u;
        int lone = (i == 1) || (i >= 2 && (code[i - 2] & 7u) != HC_SPACE);
        if (hut_is_start(c, i == 0, p1, pc, lone)) {
            if (cnt < cap) starts[cnt] = (uint32_t)i;
            cnt++;
        }
    }
    free(code);
    return cnt;
}

static uint32_t rank_of(const hut_model* m, uint32_t mg) { return mg == NONE ? 0xFFFFFFFFu : (uint32_t)m->rank_of[mg]; }

/* symbols of pretoken t[s..e) into u[]; returns count */
static size_t symbols(const hut_model* m, const uint8_t* t, size_t s, size_t e, uint32_t* u) {
    size_t cnt = 0;
    for (size_t p = s; p < e; ++p) {
        uint8_t b = t[p];
        if (!m->is_byte_encoder && p != s && (b & 0xC0) == 0x80) continue;
        uint32_t lo = m->sym_off[b], hi = m->sym_off[b + 1];
        if (hi > lo) {
            for (uint32_t k = lo; k < hi; ++k) u[cnt++] = m->sym_pool[k];
        } else if (m->is_byte_encoder) {
            u[cnt++] = m->inert;
        } else {
            size_t l = 1;
            while (l < 4 && p + l < e && (t[p + l] & 0xC0) == 0x80) l++;
            u[cnt++] = hut_model_char_lookup(m, t + p, (uint32_t)l);
        }
    }
    return cnt;
}

static size_t merge(const hut_model* m, uint32_t* u, size_t n) {
    for (;;) {
        uint32_t best = 0xFFFFFFFFu, bm = NONE;
        size_t bi = 0;
        for (size_t i = 0; i + 1 < n; ++i) {
            if (u[i] >= m->inert || u[i + 1] >= m->inert) continue;
            uint32_t mg = hut_model_pair_lookup(m, u[i], u[i + 1]);
            uint32_t rk = rank_of(m, mg);
            if (rk < best) {
                best = rk;
                bi = i;
                bm = mg;
            }
        }
        if (bm == NONE) return n;
        u[bi] = bm;
        memmove(u + bi + 1, u + bi + 2, (n - bi - 2) * sizeof *u);
        n--;
    }
}

/* table-driven encode of one document; returns id count (ids truncated to cap) */
size_t harness_encode(const hut_model* m, const uint8_t* t, size_t n, int32_t* out, size_t cap) {
    const uint8_t* z = memchr(t, 0, n);
    if (z) n = (size_t)(z - t);
    uint32_t* starts = malloc((n + 1) * sizeof *starts);
    size_t np = harness_pretokens(t, n, starts, n + 1);
    uint32_t* u = malloc((n * m->max_sym_per_char + m->n_prefix_syms + 8) * sizeof *u);
    size_t w = 0;
    for (size_t k = 0; k < np; ++k) {
        size_t s = starts[k], e = k + 1 < np ? starts[k + 1] : n;
        if (e - s > 262144) break; /* core.c:402-407 */
        size_t cnt = 0;
        int first = m->has_prefix && k == 0;
        if (first && t[s] == ' ')
            for (uint32_t i = 0; i < m->n_prefix_alone; ++i, ++w)
                if (w < cap) out[w] = m->prefix_alone_ids[i];
        if (first && t[s] != ' ')
            for (uint32_t i = 0; i < m->n_prefix_syms; ++i) u[cnt++] = m->prefix_syms[i];
        cnt += symbols(m, t, s, e, u + cnt);
        cnt = merge(m, u, cnt);
        for (size_t i = 0; i < cnt; ++i, ++w)
            if (w < cap) out[w] = m->out_id[u[i]];
    }
    free(u);
    free(starts);
    return w;
}

hut_model* harness_load(const char* vocab, const char* special, const char* prefix, int byte_mode, const char* merges,
                        char* msg, size_t msglen, int* status) {
    hut_model* m = NULL;
    *status = hut_model_load(&m, vocab, special, prefix, byte_mode, merges, msg, msglen);
    return m;
}

void harness_free(hut_model* m) { hut_model_free(m); }

/* table-driven decode of one id list with the decode tables the device reads (dec_off / dec_len / dec_pool, trie
 * edges / outputs / depths, prefix): returns the byte count, or (size_t)-1 for an undecodable id */
size_t harness_decode(const hut_model* m, const int32_t* ids, size_t n, uint8_t* out, size_t cap) {
    size_t total = 0;
    for (size_t i = 0; i < n; ++i) {
        if (ids[i] < 0 || (uint32_t)ids[i] >= m->vocab_lines || m->dec_len[ids[i]] == HUT_DEC_MISSING) return (size_t)-1;
        total += m->dec_len[ids[i]];
    }
    uint8_t* raw = calloc(total + 8, 1);
    size_t at = 0;
    for (size_t i = 0; i < n; ++i) {
        memcpy(raw + at, m->dec_pool + m->dec_off[ids[i]], m->dec_len[ids[i]]);
        at += m->dec_len[ids[i]];
    }
    size_t p = 0, w = 0;
    if (m->prefix_len && total >= m->prefix_len && memcmp(raw, m->prefix_bytes, m->prefix_len) == 0) p = m->prefix_len;
    while (p < total) {
        uint32_t node = m->trie_root_child[raw[p]], step = 0, outb = 0;
        size_t q = p + 1;
        while (node) {
            if (m->trie_out[node] >= 0) {
                step = m->trie_depth[node];
                outb = (uint32_t)m->trie_out[node];
            }
            if (q >= total) break;
            node = hut_model_trie_child(m, node, raw[q++]);
        }
        if (step) {
            if (w < cap) out[w] = (uint8_t)outb;
            w++;
            p += step;
            continue;
        }
        const uint8_t* c = raw + p;
        uint32_t cl = (c[0] & 0x80) == 0 ? 1 : (c[0] & 0xE0) == 0xC0 ? 2 : (c[0] & 0xF0) == 0xE0 ? 3 : (c[0] & 0xF8) == 0xF0 ? 4 : 1;
        if (m->is_byte_encoder) {
            uint32_t cp = cl == 1 ? ((c[0] & 0x80) ? 0xFFFD : c[0])
                          : cl == 2 ? ((c[0] & 0x1Fu) << 6) | (c[1] & 0x3Fu)
                          : cl == 3 ? ((c[0] & 0x0Fu) << 12) | ((c[1] & 0x3Fu) << 6) | (c[2] & 0x3Fu)
                                    : ((c[0] & 0x07u) << 18) | ((c[1] & 0x3Fu) << 12) | ((c[2] & 0x3Fu) << 6) | (c[3] & 0x3Fu);
            if (w < cap) out[w] = cp < 256 ? (uint8_t)cp : (uint8_t)'?';
            w++;
        } else {
            for (uint32_t k = 0; k < cl && p + k < total; ++k, ++w)
                if (w < cap) out[w] = c[k];
        }
        p += cl;
    }
    free(raw);
    return w;
}

void harness_info(const hut_model* m, uint32_t* out) {
    out[0] = m->n_vocab;
    out[1] = m->n_units;
    out[2] = m->pair_count;
    out[3] = m->pair_cap;
    out[4] = (uint32_t)m->distinct_values;
    out[5] = (uint32_t)m->identity_ids;
    out[6] = m->max_sym_per_char;
    out[7] = m->vocab_lines;
    out[8] = (uint32_t)m->monotone;
}

/* ---- bit-sliced classifier (hut_bitslice.h), chunk by chunk the way the warp lanes run it ---- */
#include "../hutoken_b200/csrc/hut_bitslice.h"

/* t[0..n): packed documents; isds[i] != 0 marks a document start.  Writes every pretoken start (over all
 * documents) into starts[]; *exotic = 1 if an over-long form was seen (the kernel then uses the generic path).
 * NUL bytes end nothing here: the caller passes NUL-free text. */
size_t harness_pretokens_bs(const uint8_t* t, size_t n, const uint8_t* isds, uint32_t* starts, size_t cap,
                            int* exotic) {
    size_t nch = (n + 31) / 32 + 1;
    uint8_t* buf = calloc(nch * 32 + 32, 1);
    memcpy(buf, t, n);
    hut_bs_local* loc = calloc(nch + 1, sizeof *loc);
    hut_bs_mid* mid = calloc(nch + 1, sizeof *mid);
    for (size_t c = 0; c < nch; ++c) {
        uint32_t w[8], p[8], ds = 0;
        memcpy(w, buf + c * 32, 32);
        for (int i = 0; i < 32; ++i)
            if (c * 32 + i < n && isds[c * 32 + i]) ds |= 1u << i;
        hut_bitslice32(w, p);
        hut_bs_classify_local(p, ds, &loc[c]);
    }
    memset(&loc[nch], 0, sizeof loc[nch]);
    *exotic = 0;
    size_t cnt = 0;
    for (size_t c = 0; c < nch; ++c) {
        hut_bs_lookahead(&loc[c], hut_bs_pack_next(&loc[c + 1]), &mid[c]);
        if (mid[c].EX) *exotic = 1;
    }
    for (size_t c = 0; c < nch; ++c) {
        uint32_t px = c ? hut_bs_pack_prev(&loc[c - 1], &mid[c - 1]) : 0u, st, nu;
        hut_bs_starts(&loc[c], &mid[c], px, &st, &nu);
        for (int i = 0; i < 32; ++i)
            if (((st >> i) & 1u) && c * 32 + i < n) {
                if (cnt < cap) starts[cnt] = (uint32_t)(c * 32 + i);
                cnt++;
            }
    }
    free(buf);
    free(loc);
    free(mid);
    return cnt;
}

/* exhaustive check of hut_bitslice32 on caller-provided bytes: returns 0 if every plane bit is right */
int harness_bitslice_check(const uint8_t* bytes32) {
    uint32_t w[8], p[8];
    memcpy(w, bytes32, 32);
    hut_bitslice32(w, p);
    for (int k = 0; k < 8; ++k)
        for (int i = 0; i < 32; ++i)
            if (((p[k] >> i) & 1u) != (uint32_t)((bytes32[i] >> k) & 1)) return 1;
    return 0;
}
