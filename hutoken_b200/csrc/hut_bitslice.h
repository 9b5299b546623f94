/* Bit-sliced pretoken-boundary classifier: 32 text bytes per lane, every later operation handles all 32
 * positions with one 32-bit logic instruction.
 *
 * Replaces the reference's sequential parser (parser.c:24-183: parser_next_token, consume_while, decode_utf8,
 * is_custom_alpha / is_digit / is_other / is_whitespace) for the fused tile kernel.  Same LOCAL rule as
 * hut_classify.h (which stays the scalar restatement used for the rare generic path and by the tests):
 *   1. the 32 bytes are transposed into 8 bit planes (plane k, bit i = bit k of byte i),
 *   2. character classes are boolean functions of the planes,
 *   3. multi-byte validity / coverage and the start rule are shifts + logic on the class masks, with a few
 *      carry bits exchanged with the neighbouring lanes (one packed word each way).
 * Over-long UTF-8 forms (C0/C1 xx, E0 80..9F xx, F0 80..8F xx xx), which decode_utf8 accepts and classifies by
 * their decoded value, are only DETECTED here (mask EX): the caller then falls back to hut_classify.h.
 *
 * Shared by the CUDA kernels and tests/host_harness.c (CPU check against the oracle's parser).
 */
#ifndef HUT_BITSLICE_H
#define HUT_BITSLICE_H

#include <stdint.h>

#ifdef __CUDACC__
#define HUT_BS __host__ __device__ __forceinline__
#else
#define HUT_BS static inline
#endif

HUT_BS uint32_t hut_byte_perm(uint32_t a, uint32_t b, uint32_t s) {
#ifdef __CUDA_ARCH__
    return __byte_perm(a, b, s);
#else
    uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((s >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
    return r;
#endif
}

/* 8x8 bit-matrix transpose of the 8 bytes in (lo, hi): afterwards byte k of lo holds bit k (k = 0..3) and byte k
 * of hi bit 4+k of each of the 8 input bytes (input byte i -> bit i).  Three butterfly stages. */
HUT_BS void hut_transpose8(uint32_t* lo, uint32_t* hi) {
    uint32_t x = *lo, y = *hi, t;
    t = (x ^ (x >> 7)) & 0x00AA00AAu; x = x ^ t ^ (t << 7);
    t = (y ^ (y >> 7)) & 0x00AA00AAu; y = y ^ t ^ (t << 7);
    t = (x ^ (x >> 14)) & 0x0000CCCCu; x = x ^ t ^ (t << 14);
    t = (y ^ (y >> 14)) & 0x0000CCCCu; y = y ^ t ^ (t << 14);
    t = (x ^ (y << 4)) & 0xF0F0F0F0u;
    *lo = x ^ t;
    *hi = y ^ (t >> 4);
}

/* planes[k] bit i = bit k of byte i of the 32 bytes in w[0..7] (little-endian words) */
HUT_BS void hut_bitslice32(const uint32_t w[8], uint32_t planes[8]) {
    uint32_t x[4], y[4];
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int j = 0; j < 4; ++j) {
        x[j] = w[2 * j];
        y[j] = w[2 * j + 1];
        hut_transpose8(&x[j], &y[j]);
    }
    uint32_t t0, t1;
    t0 = hut_byte_perm(x[0], x[1], 0x5140u); t1 = hut_byte_perm(x[2], x[3], 0x5140u);
    planes[0] = hut_byte_perm(t0, t1, 0x5410u); planes[1] = hut_byte_perm(t0, t1, 0x7632u);
    t0 = hut_byte_perm(x[0], x[1], 0x7362u); t1 = hut_byte_perm(x[2], x[3], 0x7362u);
    planes[2] = hut_byte_perm(t0, t1, 0x5410u); planes[3] = hut_byte_perm(t0, t1, 0x7632u);
    t0 = hut_byte_perm(y[0], y[1], 0x5140u); t1 = hut_byte_perm(y[2], y[3], 0x5140u);
    planes[4] = hut_byte_perm(t0, t1, 0x5410u); planes[5] = hut_byte_perm(t0, t1, 0x7632u);
    t0 = hut_byte_perm(y[0], y[1], 0x7362u); t1 = hut_byte_perm(y[2], y[3], 0x7362u);
    planes[6] = hut_byte_perm(t0, t1, 0x5410u); planes[7] = hut_byte_perm(t0, t1, 0x7632u);
}

/* Class masks of one 32-byte chunk that depend only on the chunk itself (+ its document-start mask DS). */
typedef struct hut_bs_local {
    uint32_t AL, DG, SP, WS, NU, OA; /* ASCII letter / digit / 0x20 / \t\n\v\f\r / 0x00 / any other ASCII */
    uint32_t HI;                     /* byte >= 0x80 */
    uint32_t CT;                     /* continuation-type byte (10xxxxxx) that is not a document start */
    uint32_t L2, L3, L4;             /* lead-type bytes 110xxxxx / 1110xxxx / 11110xxx */
    uint32_t C0C1, E0, F0;           /* leads whose over-long forms need the generic path */
    uint32_t C3, C5;                 /* leads of the 18 Hungarian letters (parser.c:102-129) */
    uint32_t S1, S2;                 /* continuation bytes completing a Hungarian letter after C3 / C5 */
    uint32_t NA0, N90;               /* continuation bytes < 0xA0 / < 0x90 (over-long after E0 / F0) */
    uint32_t DS;
} hut_bs_local;

HUT_BS void hut_bs_classify_local(const uint32_t p[8], uint32_t ds, hut_bs_local* o) {
    const uint32_t b0 = p[0], b1 = p[1], b2 = p[2], b3 = p[3], b4 = p[4], b5 = p[5], b6 = p[6], b7 = p[7];
    const uint32_t asc = ~b7;
    const uint32_t low5nz = b4 | b3 | b2 | b1 | b0;
    /* letters: 0x41..0x5A, 0x61..0x7A  <=> 01x + low5 in 1..26 */
    o->AL = asc & b6 & low5nz & ~(b4 & b3 & (b2 | (b1 & b0)));
    /* digits: 0x30..0x39 <=> 0011 + low4 <= 9 */
    o->DG = asc & ~b6 & b5 & b4 & ~(b3 & (b2 | b1));
    o->SP = asc & ~b6 & b5 & ~low5nz;
    /* 0x09..0x0D <=> 0000 1 + low3 in 1..5 */
    o->WS = asc & ~b6 & ~b5 & ~b4 & b3 & ~(b2 & b1) & (b2 | b1 | b0);
    o->NU = asc & ~b6 & ~b5 & ~low5nz;
    o->OA = asc & ~(o->AL | o->DG | o->SP | o->WS | o->NU);
    o->HI = b7;
    o->CT = b7 & ~b6 & ~ds;
    o->L2 = b7 & b6 & ~b5;
    o->L3 = b7 & b6 & b5 & ~b4;
    o->L4 = b7 & b6 & b5 & b4 & ~b3;
    o->C0C1 = o->L2 & ~(b4 | b3 | b2 | b1);
    o->E0 = o->L3 & ~(b3 | b2 | b1 | b0);
    o->F0 = o->L4 & ~(b2 | b1 | b0);
    o->C3 = o->L2 & ~(b4 | b3 | b2) & b1 & b0;
    o->C5 = o->L2 & ~(b4 | b3 | b1) & b2 & b0;
    /* after C3: (byte & 0xDF) in {81, 89, 8D, 93, 96, 9A, 9C}: low5 in {1, 9, 13, 19, 22, 26, 28} */
    {
        const uint32_t n4 = ~b4, n3 = ~b3, n2 = ~b2, n1 = ~b1, n0 = ~b0;
        const uint32_t s = (n4 & n3 & n2 & n1 & b0) | (n4 & b3 & n2 & n1 & b0) | (n4 & b3 & b2 & n1 & b0) |
                           (b4 & n3 & n2 & b1 & b0) | (b4 & n3 & b2 & b1 & n0) | (b4 & b3 & n2 & b1 & n0) |
                           (b4 & b3 & b2 & n1 & n0);
        o->S1 = o->CT & s;
        /* after C5: 0x90, 0x91, 0xB0, 0xB1 */
        o->S2 = o->CT & b4 & n3 & n2 & n1;
    }
    o->NA0 = o->CT & ~b5;
    o->N90 = o->CT & ~b5 & ~b4;
    o->DS = ds;
}

/* carry word a lane needs from the NEXT chunk (lane + 1): built from that chunk's local masks */
HUT_BS uint32_t hut_bs_pack_next(const hut_bs_local* n) {
    return (n->CT & 7u) | ((n->S1 & 1u) << 3) | ((n->S2 & 1u) << 4) | ((n->NA0 & 1u) << 5) | ((n->N90 & 1u) << 6);
}

/* masks that need look-ahead only */
typedef struct hut_bs_mid {
    uint32_t V2, V3, V4; /* valid multi-byte leads */
    uint32_t HA;         /* leads of 2-byte Hungarian letters */
    uint32_t EX;         /* over-long forms: generic path required */
} hut_bs_mid;

HUT_BS void hut_bs_lookahead(const hut_bs_local* c, uint32_t nx, hut_bs_mid* m) {
    const uint32_t nct = nx & 7u;
    const uint32_t ct1 = (c->CT >> 1) | (nct << 31);
    const uint32_t ct2 = (c->CT >> 2) | (nct << 30);
    const uint32_t ct3 = (c->CT >> 3) | (nct << 29);
    m->V2 = c->L2 & ct1;
    m->V3 = c->L3 & ct1 & ct2;
    m->V4 = c->L4 & ct1 & ct2 & ct3;
    const uint32_t s1n = (c->S1 >> 1) | (((nx >> 3) & 1u) << 31);
    const uint32_t s2n = (c->S2 >> 1) | (((nx >> 4) & 1u) << 31);
    const uint32_t a0n = (c->NA0 >> 1) | (((nx >> 5) & 1u) << 31);
    const uint32_t n9n = (c->N90 >> 1) | (((nx >> 6) & 1u) << 31);
    m->HA = m->V2 & ((c->C3 & s1n) | (c->C5 & s2n));
    m->EX = (m->V2 & c->C0C1) | (m->V3 & c->E0 & a0n) | (m->V4 & c->F0 & n9n);
}

/* carry word a lane needs from the PREVIOUS chunk (lane - 1).  Every field is a function of that chunk's bits
 * 28..31 only, so it does not depend on the previous chunk's own carry-in. */
HUT_BS uint32_t hut_bs_pack_prev(const hut_bs_local* c, const hut_bs_mid* m) {
    const uint32_t va = m->V2 | m->V3 | m->V4, v34 = m->V3 | m->V4, v4 = m->V4;
    /* cover class of byte 31 */
    const uint32_t cov31 = ((va >> 30) | (v34 >> 29) | (v4 >> 28)) & 1u;
    const uint32_t pha31 = (m->HA >> 30) & 1u;
    const uint32_t a31 = ((c->AL | m->HA) >> 31) | pha31;
    const uint32_t d31 = c->DG >> 31;
    const uint32_t o31 = ((c->OA | (va & ~m->HA)) >> 31) | (cov31 & ~pha31);
    return (va >> 31) | ((v34 >> 30) << 1) | ((v4 >> 29) << 3) | ((m->HA >> 31) << 6) | ((a31 & 1u) << 7) |
           ((d31 & 1u) << 8) | ((o31 & 1u) << 9) | ((c->SP >> 30) << 10) | ((c->DS >> 31) << 12);
}

/* Final masks of the chunk: ST = pretoken starts, NU = 0x00 bytes (end of text).  px = hut_bs_pack_prev of the
 * previous chunk (0 at the very beginning of the text: the first byte is a document start anyway). */
HUT_BS void hut_bs_starts(const hut_bs_local* c, const hut_bs_mid* m, uint32_t px, uint32_t* ST, uint32_t* NU) {
    const uint32_t va = m->V2 | m->V3 | m->V4, v34 = m->V3 | m->V4, v4 = m->V4;
    const uint32_t cov = ((va << 1) | (px & 1u)) | ((v34 << 2) | ((px >> 1) & 3u)) | ((v4 << 3) | ((px >> 3) & 7u));
    const uint32_t bad = c->HI & ~va & ~cov; /* stray continuation bytes, truncated leads, 0xF8..0xFF */
    const uint32_t pha = (m->HA << 1) | ((px >> 6) & 1u);
    const uint32_t la = c->AL | m->HA;            /* letter leads */
    const uint32_t lo = c->OA | (va & ~m->HA);    /* OTHER leads */
    const uint32_t ca = la | pha;                 /* bytes belonging to a letter */
    const uint32_t co = lo | (cov & ~pha);        /* bytes belonging to an OTHER char */
    const uint32_t pA = (ca << 1) | ((px >> 7) & 1u);
    const uint32_t pD = (c->DG << 1) | ((px >> 8) & 1u);
    const uint32_t pO = (co << 1) | ((px >> 9) & 1u);
    const uint32_t pSP = (c->SP << 1) | ((px >> 11) & 1u);
    const uint32_t ppSP = (c->SP << 2) | ((px >> 10) & 3u);
    const uint32_t pDS = (c->DS << 1) | ((px >> 12) & 1u);
    const uint32_t glue = pSP & (~ppSP | pDS); /* the previous byte is a lone space: it carries this run */
    const uint32_t st = (la & ~pA & ~glue) | (c->DG & ~pD & ~glue) | (lo & ~pO & ~glue) | (c->SP & ~pSP) | c->WS |
                        bad | c->DS;
    *ST = st & ~c->NU;
    *NU = c->NU;
}

#endif
