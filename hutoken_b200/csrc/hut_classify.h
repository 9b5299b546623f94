/* Per-byte character classes and the LOCAL pretoken-start rule.
 *
 * Replaces the sequential state machine of the reference's parser (parser.c:24-183: parser_next_token,
 * consume_while, decode_utf8, is_custom_alpha / is_digit / is_other / is_whitespace) with a formulation
 * in which every byte decides from a bounded neighbourhood (4 bytes back, 3 forward) whether a pretoken
 * starts at it -- the property a data-parallel pass needs.
 *
 * Shared by the CUDA kernels (hut_engine.cu) and by tests/host_harness.c, which checks this header against
 * the oracle on the CPU.  Pure functions, no state.
 */
#ifndef HUT_CLASSIFY_H
#define HUT_CLASSIFY_H

#include <stdint.h>

#ifdef __CUDACC__
#define HUT_HD __host__ __device__ __forceinline__
#else
#define HUT_HD static inline
#endif

/* class codes, one per byte */
enum {
    HC_END = 0,   /* NUL / outside the document */
    HC_ALPHA = 1, /* [A-Za-z] + 18 Hungarian letters (parser.c:102-129) */
    HC_DIGIT = 2, /* [0-9] (parser.c:131-133) */
    HC_OTHER = 3, /* any other valid non-space code point (parser.c:135-138) */
    HC_SPACE = 4, /* 0x20 */
    HC_WS = 5,    /* \t \n \v \f \r : single-byte pretokens (parser.c:84-87, 140-142 in the C locale) */
    HC_BAD = 6,   /* byte that decode_utf8 rejects at this position: single-byte pretoken */
    HC_CONT = 7   /* continuation byte consumed by a valid multi-byte char */
};

HUT_HD uint32_t hut_cp_class(uint32_t cp) {
    if (cp == 0) return HC_BAD;
    if (((cp | 0x20u) - 'a') < 26u) return HC_ALPHA;
    if ((cp - '0') < 10u) return HC_DIGIT;
    if (cp == 0x20u || (cp - 9u) < 5u) return HC_BAD; /* only reachable through over-long forms */
    switch (cp) {
        case 0xE1: case 0xE9: case 0xED: case 0xF3: case 0xFA: case 0x151: case 0x171: case 0xFC: case 0xF6:
        case 0xC1: case 0xC9: case 0xCD: case 0xD3: case 0xDA: case 0x150: case 0x170: case 0xDC: case 0xD6:
            return HC_ALPHA;
        default:
            return HC_OTHER;
    }
}

/* Class of the character whose first byte is b0, with b1..b3 the following bytes of the SAME document (0 past
 * its end).  Returns code | (length << 4); continuation bytes come back as HC_BAD (stray) -- the caller
 * overwrites the bytes covered by a valid lead with HC_CONT.  decode_utf8: parser.c:144-183. */
HUT_HD uint32_t hut_lead_class(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3) {
    if (b0 < 0x80u) {
        uint32_t c;
        if (b0 == 0) c = HC_END;
        else if (((b0 | 0x20u) - 'a') < 26u) c = HC_ALPHA;
        else if ((b0 - '0') < 10u) c = HC_DIGIT;
        else if (b0 == 0x20u) c = HC_SPACE;
        else if ((b0 - 9u) < 5u) c = HC_WS;
        else c = HC_OTHER;
        return c | (1u << 4);
    }
    uint32_t cp, len;
    if ((b0 & 0xE0u) == 0xC0u) {
        if ((b1 & 0xC0u) != 0x80u) return HC_BAD | (1u << 4);
        cp = ((b0 & 0x1Fu) << 6) | (b1 & 0x3Fu);
        len = 2;
    } else if ((b0 & 0xF0u) == 0xE0u) {
        if ((b1 & 0xC0u) != 0x80u || (b2 & 0xC0u) != 0x80u) return HC_BAD | (1u << 4);
        cp = ((b0 & 0x0Fu) << 12) | ((b1 & 0x3Fu) << 6) | (b2 & 0x3Fu);
        len = 3;
    } else if ((b0 & 0xF8u) == 0xF0u) {
        if ((b1 & 0xC0u) != 0x80u || (b2 & 0xC0u) != 0x80u || (b3 & 0xC0u) != 0x80u) return HC_BAD | (1u << 4);
        cp = ((b0 & 0x07u) << 18) | ((b1 & 0x3Fu) << 12) | ((b2 & 0x3Fu) << 6) | (b3 & 0x3Fu);
        len = 4;
    } else {
        return HC_BAD | (1u << 4); /* stray continuation byte or 0xF8..0xFF */
    }
    uint32_t c = hut_cp_class(cp);
    if (c == HC_BAD) return HC_BAD | (1u << 4); /* cp == 0 or over-long whitespace: the lead is a 1-byte token */
    return c | (len << 4);
}

/* Does a pretoken start at a byte?  Inputs, all for the same document:
 *   c        class code of this byte (not HC_CONT / HC_END)
 *   docstart this byte is the first of its document
 *   p1       code of the byte before (HC_END if none)
 *   pc       code of the CHARACTER before (p1 unless p1 is HC_CONT: then the code of its lead byte)
 *   lone     meaningful when p1 == HC_SPACE: that space is not preceded by another space in the document
 * Rule (parser.c:24-88): a run of ALPHA / DIGIT / OTHER absorbs exactly one preceding space when that space
 * is alone; two or more spaces form their own pretoken; \t\n\v\f\r and rejected bytes stand alone. */
HUT_HD int hut_is_start(uint32_t c, int docstart, uint32_t p1, uint32_t pc, int lone) {
    if (docstart) return 1;
    if (c == HC_SPACE) return p1 != HC_SPACE;
    if (c >= HC_WS) return 1; /* HC_WS, HC_BAD */
    if (pc == c) return 0;
    if (p1 == HC_SPACE) return !lone;
    return 1;
}

#endif
