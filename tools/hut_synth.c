/* Deterministic synthetic corpus generator for tests and bench.py.
 *
 * Lexicon: n word types built from random Hungarian-looking syllables over a-z + accented vowels
 * (10 % upper-case initials), 1..14 characters, length growing with frequency rank.
 * Documents: log-normal length (median 2 KB, clipped to [64 B, 64 KB]); Zipf(s) word draws joined by single
 * spaces; 12 % punctuation, 4 % digit runs (1-6 digits), 2 % newlines, 0.5 % double spaces / tabs;
 * optionally one "special" snippet (control bytes, multi-byte UTF-8, emoji) every ~special_every bytes.
 *
 * Not part of the engine: a data tool shared by the oracle side and the CUDA side so both see the
 * same bytes.  Build: gcc -O2 -fPIC -shared tools/hut_synth.c -o tools/_build/libhut_synth.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    uint64_t s;
} rng_t;

static inline uint64_t rng_next(rng_t* r) {
    uint64_t z = (r->s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline double rng_unit(rng_t* r) { return (double)(rng_next(r) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint32_t rng_below(rng_t* r, uint32_t n) { return (uint32_t)(((rng_next(r) >> 32) * (uint64_t)n) >> 32); }

static const char* ONSETS[] = {"b", "c", "d", "f", "g", "h", "j", "k", "l", "m", "n", "p", "r", "s", "t", "v", "z",
                               "sz", "gy", "cs", "ny", "ly", "ty", "zs", "", "", "", "k", "t", "l", "n", "r", "m"};
static const char* VOWELS[] = {"a", "e", "i", "o", "u", "a", "e", "e", "o", "i", "\xC3\xA1", "\xC3\xA9", "\xC3\xAD",
                               "\xC3\xB3", "\xC3\xB6", "\xC5\x91", "\xC3\xBA", "\xC3\xBC", "\xC5\xB1"};
static const char* CODAS[] = {"", "", "", "", "n", "k", "t", "l", "r", "s", "m", "g", "d"};
#define NEL(a) (sizeof(a) / sizeof((a)[0]))

/* number of characters (not bytes) */
static int u8len(const char* s, int n) {
    int c = 0;
    for (int i = 0; i < n; ++i)
        if (((unsigned char)s[i] & 0xC0) != 0x80) c++;
    return c;
}

static int upper_first(char* w, int n) {
    unsigned char b = (unsigned char)w[0];
    if (b >= 'a' && b <= 'z') {
        w[0] = (char)(b - 32);
    } else if (b == 0xC3 && n >= 2) {
        unsigned char c = (unsigned char)w[1]; /* á é í ó ö ú ü -> upper = -0x20 */
        if (c >= 0xA0) w[1] = (char)(c - 0x20);
    } else if (b == 0xC5 && n >= 2) {
        unsigned char c = (unsigned char)w[1]; /* ő (0x91) -> Ő (0x90); ű (0xB1) -> Ű (0xB0) */
        if (c == 0x91 || c == 0xB1) w[1] = (char)(c - 1);
    }
    return n;
}

/* open-addressing set of words for de-duplication */
typedef struct {
    uint32_t* slot; /* index+1 into offs */
    size_t cap;
} wset;

static uint64_t hbytes(const uint8_t* p, size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) {
        h ^= p[i];
        h *= 1099511628211ull;
    }
    return h ^ (h >> 29);
}

/* Fills buf with the concatenated words and offs[0..n] with their byte offsets; returns bytes used, or 0
 * if buf_cap is too small.  Word i is the rank-i word of the Zipf distribution. */
uint64_t hut_synth_lexicon(uint64_t seed, uint32_t n, uint8_t* buf, uint64_t buf_cap, uint32_t* offs) {
    rng_t r = {seed ^ 0x2545F4914F6CDD1Dull};
    r.s = rng_next(&r) * 0xD6E8FEB86659FD93ull + seed;
    wset ws;
    ws.cap = 1;
    while (ws.cap < (size_t)n * 4) ws.cap <<= 1;
    ws.slot = calloc(ws.cap, sizeof(uint32_t));
    uint64_t used = 0;
    uint32_t made = 0;
    offs[0] = 0;
    while (made < n) {
        double lr = log(1.0 + (double)made);
        int target = (int)floor(1.2 + lr * (0.45 + 0.55 * rng_unit(&r)));
        if (target < 1) target = 1;
        if (target > 14) target = 14;
        char w[96];
        int wn = 0;
        while (u8len(w, wn) < target) {
            const char* on = ONSETS[rng_below(&r, NEL(ONSETS))];
            const char* vo = VOWELS[rng_below(&r, NEL(VOWELS))];
            const char* co = CODAS[rng_below(&r, NEL(CODAS))];
            for (const char* p = on; *p; ++p) w[wn++] = *p;
            for (const char* p = vo; *p; ++p) w[wn++] = *p;
            for (const char* p = co; *p; ++p) w[wn++] = *p;
        }
        { /* cut back to exactly `target` characters, at a character boundary */
            int chars = 0, k = 0;
            while (k < wn) {
                if (((unsigned char)w[k] & 0xC0) != 0x80) {
                    if (chars == target) break;
                    chars++;
                }
                k++;
            }
            wn = k;
        }
        if (rng_below(&r, 10) == 0) upper_first(w, wn);
        /* dedupe */
        uint64_t h = hbytes((const uint8_t*)w, (size_t)wn);
        size_t i = h & (ws.cap - 1);
        int dup = 0;
        while (ws.slot[i]) {
            uint32_t k = ws.slot[i] - 1;
            if (offs[k + 1] - offs[k] == (uint32_t)wn && memcmp(buf + offs[k], w, (size_t)wn) == 0) {
                dup = 1;
                break;
            }
            i = (i + 1) & (ws.cap - 1);
        }
        if (dup) continue;
        if (used + (uint64_t)wn > buf_cap) {
            free(ws.slot);
            return 0;
        }
        memcpy(buf + used, w, (size_t)wn);
        used += (uint64_t)wn;
        ws.slot[i] = made + 1;
        made++;
        offs[made] = (uint32_t)used;
    }
    free(ws.slot);
    return used;
}

/* Walker alias table for Zipf(s) over n ranks */
typedef struct {
    double* prob;
    uint32_t* alias;
    uint32_t n;
} alias_t;

static void alias_build(alias_t* a, uint32_t n, double s) {
    a->n = n;
    a->prob = malloc(sizeof(double) * n);
    a->alias = malloc(sizeof(uint32_t) * n);
    double* p = malloc(sizeof(double) * n);
    double z = 0;
    for (uint32_t i = 0; i < n; ++i) {
        p[i] = pow((double)(i + 1), -s);
        z += p[i];
    }
    uint32_t* small = malloc(sizeof(uint32_t) * n);
    uint32_t* large = malloc(sizeof(uint32_t) * n);
    uint32_t ns = 0, nl = 0;
    for (uint32_t i = 0; i < n; ++i) {
        p[i] = p[i] / z * n;
        if (p[i] < 1.0) small[ns++] = i; else large[nl++] = i;
    }
    while (ns && nl) {
        uint32_t sm = small[--ns], lg = large[--nl];
        a->prob[sm] = p[sm];
        a->alias[sm] = lg;
        p[lg] = (p[lg] + p[sm]) - 1.0;
        if (p[lg] < 1.0) small[ns++] = lg; else large[nl++] = lg;
    }
    while (nl) { uint32_t i = large[--nl]; a->prob[i] = 1.0; a->alias[i] = i; }
    while (ns) { uint32_t i = small[--ns]; a->prob[i] = 1.0; a->alias[i] = i; }
    free(p); free(small); free(large);
}

static inline uint32_t alias_draw(const alias_t* a, rng_t* r) {
    uint64_t x = rng_next(r);
    uint32_t i = (uint32_t)(((x >> 32) * (uint64_t)a->n) >> 32);
    double u = (double)(x & 0xFFFFFFFFu) * (1.0 / 4294967296.0);
    return u < a->prob[i] ? i : a->alias[i];
}

static double rng_normal(rng_t* r) {
    double u1 = rng_unit(r), u2 = rng_unit(r);
    if (u1 < 1e-300) u1 = 1e-300;
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

static const char* PUNCT[] = {".", ",", ";", ":", "!", "?", "-", "(", ")", "\"", ".", ",", ",", "."};
/* snippets whose bytes exercise the replacement table / multi-byte paths (config C4) */
static const char* SPECIALS[] = {"\n", "\t", "\r\n", "\x7F", "\xC5\x91", "\xC5\xB1", "\xE2\x82\xAC",
                                 "\xF0\x9F\x98\x82", "\xE6\x97\xA5\xE6\x9C\xAC", "\xC2\xA0", "\xC2\xAD",
                                 "\xC3\xB1", "\xE2\x80\x94", "\x0B", "\x0C", "\x01"};

/* Generates documents until target_bytes are produced.  out must hold target_bytes + 70000 bytes;
 * doc_offs must hold max_docs + 1 entries.  Returns the number of documents; *out_bytes = bytes written.
 * lex/lex_offs come from hut_synth_lexicon. */
uint64_t hut_synth_corpus(uint64_t seed, const uint8_t* lex, const uint32_t* lex_offs, uint32_t n_types, double zipf_s,
                          uint64_t target_bytes, uint32_t special_every, uint8_t* out, uint64_t* doc_offs,
                          uint64_t max_docs, uint64_t* out_bytes) {
    rng_t r = {seed ^ 0x5DEECE66Dull};
    r.s = rng_next(&r) * 0xD6E8FEB86659FD93ull + seed; /* decorrelate nearby seeds */
    r.s ^= rng_next(&r) >> 7;
    alias_t al;
    alias_build(&al, n_types, zipf_s);
    uint64_t pos = 0, nd = 0;
    doc_offs[0] = 0;
    uint64_t next_special = special_every ? special_every / 2 + rng_below(&r, special_every) : ~0ull;
    while (pos < target_bytes && nd < max_docs) {
        double ln = log(2048.0) + 1.0 * rng_normal(&r);
        double dl = exp(ln);
        if (dl < 64) dl = 64;
        if (dl > 65536) dl = 65536;
        uint64_t end = pos + (uint64_t)dl;
        if (end > target_bytes) end = target_bytes;
        if (end < pos + 16) end = pos + 16;
        int first = 1;
        while (pos < end) {
            uint32_t roll = rng_below(&r, 1000);
            if (!first) {
                if (roll < 20) { /* 2 % newline(s) */
                    out[pos++] = '\n';
                    if (roll < 8) out[pos++] = '\n';
                } else if (roll < 25) { /* 0.5 % double space / tab */
                    if (roll < 23) { out[pos++] = ' '; out[pos++] = ' '; } else out[pos++] = '\t';
                } else {
                    out[pos++] = ' ';
                }
            }
            first = 0;
            if (pos >= next_special) {
                const char* sp = SPECIALS[rng_below(&r, NEL(SPECIALS))];
                size_t sl = strlen(sp);
                memcpy(out + pos, sp, sl);
                pos += sl;
                next_special = pos + special_every / 2 + rng_below(&r, special_every);
                continue;
            }
            uint32_t kind = rng_below(&r, 100);
            if (kind < 4) { /* digit run */
                uint32_t nd_ = 1 + rng_below(&r, 6);
                for (uint32_t k = 0; k < nd_; ++k) out[pos++] = (uint8_t)('0' + rng_below(&r, 10));
            } else {
                uint32_t w = alias_draw(&al, &r);
                uint32_t wl = lex_offs[w + 1] - lex_offs[w];
                memcpy(out + pos, lex + lex_offs[w], wl);
                pos += wl;
            }
            if (rng_below(&r, 100) < 12) { /* punctuation, attached or spaced */
                const char* pc = PUNCT[rng_below(&r, NEL(PUNCT))];
                if (rng_below(&r, 4) == 0) out[pos++] = ' ';
                out[pos++] = (uint8_t)pc[0];
            }
        }
        nd++;
        doc_offs[nd] = pos;
    }
    free(al.prob);
    free(al.alias);
    *out_bytes = pos;
    return nd;
}
