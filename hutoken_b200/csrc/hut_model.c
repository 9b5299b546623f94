/* Host-side model loader and derived-table builder.  See hut_model.h.
 *
 * File formats and loader quirks follow the reference's p_initialize (lib.c:243-566) and
 * hex_str_to_ascii (helper.c:82-128); SURVEY.md Appendix A.5 lists them.  The pair table is new: the
 * reference ranks a pair by hashing the concatenated bytes into its string map (core.c:700-722); here
 * every split of every vocab string into two possible tokens becomes one (L,R)->M slot.
 */
#define _GNU_SOURCE
#include "hut_model.h"

#include <errno.h>
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------
 * bytes -> index map (open addressing, FNV-1a)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    uint32_t* slot; /* index + 1 */
    size_t cap;
    size_t count;
    uint8_t*** keys; /* -> array of key pointers owned elsewhere */
    uint32_t** lens;
} smap;

static uint64_t fnv(const uint8_t* p, size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) {
        h ^= p[i];
        h *= 1099511628211ull;
    }
    return h ^ (h >> 31);
}

typedef struct {
    uint8_t** str;
    uint32_t* len;
    int32_t* val;
    size_t n, cap;
    uint32_t* slot;
    size_t scap;
} strtab;

static int strtab_init(strtab* t) {
    memset(t, 0, sizeof *t);
    t->scap = 1024;
    t->slot = calloc(t->scap, sizeof(uint32_t));
    return t->slot != NULL;
}

static void strtab_free(strtab* t, int free_strings) {
    if (free_strings)
        for (size_t i = 0; i < t->n; ++i) free(t->str[i]);
    free(t->str);
    free(t->len);
    free(t->val);
    free(t->slot);
    memset(t, 0, sizeof *t);
}

static long strtab_find(const strtab* t, const uint8_t* k, size_t n) {
    size_t mask = t->scap - 1, i = fnv(k, n) & mask;
    while (t->slot[i]) {
        uint32_t j = t->slot[i] - 1;
        if (t->len[j] == n && memcmp(t->str[j], k, n) == 0) return (long)j;
        i = (i + 1) & mask;
    }
    return -1;
}

static int strtab_rehash(strtab* t, size_t ncap) {
    uint32_t* ns = calloc(ncap, sizeof(uint32_t));
    if (!ns) return 0;
    for (size_t j = 0; j < t->n; ++j) {
        size_t i = fnv(t->str[j], t->len[j]) & (ncap - 1);
        while (ns[i]) i = (i + 1) & (ncap - 1);
        ns[i] = (uint32_t)j + 1;
    }
    free(t->slot);
    t->slot = ns;
    t->scap = ncap;
    return 1;
}

/* insert or overwrite ("last line wins", hashmap.c:208-214); returns index or -1 */
static long strtab_set(strtab* t, const uint8_t* k, size_t n, int32_t v) {
    long j = strtab_find(t, k, n);
    if (j >= 0) {
        t->val[j] = v;
        return j;
    }
    if (t->n == t->cap) {
        size_t nc = t->cap ? t->cap * 2 : 1024;
        uint8_t** a = realloc(t->str, nc * sizeof *a);
        if (!a) return -1;
        t->str = a;
        uint32_t* b = realloc(t->len, nc * sizeof *b);
        if (!b) return -1;
        t->len = b;
        int32_t* c = realloc(t->val, nc * sizeof *c);
        if (!c) return -1;
        t->val = c;
        t->cap = nc;
    }
    if ((t->n + 1) * 2 > t->scap && !strtab_rehash(t, t->scap * 2)) return -1;
    uint8_t* copy = malloc(n ? n : 1);
    if (!copy) return -1;
    memcpy(copy, k, n);
    j = (long)t->n++;
    t->str[j] = copy;
    t->len[j] = (uint32_t)n;
    t->val[j] = v;
    size_t mask = t->scap - 1, i = fnv(k, n) & mask;
    while (t->slot[i]) i = (i + 1) & mask;
    t->slot[i] = (uint32_t)j + 1;
    return j;
}

/* ------------------------------------------------------------------------------------------------ */
static int fail(char* msg, size_t msglen, int cls, const char* text) {
    if (msg && msglen) snprintf(msg, msglen, "%s", text);
    return cls;
}

/* helper.c:82-128.  Each "0x" followed by two characters yields strtol(those two, 16) as one byte; other
 * characters are skipped.  Returns bytes produced or -1 when more than size-1 bytes would be produced. */
static long hex_to_bytes(const char* hex, uint8_t* out, size_t size) {
    size_t n = 0;
    while (*hex) {
        if (hex[0] == '0' && hex[1] == 'x') {
            hex += 2;
            if (hex[0] == '\0' || hex[1] == '\0') break; /* reference walks past the NUL here (undefined) */
            char pair[3] = {hex[0], hex[1], '\0'};
            unsigned v = (unsigned)strtol(pair, NULL, 16);
            if (n >= size - 1) return -1;
            out[n++] = (uint8_t)v;
            hex += 2;
        } else {
            hex++;
        }
    }
    return (long)n;
}

/* lib.c:243-378 */
static int load_vocab(const char* path, strtab* tab, uint32_t* lines, char* msg, size_t msglen) {
    FILE* f = fopen(path, "r");
    if (!f) return fail(msg, msglen, HUT_ERR_FILE_NOT_FOUND, "Could not open vocab file.");
    size_t cap = 8192, len;
    char* line = malloc(cap);
    char chunk[1024];
    int rc = HUT_OK;
    *lines = 0;
    if (!line) {
        fclose(f);
        return fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
    }
    for (;;) {
        len = 0;
        int eof = 0;
        for (;;) { /* a line is assembled from 1 KiB chunks; an unterminated last line is dropped */
            if (!fgets(chunk, sizeof chunk, f)) {
                eof = 1;
                break;
            }
            size_t cl = strlen(chunk);
            if (len + cl + 1 > cap) {
                cap = (len + cl + 1) * 2;
                char* nl = realloc(line, cap);
                if (!nl) {
                    rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
                    goto done;
                }
                line = nl;
            }
            memcpy(line + len, chunk, cl + 1);
            len += cl;
            if (strchr(chunk, '\n')) break;
        }
        if (eof || len == 0) break;
        char* sep = strstr(line, " == ");
        if (!sep) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Invalid format in vocab file.");
            break;
        }
        *sep = '\0';
        const char* vs = sep + 4;
        char* endp = NULL;
        errno = 0;
        long value = strtol(vs, &endp, 10);
        if (endp == vs) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Invalid vocab format: could not parse integer value.");
            break;
        }
        if (errno == ERANGE || value > INT_MAX || value < INT_MIN) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Integer value in vocab file is out of range.");
            break;
        }
        uint8_t key[2048];
        long kn = hex_to_bytes(line, key, sizeof key);
        size_t klen = 0;
        if (kn > 0) { /* strdup (lib.c:358) cuts the key at its first 0x00 */
            const uint8_t* z = memchr(key, 0, (size_t)kn);
            klen = z ? (size_t)(z - key) : (size_t)kn;
        }
        if (klen == 0) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Failed to convert hex string to ASCII.");
            break;
        }
        if (value < 0) {
            /* the reference indexes its decode arrays with the id (lib.c:427): a negative id writes out of
             * bounds there, and -1 doubles as "no such pair" (core.c:100).  Refuse instead. */
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Integer value in vocab file is out of range.");
            break;
        }
        if (strtab_set(tab, key, klen, (int32_t)value) < 0) {
            rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
            break;
        }
        (*lines)++;
    }
done:
    free(line);
    fclose(f);
    if (rc == HUT_OK && *lines == 0) rc = fail(msg, msglen, HUT_ERR_VALUE, "Vocab file is empty.");
    return rc;
}

/* lib.c:460-566 */
typedef struct {
    uint8_t bytes[32];
    uint32_t len;
    int byte;
} special_line;

static int load_special(const char* path, uint8_t* special[256], uint32_t special_len[256], special_line** lines,
                        size_t* n_lines, char* msg, size_t msglen) {
    size_t cap_lines = 0;
    *lines = NULL;
    *n_lines = 0;
    FILE* f = fopen(path, "r");
    if (!f) return fail(msg, msglen, HUT_ERR_FILE_NOT_FOUND, "Could not open special characters file.");
    char buf[32];
    int rc = HUT_OK;
    while (fgets(buf, sizeof buf, f)) {
        char* sep = strstr(buf, " == ");
        if (!sep) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Invalid format in special character file.");
            break;
        }
        char* endp = NULL;
        errno = 0;
        long index = strtol(buf, &endp, 10);
        if (endp == buf) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Invalid vocab format: could not parse integer value.");
            break;
        }
        /* the reference also lets 256 through (lib.c:516), which overruns its 256-entry arrays */
        if (errno == ERANGE || index > 255 || index < 0) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Integer value in vocab file is out of range.");
            break;
        }
        const char* vs = sep + 4;
        size_t vl = strlen(vs);
        size_t keep = vl ? vl - 1 : 0; /* last character always dropped as the presumed '\n' (lib.c:528-531) */
        if (keep == 0) {
            rc = fail(msg, msglen, HUT_ERR_VALUE, "Failed to convert hex string to ASCII.");
            break;
        }
        free(special[index]);
        special[index] = malloc(keep);
        if (!special[index]) {
            rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
            break;
        }
        memcpy(special[index], vs, keep);
        special_len[index] = (uint32_t)keep;
        if (*n_lines == cap_lines) { /* every line goes into the decode trie, in file order (lib.c:555-556) */
            cap_lines = cap_lines ? cap_lines * 2 : 128;
            special_line* nl = realloc(*lines, cap_lines * sizeof **lines);
            if (!nl) {
                rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
                break;
            }
            *lines = nl;
        }
        special_line* L = &(*lines)[(*n_lines)++];
        /* the reference stores the value as a C string: it ends at an embedded NUL */
        L->len = (uint32_t)strnlen(vs, keep);
        memcpy(L->bytes, vs, L->len);
        L->byte = (int)index;
    }
    fclose(f);
    return rc;
}

/* ------------------------------------------------------------------------------------------------
 * initial-symbol segmentation of an encoded string (core.c:35-55, pretokenizer.c:14-28)
 * ---------------------------------------------------------------------------------------------- */
static int lead_len(uint8_t c) {
    if ((c & 0x80) == 0x00) return 1;
    if ((c & 0xE0) == 0xC0) return 2;
    if ((c & 0xF0) == 0xE0) return 3;
    if ((c & 0xF8) == 0xF0) return 4;
    return 1;
}

static int is_hex(uint8_t c) {
    return (c >= '0' && c <= '9') || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F');
}

static size_t hex_symbol_len(const uint8_t* p, size_t n) {
    if (n >= 4 && p[0] == '<' && p[1] == '0' && (p[2] == 'x' || p[2] == 'X')) {
        size_t k = 3;
        while (k < n && is_hex(p[k])) k++;
        if (k < n && p[k] == '>') return k + 1;
    }
    return 0;
}

/* Splits e[0..n) into symbols; returns count, or -1 if the last symbol's lead byte claims more bytes than
 * remain (the segmentation would then depend on what follows). */
static long segment(const uint8_t* e, size_t n, int use_hex, uint32_t* starts, uint32_t* lens, size_t cap) {
    size_t i = 0, cnt = 0;
    while (i < n) {
        size_t l = use_hex ? hex_symbol_len(e + i, n - i) : 0;
        if (l == 0) l = (size_t)lead_len(e[i]);
        if (l > n - i) return -1;
        if (cnt < cap) {
            starts[cnt] = (uint32_t)i;
            lens[cnt] = (uint32_t)l;
        }
        cnt++;
        i += l;
    }
    return (long)cnt;
}

/* ------------------------------------------------------------------------------------------------
 * "Can the <0x..> rule (core.c:35-47) ever glue symbols coming from DIFFERENT input characters?"
 * If it could, initial symbols would not be a per-character function and the per-byte symbol tables
 * would be wrong, so such special-chars tables are refused at init.  Conservative reachability search
 * over (hex-pattern state, set of pretoken classes still possible).
 * ---------------------------------------------------------------------------------------------- */
enum { HX_LT = 1, HX_ZERO = 2, HX_DIGITS = 3 }; /* saw "<", "<0", "<0x" hex*  */

/* classes a byte may belong to inside a pretoken run (parser.c:24-88): bit0 ALPHA, bit1 DIGIT, bit2 OTHER */
static unsigned byte_run_classes(unsigned b) {
    if ((b >= 'a' && b <= 'z') || (b >= 'A' && b <= 'Z')) return 1;
    if (b >= '0' && b <= '9') return 2;
    if (b == ' ' || (b >= 9 && b <= 13) || b == 0) return 0;
    if (b < 0x80) return 4;
    return 1 | 4; /* bytes of multi-byte chars: Hungarian letters (ALPHA) or OTHER */
}

/* advance the hex DFA over s[0..n); returns new state, 0 = dead, -1 = pattern completed */
static int hex_advance(int st, const uint8_t* s, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        uint8_t c = s[i];
        switch (st) {
            case HX_LT: st = (c == '0') ? HX_ZERO : 0; break;
            case HX_ZERO: st = (c == 'x' || c == 'X') ? HX_DIGITS : 0; break;
            case HX_DIGITS:
                if (c == '>') return -1;
                st = is_hex(c) ? HX_DIGITS : 0;
                break;
            default: st = 0;
        }
        if (st == 0) return 0;
    }
    return st;
}

/* state in which an encoded string leaves the DFA open at its end (0 if it does not) */
static int open_hex_tail(const uint8_t* e, size_t n) {
    uint32_t st[64], ln[64];
    long c = segment(e, n, 1, st, ln, 64);
    if (c <= 0 || c > 64) return 0;
    /* an open pattern can only begin at a symbol boundary with '<' */
    for (long k = c - 1; k >= 0; --k) {
        if (e[st[k]] == '<' && ln[k] == 1) {
            int s = hex_advance(HX_LT, e + st[k] + 1, n - st[k] - 1);
            if (s > 0) return s;
        }
    }
    return 0;
}

typedef struct {
    const uint8_t* enc[256];
    uint32_t len[256];
    int present[256]; /* this byte can start an input character whose encoding is enc[b] */
} enc_table;

static int hex_can_span(const enc_table* t) {
    /* visited[state][classmask] */
    unsigned char seen[4][8];
    memset(seen, 0, sizeof seen);
    int stack[64][2], sp = 0;
    for (int b = 1; b < 256; ++b) {
        if (!t->present[b]) continue;
        int s = open_hex_tail(t->enc[b], t->len[b]);
        if (s <= 0) continue;
        unsigned cm = (b == ' ') ? 7u : byte_run_classes((unsigned)b);
        if (cm == 0) continue; /* whitespace other than ' ' is a pretoken by itself */
        if (!seen[s][cm]) {
            seen[s][cm] = 1;
            stack[sp][0] = s;
            stack[sp][1] = (int)cm;
            sp++;
        }
    }
    while (sp) {
        sp--;
        int s = stack[sp][0];
        unsigned cm = (unsigned)stack[sp][1];
        for (int b = 1; b < 256; ++b) {
            if (!t->present[b]) continue;
            unsigned c2 = byte_run_classes((unsigned)b) & cm;
            if (c2 == 0) continue;
            int ns = hex_advance(s, t->enc[b], t->len[b]);
            if (ns < 0) return 1;
            if (ns > 0 && !seen[ns][c2] && sp < 64) {
                seen[ns][c2] = 1;
                stack[sp][0] = ns;
                stack[sp][1] = (int)c2;
                sp++;
            }
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * builder
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    hut_model* m;
    strtab units;      /* unit string -> unit id (vocab first, virtual appended) */
    strtab symbols;    /* explicit initial symbols (replacement / expansion / prefix pieces) */
    uint64_t* pairs;   /* packed entries, grown */
    size_t n_pairs, cap_pairs;
} builder;

static int is_initial_symbol(const builder* b, const uint8_t* s, size_t n) {
    if (strtab_find(&b->symbols, s, n) >= 0) return 1;
    /* plain mode: any raw character (one lead-byte-length group) is a symbol */
    if (!b->m->is_byte_encoder && n >= 1 && (size_t)lead_len(s[0]) == n) return 1;
    return 0;
}

/* unit id of a possible token, creating a virtual unit for non-vocab initial symbols; -1 if impossible */
static long unit_of(builder* b, const uint8_t* s, size_t n, int create) {
    long u = strtab_find(&b->units, s, n);
    if (u >= 0) return u;
    if (!is_initial_symbol(b, s, n)) return -1;
    if (!create) return -1;
    return strtab_set(&b->units, s, n, -1);
}

static int add_pair(builder* b, uint32_t l, uint32_t r, uint32_t mrg) {
    if (b->n_pairs == b->cap_pairs) {
        size_t nc = b->cap_pairs ? b->cap_pairs * 2 : 65536;
        uint64_t* p = realloc(b->pairs, nc * sizeof *p);
        if (!p) return 0;
        b->pairs = p;
        b->cap_pairs = nc;
    }
    b->pairs[b->n_pairs++] = ((uint64_t)l << 42) | ((uint64_t)r << 21) | mrg;
    return 1;
}

uint32_t hut_model_pair_lookup(const hut_model* m, uint32_t l, uint32_t r) {
    uint64_t key = ((uint64_t)l << 21) | r;
    uint32_t mask = m->pair_cap - 1;
    uint32_t i = hut_pair_hash(l, r, mask);
    for (;;) {
        uint64_t s = m->pair_tab[i];
        if (s == HUT_PAIR_EMPTY) return HUT_UNIT_MASK;
        if ((s >> 21) == key) return (uint32_t)(s & HUT_UNIT_MASK);
        i = (i + 1) & mask;
    }
}

uint32_t hut_model_char_lookup(const hut_model* m, const uint8_t* p, uint32_t len) {
    if (len == 1 && p[0] < 0x80) return m->ascii_unit[p[0]];
    uint32_t key = 0;
    for (uint32_t i = 0; i < len && i < 4; ++i) key |= (uint32_t)p[i] << (8 * i);
    if (!m->char_cap) return m->inert;
    uint32_t mask = m->char_cap - 1;
    uint32_t i = (uint32_t)(hut_mix64(key) >> 32) & mask;
    for (;;) {
        uint64_t s = m->char_tab[i];
        if (s == HUT_CHAR_EMPTY) return m->inert;
        if ((uint32_t)(s >> 32) == key) return (uint32_t)s;
        i = (i + 1) & mask;
    }
}

static int cmp_vocab_order(const void* a, const void* b, void* ctx) {
    const strtab* t = ctx;
    uint32_t i = *(const uint32_t*)a, j = *(const uint32_t*)b;
    if (t->val[i] != t->val[j]) return t->val[i] < t->val[j] ? -1 : 1;
    uint32_t n = t->len[i] < t->len[j] ? t->len[i] : t->len[j];
    int c = memcmp(t->str[i], t->str[j], n);
    if (c) return c;
    return (t->len[i] > t->len[j]) - (t->len[i] < t->len[j]);
}

/* greedy merge of a short unit sequence with the finished tables (used once, for the constant
 * "prefix alone" ids of core.c:421-446); same order as core.c:66-209: min (rank, left position). */
static uint32_t merge_small(const hut_model* m, uint32_t* u, uint32_t n) {
    for (;;) {
        int64_t best = -1;
        uint32_t bi = 0, bm = 0;
        for (uint32_t i = 0; i + 1 < n; ++i) {
            uint32_t mg = hut_model_pair_lookup(m, u[i], u[i + 1]);
            if (mg == HUT_UNIT_MASK) continue;
            int64_t rk = m->rank_of[mg];
            if (best < 0 || rk < best) {
                best = rk;
                bi = i;
                bm = mg;
            }
        }
        if (best < 0) return n;
        u[bi] = bm;
        memmove(u + bi + 1, u + bi + 2, (n - bi - 2) * sizeof *u);
        n--;
    }
}

/* ------------------------------------------------------------------------------------------------
 * decode tables.  vocab_decode[value] = key (lib.c:423-433): when several keys carry one id the reference keeps
 * whichever its hash map iterates last; here the key stated last in the file wins (ids are unique in every real
 * vocabulary).  Ids >= vocab_lines are never decodable (core.c:522).
 * Trie = ac_automaton_add_string per special-chars line, in file order (ac.c:27-50): a repeated pattern keeps the
 * later byte.  The failure links the reference also builds are not used by its decoder (pretokenizer.c:224-240).
 * ---------------------------------------------------------------------------------------------- */
uint32_t hut_model_trie_child(const hut_model* m, uint32_t node, uint8_t byte) {
    if (!m->trie_cap) return 0;
    const uint32_t key = (node << 8) | byte, mask = m->trie_cap - 1;
    uint32_t i = hut_trie_hash(key, mask);
    for (;;) {
        uint64_t s = m->trie_edge[i];
        if (s == HUT_TRIE_EMPTY) return 0;
        if ((uint32_t)(s >> 32) == key) return (uint32_t)s;
        i = (i + 1) & mask;
    }
}

static int build_decode_tables(hut_model* m, const strtab* vocab, const special_line* lines, size_t n_lines,
                               const char* prefix, size_t plen) {
    const uint32_t V = m->vocab_lines;
    m->dec_off = calloc((size_t)V + 1, sizeof(uint32_t));
    m->dec_len = malloc(((size_t)V ? V : 1) * sizeof(uint32_t));
    uint32_t* who = malloc(((size_t)V ? V : 1) * sizeof(uint32_t));
    if (!m->dec_off || !m->dec_len || !who) {
        free(who);
        return 0;
    }
    memset(who, 0xFF, ((size_t)V ? V : 1) * sizeof(uint32_t));
    for (size_t i = 0; i < vocab->n; ++i)
        if (vocab->val[i] >= 0 && (uint32_t)vocab->val[i] < V) who[vocab->val[i]] = (uint32_t)i;
    size_t total = 0;
    for (uint32_t id = 0; id < V; ++id) total += who[id] != 0xFFFFFFFFu ? vocab->len[who[id]] : 0;
    m->dec_pool = malloc(total + 16);
    if (!m->dec_pool || total > 0xFFFFFFF0u) {
        free(who);
        return 0;
    }
    size_t at = 0;
    for (uint32_t id = 0; id < V; ++id) {
        m->dec_off[id] = (uint32_t)at;
        if (who[id] == 0xFFFFFFFFu) {
            m->dec_len[id] = HUT_DEC_MISSING;
            continue;
        }
        m->dec_len[id] = vocab->len[who[id]];
        memcpy(m->dec_pool + at, vocab->str[who[id]], vocab->len[who[id]]);
        at += vocab->len[who[id]];
    }
    m->dec_off[V] = (uint32_t)at;
    m->dec_pool_len = (uint32_t)at;
    memset(m->dec_pool + at, 0, 16);
    free(who);

    size_t max_nodes = 1;
    for (size_t k = 0; k < n_lines; ++k) max_nodes += lines[k].len;
    m->trie_out = malloc(max_nodes * sizeof(int16_t));
    m->trie_depth = malloc(max_nodes);
    m->trie_cap = 64;
    while (m->trie_cap < max_nodes * 2 + 8) m->trie_cap <<= 1;
    m->trie_edge = malloc((size_t)m->trie_cap * sizeof(uint64_t));
    if (!m->trie_out || !m->trie_depth || !m->trie_edge) return 0;
    memset(m->trie_edge, 0xFF, (size_t)m->trie_cap * sizeof(uint64_t));
    m->trie_out[0] = -1;
    m->trie_depth[0] = 0;
    m->trie_nodes = 1;
    for (size_t k = 0; k < n_lines; ++k) {
        uint32_t node = 0;
        for (uint32_t i = 0; i < lines[k].len; ++i) {
            uint32_t child = hut_model_trie_child(m, node, lines[k].bytes[i]);
            if (!child) {
                child = m->trie_nodes++;
                m->trie_out[child] = -1;
                m->trie_depth[child] = (uint8_t)(i + 1);
                const uint32_t key = (node << 8) | lines[k].bytes[i], mask = m->trie_cap - 1;
                uint32_t h = hut_trie_hash(key, mask);
                while (m->trie_edge[h] != HUT_TRIE_EMPTY) h = (h + 1) & mask;
                m->trie_edge[h] = ((uint64_t)key << 32) | child;
                if (node == 0) m->trie_root_child[lines[k].bytes[i]] = child;
            }
            node = child;
        }
        if (node) m->trie_out[node] = (int16_t)lines[k].byte;
    }
    if (plen) {
        m->prefix_bytes = malloc(plen);
        if (!m->prefix_bytes) return 0;
        memcpy(m->prefix_bytes, prefix, plen);
        m->prefix_len = (uint32_t)plen;
    }
    return 1;
}

/* ------------------------------------------------------------------------------------------------
 * merges file (lib.c:573-663): lines "left right"; '#' lines skipped; a rule is accepted when left, right and
 * left+right are all vocab keys; rank = index among the accepted rules; a later rule with the same
 * (left_id, right_id) replaces the earlier one (hashmap_set, hashmap.c:207-214) but keeps its own rank.
 * *active = 0 when the file has no candidate line at all: the reference then leaves merges_map NULL and
 * encodes through the string path (lib.c:593,653-655).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int32_t l, r, m;
    uint32_t rank;
} merge_rule;

#define HUT_MERGES_LINE 10000 /* MAX_LINE_LENGTH, lib.c:71: longer lines are seen in pieces, as by the reference */

static int load_merges(const char* path, const strtab* vocab, merge_rule** out, size_t* n_out, int* active,
                       char* msg, size_t msglen) {
    *out = NULL;
    *n_out = 0;
    *active = 0;
    FILE* f = fopen(path, "r");
    if (!f) return fail(msg, msglen, HUT_ERR_FILE_NOT_FOUND, "Could not open merges file.");
    char* line = malloc(HUT_MERGES_LINE);
    char* cat = malloc(2 * HUT_MERGES_LINE);
    if (!line || !cat) {
        free(line);
        free(cat);
        fclose(f);
        return fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
    }
    size_t line_count = 0;
    while (fgets(line, HUT_MERGES_LINE, f))
        if (line[0] != '#' && strchr(line, ' ') != NULL) line_count++;
    int rc = HUT_OK;
    if (line_count > 0) {
        *active = 1;
        merge_rule* rules = malloc(line_count * sizeof *rules);
        size_t cap = 64;
        while (cap < line_count * 2 + 8) cap <<= 1;
        uint32_t* slot = malloc(cap * sizeof *slot); /* (l, r) -> index into rules, for "last wins" */
        if (!rules || !slot) {
            free(rules);
            free(slot);
            rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
            goto done;
        }
        memset(slot, 0xFF, cap * sizeof *slot);
        size_t n = 0;
        uint32_t rank = 0;
        fseek(f, 0L, SEEK_SET);
        while (fgets(line, HUT_MERGES_LINE, f) && rank < line_count) {
            if (line[0] == '#') continue;
            line[strcspn(line, "\r\n")] = 0;
            char* save = NULL;
            char* ls = strtok_r(line, " ", &save);
            char* rs = strtok_r(NULL, " ", &save);
            if (!ls || !rs) continue;
            size_t ll = strlen(ls), rl = strlen(rs);
            memcpy(cat, ls, ll);
            memcpy(cat + ll, rs, rl);
            long li = strtab_find(vocab, (const uint8_t*)ls, ll);
            long ri = strtab_find(vocab, (const uint8_t*)rs, rl);
            long mi = strtab_find(vocab, (const uint8_t*)cat, ll + rl);
            if (li < 0 || ri < 0 || mi < 0) continue;
            merge_rule R = {vocab->val[li], vocab->val[ri], vocab->val[mi], rank++};
            uint64_t key = ((uint64_t)(uint32_t)R.l << 32) | (uint32_t)R.r;
            size_t i = (size_t)hut_mix64(key) & (cap - 1);
            while (slot[i] != 0xFFFFFFFFu && !(rules[slot[i]].l == R.l && rules[slot[i]].r == R.r)) i = (i + 1) & (cap - 1);
            if (slot[i] == 0xFFFFFFFFu) {
                slot[i] = (uint32_t)n;
                rules[n++] = R;
            } else {
                rules[slot[i]] = R;
            }
        }
        free(slot);
        *out = rules;
        *n_out = n;
    }
done:
    free(line);
    free(cat);
    fclose(f);
    return rc;
}

static int cmp_rule_rank(const void* a, const void* b) {
    const merge_rule *x = a, *y = b;
    return (x->rank > y->rank) - (x->rank < y->rank);
}

static int compute_monotone(hut_model* m);

/* open-addressing table from the builder's packed (L,R,M) entries (rank_of is final when this is called) */
static int build_pair_tab(hut_model* m, const builder* b) {
    free(m->pair_tab);
    m->pair_count = 0;
    m->pair_cap = 1024;
    while (m->pair_cap < b->n_pairs * 4 + 16) m->pair_cap <<= 1; /* load <= 0.25: ~1.2 probes per lookup */
    m->pair_tab = malloc((size_t)m->pair_cap * sizeof(uint64_t));
    if (!m->pair_tab) return 0;
    memset(m->pair_tab, 0xFF, (size_t)m->pair_cap * sizeof(uint64_t));
    for (size_t k = 0; k < b->n_pairs; ++k) {
        uint64_t e = b->pairs[k], key = e >> 21;
        uint32_t mask = m->pair_cap - 1;
        uint32_t i = hut_pair_hash((uint32_t)(key >> 21), (uint32_t)key & HUT_UNIT_MASK, mask);
        int dup = 0;
        while (m->pair_tab[i] != HUT_PAIR_EMPTY) {
            if ((m->pair_tab[i] >> 21) == key) { /* same (L,R) => same concatenation => same M */
                dup = 1;
                break;
            }
            i = (i + 1) & mask;
        }
        if (!dup) {
            m->pair_tab[i] = e;
            m->pair_count++;
        }
    }
    return compute_monotone(m);
}

/* monotone ranks (hut_model.h): rank(M) > rank(L), rank(R) for every pair (L, R) -> M whose L / R is itself produced
 * by some pair */
static int compute_monotone(hut_model* m) {
    uint8_t* produced = calloc((size_t)m->n_units + 2, 1);
    if (!produced) return 0;
    for (uint32_t i = 0; i < m->pair_cap; ++i)
        if (m->pair_tab[i] != HUT_PAIR_EMPTY) produced[m->pair_tab[i] & HUT_UNIT_MASK] = 1;
    int mono = 1;
    for (uint32_t i = 0; i < m->pair_cap && mono; ++i) {
        const uint64_t e = m->pair_tab[i];
        if (e == HUT_PAIR_EMPTY) continue;
        const uint32_t l = (uint32_t)(e >> 42) & HUT_UNIT_MASK, r = (uint32_t)(e >> 21) & HUT_UNIT_MASK,
                       mg = (uint32_t)e & HUT_UNIT_MASK;
        const int32_t rk = m->rank_of[mg];
        if ((produced[l] && m->rank_of[l] >= rk) || (produced[r] && m->rank_of[r] >= rk)) mono = 0;
    }
    free(produced);
    m->monotone = mono;
    return 1;
}

/* ------------------------------------------------------------------------------------------------
 * id-keyed merge path (core.c:211-337, 453-474; rules from load_merges).  The kernels stay as they are: they
 * merge the adjacent pair with the smallest unit index M found in the pair table and print out_id[unit].  Here
 *   - symbols are ids, not strings: one "base" unit per id that a single UTF-8 character of the vocab maps to
 *     (core.c:460-474; an unknown character is the inert unit: -1 never matches a rule), then
 *   - one unit per rule, in rank order (unit index order == rank order, so distinct_values = 1), printing the
 *     rule's merge_id;
 *   - every unit that prints id x stands for x in later lookups, so rule (a, b) gets one pair-table entry per
 *     (unit printing a, unit printing b).  With one rule per merged id (every real merges file) that is one
 *     entry per rule.
 * The string-mode tables built before are replaced, except the constant "prefix alone" ids, which the
 * reference always computes through the string path (core.c:421-446).
 * ---------------------------------------------------------------------------------------------- */
static int apply_merges(hut_model* m, builder* b, const strtab* vocab, const enc_table* et, uint8_t* const special[256],
                        const char* prefix, size_t plen, merge_rule* rules, size_t n_rules, char* msg, size_t msglen) {
    int rc = HUT_OK;
    const size_t V = (size_t)m->max_value + 1;
    int32_t* base_of = malloc(V * sizeof *base_of);
    uint32_t *pstart = NULL, *plist = NULL;
    if (!base_of) return fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
    memset(base_of, 0xFF, V * sizeof *base_of);
    for (size_t i = 0; i < vocab->n; ++i)
        if (vocab->len[i] == (uint32_t)lead_len(vocab->str[i][0])) base_of[vocab->val[i]] = 0;
    uint32_t n_base = 0;
    for (size_t id = 0; id < V; ++id)
        if (base_of[id] == 0) base_of[id] = (int32_t)n_base++;
    qsort(rules, n_rules, sizeof *rules, cmp_rule_rank);
    const size_t n_units = (size_t)n_base + n_rules;
    if (n_units > HUT_MAX_UNITS) {
        rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED, "vocabulary too large for 21-bit unit ids");
        goto done;
    }
    /* units no longer have strings */
    for (size_t i = 0; i < b->units.n; ++i) free(b->units.str[i]);
    free(b->units.str);
    free(b->units.len);
    b->units.str = calloc(n_units ? n_units : 1, sizeof *b->units.str);
    b->units.len = calloc(n_units ? n_units : 1, sizeof *b->units.len);
    b->units.n = n_units;
    m->unit_str = b->units.str;
    m->unit_len = b->units.len;
    free(m->out_id);
    free(m->rank_of);
    m->out_id = malloc((n_units + 1) * sizeof(int32_t));
    m->rank_of = malloc((n_units + 1) * sizeof(int32_t));
    if (!b->units.str || !b->units.len || !m->out_id || !m->rank_of) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto done;
    }
    for (size_t id = 0; id < V; ++id)
        if (base_of[id] >= 0) m->out_id[base_of[id]] = (int32_t)id;
    for (size_t k = 0; k < n_rules; ++k) m->out_id[n_base + k] = rules[k].m;
    m->out_id[n_units] = -1;
    for (size_t u = 0; u < n_units; ++u) m->rank_of[u] = (int32_t)u;
    m->rank_of[n_units] = INT_MAX;
    m->n_vocab = m->n_units = m->inert = (uint32_t)n_units;
    m->distinct_values = 1;
    m->identity_ids = 0;
    m->max_unit_len = 0;
    m->merges_mode = 1;
    m->n_merge_rules = (uint32_t)n_rules;

    /* units printing each id (CSR) */
    pstart = calloc(V + 1, sizeof *pstart);
    plist = malloc((n_units ? n_units : 1) * sizeof *plist);
    if (!pstart || !plist) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto done;
    }
    for (size_t u = 0; u < n_units; ++u) pstart[m->out_id[u] + 1]++;
    for (size_t id = 0; id < V; ++id) pstart[id + 1] += pstart[id];
    {
        uint32_t* fill = calloc(V, sizeof *fill);
        if (!fill) {
            rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
            goto done;
        }
        for (size_t u = 0; u < n_units; ++u) {
            const int32_t id = m->out_id[u];
            plist[pstart[id] + fill[id]++] = (uint32_t)u;
        }
        free(fill);
    }
    b->n_pairs = 0;
    for (size_t k = 0; k < n_rules; ++k) {
        const int32_t a = rules[k].l, c = rules[k].r;
        for (uint32_t i = pstart[a]; i < pstart[a + 1]; ++i)
            for (uint32_t j = pstart[c]; j < pstart[c + 1]; ++j) {
                if (b->n_pairs >= ((size_t)1 << 24)) {
                    rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED, "merges file expands to more than 2^24 unit pairs");
                    goto done;
                }
                if (!add_pair(b, plist[i], plist[j], (uint32_t)(n_base + k))) {
                    rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
                    goto done;
                }
            }
    }
    if (!build_pair_tab(m, b)) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto done;
    }

#define CHAR_UNIT(P, N, OUT)                                                        \
    do {                                                                            \
        long j_ = strtab_find(vocab, (P), (N));                                     \
        int32_t bu_ = j_ >= 0 ? base_of[vocab->val[j_]] : -1;                       \
        (OUT) = bu_ >= 0 ? (uint32_t)bu_ : m->inert;                                \
    } while (0)

    /* ---- initial symbols: one per UTF-8 character of the encoded form (core.c:460-474) ---- */
    uint32_t st[64], ln[64];
    {
        size_t pool_cap = 256 * 4, pool_n = 0;
        free(m->sym_pool);
        m->sym_pool = malloc(pool_cap * sizeof(uint32_t));
        m->max_sym_per_char = 1;
        for (int c = 0; c < 256; ++c) {
            m->sym_off[c] = (uint32_t)pool_n;
            if (c == 0 || !et->present[c]) continue;
            if (!m->is_byte_encoder && !special[c]) continue;
            long cnt = segment(et->enc[c], et->len[c], 0, st, ln, 64);
            if (cnt <= 0 || cnt > 64) {
                rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED,
                          "special-chars replacement does not split into whole characters on its own");
                goto done;
            }
            if ((uint32_t)cnt > m->max_sym_per_char) m->max_sym_per_char = (uint32_t)cnt;
            if (pool_n + (size_t)cnt > pool_cap) {
                pool_cap = (pool_n + (size_t)cnt) * 2;
                m->sym_pool = realloc(m->sym_pool, pool_cap * sizeof(uint32_t));
            }
            for (long k = 0; k < cnt; ++k) CHAR_UNIT(et->enc[c] + st[k], ln[k], m->sym_pool[pool_n++]);
        }
        m->sym_off[256] = (uint32_t)pool_n;
    }
    m->ascii_unit[0] = m->inert;
    for (int c = 1; c < 128; ++c) {
        uint8_t ch = (uint8_t)c;
        CHAR_UNIT(&ch, 1, m->ascii_unit[c]);
    }
    free(m->char_tab);
    m->char_tab = NULL;
    m->char_cap = 0;
    if (!m->is_byte_encoder) {
        uint32_t n_chars = 0;
        for (size_t i = 0; i < vocab->n; ++i) {
            uint32_t l = vocab->len[i];
            if (l >= 2 && l <= 4 && (uint32_t)lead_len(vocab->str[i][0]) == l) n_chars++;
        }
        m->char_cap = 64;
        while (m->char_cap < n_chars * 2 + 8) m->char_cap <<= 1;
        m->char_tab = malloc((size_t)m->char_cap * sizeof(uint64_t));
        if (!m->char_tab) {
            rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
            goto done;
        }
        memset(m->char_tab, 0xFF, (size_t)m->char_cap * sizeof(uint64_t));
        for (size_t i = 0; i < vocab->n; ++i) {
            uint32_t l = vocab->len[i];
            if (!(l >= 2 && l <= 4 && (uint32_t)lead_len(vocab->str[i][0]) == l)) continue;
            uint32_t key = 0, u;
            for (uint32_t q = 0; q < l; ++q) key |= (uint32_t)vocab->str[i][q] << (8 * q);
            CHAR_UNIT(vocab->str[i], l, u);
            uint32_t mask = m->char_cap - 1, h = (uint32_t)(hut_mix64(key) >> 32) & mask;
            while (m->char_tab[h] != HUT_CHAR_EMPTY) h = (h + 1) & mask;
            m->char_tab[h] = ((uint64_t)key << 32) | u;
        }
    }
    if (m->has_prefix) { /* the raw prefix is prepended to the first pretoken and split with it (core.c:448-474) */
        uint32_t pst[256], pln[256];
        long n_pfx = segment((const uint8_t*)prefix, plen, 0, pst, pln, 256);
        if (n_pfx <= 0 || n_pfx > 256) {
            rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED, "prefix does not split into whole characters on its own");
            goto done;
        }
        free(m->prefix_syms);
        m->n_prefix_syms = (uint32_t)n_pfx;
        m->prefix_syms = malloc((size_t)n_pfx * sizeof(uint32_t));
        for (long k = 0; k < n_pfx; ++k) CHAR_UNIT((const uint8_t*)prefix + pst[k], pln[k], m->prefix_syms[k]);
    }
#undef CHAR_UNIT
done:
    free(base_of);
    free(pstart);
    free(plist);
    return rc;
}

void hut_model_free(hut_model* m) {
    if (!m) return;
    if (m->unit_str)
        for (uint32_t i = 0; i < m->n_units; ++i) free(m->unit_str[i]);
    free(m->unit_str);
    free(m->unit_len);
    free(m->out_id);
    free(m->rank_of);
    free(m->pair_tab);
    free(m->sym_pool);
    free(m->char_tab);
    free(m->prefix_syms);
    free(m->prefix_alone_ids);
    free(m->dec_off);
    free(m->dec_len);
    free(m->dec_pool);
    free(m->trie_edge);
    free(m->trie_out);
    free(m->trie_depth);
    free(m->prefix_bytes);
    free(m);
}

int hut_model_load(hut_model** out, const char* vocab_path, const char* special_path, const char* prefix,
                   int is_byte_encoder, const char* merges_path, char* msg, size_t msglen) {
    int rc = HUT_OK;
    hut_model* m = calloc(1, sizeof *m);
    uint8_t* special[256] = {0};
    uint32_t special_len[256] = {0};
    strtab vocab;
    builder b;
    memset(&b, 0, sizeof b);
    uint32_t* order = NULL;
    uint8_t(*expansion)[2] = NULL;
    special_line* sp_lines = NULL;
    size_t n_sp_lines = 0;
    if (!m || !strtab_init(&vocab)) {
        free(m);
        return fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
    }
    b.m = m;
    m->is_byte_encoder = is_byte_encoder != 0;
    if (!strtab_init(&b.units) || !strtab_init(&b.symbols)) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto out;
    }
    if ((rc = load_vocab(vocab_path, &vocab, &m->vocab_lines, msg, msglen)) != HUT_OK) goto out;
    if ((rc = load_special(special_path, special, special_len, &sp_lines, &n_sp_lines, msg, msglen)) != HUT_OK) goto out;
    if (vocab.n > HUT_MAX_UNITS / 2) {
        rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED, "vocabulary too large for 21-bit unit ids");
        goto out;
    }

    /* ---- units for vocab strings, ordered by (value, bytes) ---- */
    order = malloc(vocab.n * sizeof *order);
    if (!order) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto out;
    }
    for (size_t i = 0; i < vocab.n; ++i) order[i] = (uint32_t)i;
    qsort_r(order, vocab.n, sizeof *order, cmp_vocab_order, &vocab);
    m->distinct_values = 1;
    m->identity_ids = 1;
    for (size_t k = 0; k < vocab.n; ++k) {
        uint32_t j = order[k];
        if (strtab_set(&b.units, vocab.str[j], vocab.len[j], vocab.val[j]) != (long)k) {
            rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
            goto out;
        }
        if (k && vocab.val[order[k - 1]] == vocab.val[j]) m->distinct_values = 0;
        if (vocab.val[j] != (int32_t)k) m->identity_ids = 0;
        if (vocab.val[j] > m->max_value) m->max_value = vocab.val[j];
    }
    m->n_vocab = (uint32_t)vocab.n;

    /* ---- encoded form of each input byte / lead byte, explicit symbol set ---- */
    enc_table et;
    memset(&et, 0, sizeof et);
    expansion = calloc(256, sizeof *expansion);
    uint8_t ident[256];
    for (int c = 0; c < 256; ++c) ident[c] = (uint8_t)c;
    for (int c = 1; c < 256; ++c) {
        if (special[c]) {
            et.enc[c] = special[c];
            et.len[c] = special_len[c];
            et.present[c] = 1;
        } else if (m->is_byte_encoder) {
            if (c >= 0x80) { /* pretokenizer.c:138-146 */
                expansion[c][0] = (uint8_t)(0xC0 | (c >> 6));
                expansion[c][1] = (uint8_t)(0x80 | (c & 0x3F));
                et.enc[c] = expansion[c];
                et.len[c] = 2;
            } else {
                et.enc[c] = &ident[c];
                et.len[c] = 1;
            }
            et.present[c] = 1;
        } else if (c < 0x80) { /* plain mode raw ASCII char; raw multi-byte chars cannot carry hex-pattern bytes */
            et.enc[c] = &ident[c];
            et.len[c] = 1;
            et.present[c] = 1;
        }
    }
    uint32_t st[64], ln[64];
    m->max_sym_per_char = 1;
    for (int c = 1; c < 256; ++c) {
        if (!et.present[c]) continue;
        if (!m->is_byte_encoder && !special[c]) continue; /* raw chars are implicit symbols */
        long cnt = segment(et.enc[c], et.len[c], 1, st, ln, 64);
        if (cnt <= 0 || cnt > 64) {
            rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED,
                      "special-chars replacement does not split into whole symbols on its own");
            goto out;
        }
        if ((uint32_t)cnt > m->max_sym_per_char) m->max_sym_per_char = (uint32_t)cnt;
        for (long k = 0; k < cnt; ++k)
            if (strtab_set(&b.symbols, et.enc[c] + st[k], ln[k], 0) < 0) {
                rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
                goto out;
            }
    }
    if (hex_can_span(&et)) {
        rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED,
                  "special-chars table lets a <0x..> symbol span several input characters");
        goto out;
    }

    /* ---- prefix pieces ---- */
    uint8_t* pfx_alone = NULL; /* byte-encoded prefix (core.c:423-426) */
    size_t pfx_alone_len = 0;
    uint32_t pst[256], pln[256], ast[1024], aln[1024];
    long n_pfx = 0, n_alone = 0;
    size_t plen = prefix ? strlen(prefix) : 0;
    if (plen > 200) {
        rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED, "prefix longer than 200 bytes");
        goto out;
    }
    if (plen) { /* an empty prefix behaves as no prefix (the reference's own handling is undefined) */
        m->has_prefix = 1;
        n_pfx = segment((const uint8_t*)prefix, plen, 1, pst, pln, 256);
        if (n_pfx <= 0 || open_hex_tail((const uint8_t*)prefix, plen)) {
            rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED, "prefix does not split into whole symbols on its own");
            goto out;
        }
        for (long k = 0; k < n_pfx; ++k) strtab_set(&b.symbols, (const uint8_t*)prefix + pst[k], pln[k], 0);
        /* prefix alone: byte-encode it like a word (pretokenizer.c:102-168), split by lead length only */
        pfx_alone = malloc(plen * 32 + 8);
        const uint8_t* p = (const uint8_t*)prefix;
        size_t i = 0;
        while (i < plen) {
            uint8_t c = p[i];
            size_t cl = m->is_byte_encoder ? 1 : (size_t)lead_len(c);
            if (cl > plen - i) cl = plen - i;
            if (special[c]) {
                memcpy(pfx_alone + pfx_alone_len, special[c], special_len[c]);
                pfx_alone_len += special_len[c];
            } else if (m->is_byte_encoder && c >= 0x80) {
                pfx_alone[pfx_alone_len++] = (uint8_t)(0xC0 | (c >> 6));
                pfx_alone[pfx_alone_len++] = (uint8_t)(0x80 | (c & 0x3F));
            } else {
                memcpy(pfx_alone + pfx_alone_len, p + i, cl);
                pfx_alone_len += cl;
            }
            i += cl;
        }
        {
            size_t q = 0;
            while (q < pfx_alone_len && n_alone < 1024) { /* core.c:432-437: no <0x..> rule here */
                size_t l = (size_t)lead_len(pfx_alone[q]);
                if (l > pfx_alone_len - q) l = pfx_alone_len - q;
                ast[n_alone] = (uint32_t)q;
                aln[n_alone] = (uint32_t)l;
                n_alone++;
                q += l;
            }
        }
        for (long k = 0; k < n_alone; ++k) strtab_set(&b.symbols, pfx_alone + ast[k], aln[k], 0);
    }

    /* ---- pair table: every split of every vocab string into two possible tokens ---- */
    for (uint32_t u = 0; u < m->n_vocab; ++u) {
        const uint8_t* s = b.units.str[u];
        uint32_t n = b.units.len[u];
        for (uint32_t k = 1; k < n; ++k) {
            long l = unit_of(&b, s, k, 0);
            if (l < 0 && !is_initial_symbol(&b, s, k)) continue;
            long r = unit_of(&b, s + k, n - k, 0);
            if (r < 0 && !is_initial_symbol(&b, s + k, n - k)) continue;
            if (l < 0) l = unit_of(&b, s, k, 1);
            if (r < 0) r = unit_of(&b, s + k, n - k, 1);
            if (l < 0 || r < 0) {
                rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
                goto out;
            }
            /* unit_of() may have reallocated the unit arrays */
            s = b.units.str[u];
            if (!add_pair(&b, (uint32_t)l, (uint32_t)r, u)) {
                rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
                goto out;
            }
        }
    }
    if (b.units.n > HUT_MAX_UNITS) {
        rc = fail(msg, msglen, HUT_ERR_UNSUPPORTED, "vocabulary too large for 21-bit unit ids");
        goto out;
    }
    m->n_units = (uint32_t)b.units.n;
    m->inert = m->n_units;

    m->unit_str = b.units.str;
    m->unit_len = b.units.len;
    m->out_id = malloc((m->n_units + 1) * sizeof(int32_t));
    m->rank_of = malloc((m->n_units + 1) * sizeof(int32_t));
    if (!m->out_id || !m->rank_of) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto out;
    }
    /* rank_of is a DENSE rank: vocab units are ordered by (value, bytes), so the index of the first unit carrying a
     * value orders the units exactly as the values do (ties stay ties), and it fits the 21-bit unit range -- the merge
     * loops pack it with a position into one 32-bit key.  Only the order of ranks is ever used (core.c:111-176). */
    for (uint32_t u = 0; u <= m->n_units; ++u) {
        int is_vocab = u < m->n_vocab;
        m->out_id[u] = is_vocab ? b.units.val[u] : -1;
        m->rank_of[u] = !is_vocab ? INT_MAX : (u && b.units.val[u] == b.units.val[u - 1]) ? m->rank_of[u - 1] : (int32_t)u;
        if (u < m->n_units && m->unit_len[u] > m->max_unit_len) m->max_unit_len = m->unit_len[u];
    }

    if (!build_pair_tab(m, &b)) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto out;
    }

    /* ---- per-byte symbol pools, raw-char tables ---- */
    {
        size_t pool_cap = 256 * 4, pool_n = 0;
        m->sym_pool = malloc(pool_cap * sizeof(uint32_t));
        for (int c = 0; c < 256; ++c) {
            m->sym_off[c] = (uint32_t)pool_n;
            if (c == 0 || !et.present[c]) continue;
            if (!m->is_byte_encoder && !special[c]) continue;
            long cnt = segment(et.enc[c], et.len[c], 1, st, ln, 64);
            if (pool_n + (size_t)cnt > pool_cap) {
                pool_cap = (pool_n + (size_t)cnt) * 2;
                m->sym_pool = realloc(m->sym_pool, pool_cap * sizeof(uint32_t));
            }
            for (long k = 0; k < cnt; ++k) {
                long u = strtab_find(&b.units, et.enc[c] + st[k], ln[k]);
                m->sym_pool[pool_n++] = u >= 0 ? (uint32_t)u : m->inert;
            }
        }
        m->sym_off[256] = (uint32_t)pool_n;
    }
    for (int c = 0; c < 128; ++c) {
        uint8_t ch = (uint8_t)c;
        long u = c ? strtab_find(&b.units, &ch, 1) : -1;
        m->ascii_unit[c] = u >= 0 ? (uint32_t)u : m->inert;
    }
    if (!m->is_byte_encoder) {
        uint32_t n_chars = 0;
        for (uint32_t u = 0; u < m->n_units; ++u) {
            uint32_t l = m->unit_len[u];
            if (l >= 2 && l <= 4 && (uint32_t)lead_len(m->unit_str[u][0]) == l) n_chars++;
        }
        m->char_cap = 64;
        while (m->char_cap < n_chars * 2 + 8) m->char_cap <<= 1;
        m->char_tab = malloc((size_t)m->char_cap * sizeof(uint64_t));
        memset(m->char_tab, 0xFF, (size_t)m->char_cap * sizeof(uint64_t));
        for (uint32_t u = 0; u < m->n_units; ++u) {
            uint32_t l = m->unit_len[u];
            if (!(l >= 2 && l <= 4 && (uint32_t)lead_len(m->unit_str[u][0]) == l)) continue;
            uint32_t key = 0;
            for (uint32_t i = 0; i < l; ++i) key |= (uint32_t)m->unit_str[u][i] << (8 * i);
            uint32_t mask = m->char_cap - 1, i = (uint32_t)(hut_mix64(key) >> 32) & mask;
            while (m->char_tab[i] != HUT_CHAR_EMPTY) i = (i + 1) & mask;
            m->char_tab[i] = ((uint64_t)key << 32) | u;
        }
    }

    /* ---- prefix symbol units and the constant "prefix alone" ids ---- */
    if (m->has_prefix) {
        m->n_prefix_syms = (uint32_t)n_pfx;
        m->prefix_syms = malloc((size_t)n_pfx * sizeof(uint32_t));
        for (long k = 0; k < n_pfx; ++k) {
            long u = strtab_find(&b.units, (const uint8_t*)prefix + pst[k], pln[k]);
            m->prefix_syms[k] = u >= 0 ? (uint32_t)u : m->inert;
        }
        uint32_t* tmp = malloc((size_t)(n_alone ? n_alone : 1) * sizeof(uint32_t));
        for (long k = 0; k < n_alone; ++k) {
            long u = strtab_find(&b.units, pfx_alone + ast[k], aln[k]);
            tmp[k] = u >= 0 ? (uint32_t)u : m->inert;
        }
        uint32_t left = merge_small(m, tmp, (uint32_t)n_alone);
        m->n_prefix_alone = left;
        m->prefix_alone_ids = malloc((size_t)(left ? left : 1) * sizeof(int32_t));
        for (uint32_t k = 0; k < left; ++k) m->prefix_alone_ids[k] = m->out_id[tmp[k]];
        free(tmp);
    }
    free(pfx_alone);

    /* ---- decode tables ---- */
    if (!build_decode_tables(m, &vocab, sp_lines, n_sp_lines, prefix, plen)) {
        rc = fail(msg, msglen, HUT_ERR_MEMORY, "out of memory");
        goto out;
    }

    /* ---- optional merges file: switch the tables to the id-keyed path ---- */
    if (merges_path) {
        merge_rule* rules = NULL;
        size_t n_rules = 0;
        int active = 0;
        rc = load_merges(merges_path, &vocab, &rules, &n_rules, &active, msg, msglen);
        if (rc == HUT_OK && active)
            rc = apply_merges(m, &b, &vocab, &et, special, prefix, plen, rules, n_rules, msg, msglen);
        free(rules);
        if (rc != HUT_OK) goto out;
    }

    /* ownership of the unit arrays moved into the model */
    b.units.str = NULL;
    b.units.len = NULL;

out:
    free(sp_lines);
    for (int c = 0; c < 256; ++c) free(special[c]);
    free(expansion);
    free(order);
    free(b.pairs);
    if (b.units.str) strtab_free(&b.units, 1);
    else {
        free(b.units.val);
        free(b.units.slot);
    }
    strtab_free(&b.symbols, 1);
    strtab_free(&vocab, 1);
    if (rc != HUT_OK) {
        m->unit_str = NULL; /* freed with the builder above */
        m->unit_len = NULL;
        hut_model_free(m);
        return rc;
    }
    *out = m;
    return HUT_OK;
}
