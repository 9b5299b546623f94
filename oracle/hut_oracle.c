/* TEST INFRASTRUCTURE ONLY -- see hut_oracle.h.
 *
 * Plain-C restatement of the reference's encode path.  It is written from the behaviour described in
 * SURVEY.md Appendix A, not translated from the reference sources: different containers (FNV open-addressing
 * map, explicit binary heap on malloc'd storage, no arena, no VLAs), same observable results inside the
 * reference's defined domain.  Reference citations (matyasosvath/hutoken @ 2d23def) are given per function.
 *
 * Pinning: tests/test_oracle_golden.py runs this file against the UNMODIFIED reference compiled into
 * oracle/_ref on randomised inputs, and tests/golden/ holds vectors produced by that reference.
 */
#define _GNU_SOURCE
#include "hut_oracle.h"

#include <errno.h>
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------
 * string -> id map.  Contents follow hashmap_set semantics (hashmap.c:164-229): inserting an existing
 * key overwrites the value ("last line wins", hashmap.c:208-214).  Layout is ours.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    uint8_t* key; /* malloc'd, not NUL-terminated */
    uint32_t len;
    int32_t value;
    uint32_t seq; /* order of first insertion (decode: among keys sharing an id the one stated last wins) */
} omap_slot;

typedef struct {
    omap_slot* slots;
    size_t cap; /* power of two */
    size_t count;
    uint32_t max_key_len;
} omap;

static uint64_t fnv1a(const uint8_t* p, size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) {
        h ^= p[i];
        h *= 1099511628211ull;
    }
    h ^= h >> 32;
    return h;
}

static int omap_init(omap* m, size_t cap) {
    m->cap = 256;
    while (m->cap < cap) m->cap <<= 1;
    m->slots = calloc(m->cap, sizeof(omap_slot));
    m->count = 0;
    m->max_key_len = 0;
    return m->slots != NULL;
}

static void omap_free(omap* m) {
    if (!m->slots) return;
    for (size_t i = 0; i < m->cap; ++i) free(m->slots[i].key);
    free(m->slots);
    m->slots = NULL;
}

static omap_slot* omap_find(const omap* m, const uint8_t* key, size_t len) {
    size_t mask = m->cap - 1;
    size_t i = fnv1a(key, len) & mask;
    for (;;) {
        omap_slot* s = &m->slots[i];
        if (!s->key) return s;
        if (s->len == len && memcmp(s->key, key, len) == 0) return s;
        i = (i + 1) & mask;
    }
}

static int omap_set(omap* m, const uint8_t* key, size_t len, int32_t value);

static int omap_grow(omap* m) {
    omap old = *m;
    if (!omap_init(m, old.cap * 2)) return 0;
    m->max_key_len = old.max_key_len;
    for (size_t i = 0; i < old.cap; ++i) {
        if (old.slots[i].key) {
            omap_slot* s = omap_find(m, old.slots[i].key, old.slots[i].len);
            *s = old.slots[i];
            m->count++;
        }
    }
    free(old.slots);
    return 1;
}

static int omap_set(omap* m, const uint8_t* key, size_t len, int32_t value) {
    if ((m->count + 1) * 2 > m->cap && !omap_grow(m)) return 0;
    omap_slot* s = omap_find(m, key, len);
    if (s->key) {
        s->value = value; /* duplicate key: last wins */
        return 1;
    }
    s->key = malloc(len ? len : 1);
    if (!s->key) return 0;
    memcpy(s->key, key, len);
    s->len = (uint32_t)len;
    s->value = value;
    s->seq = (uint32_t)m->count;
    m->count++;
    if (len > m->max_key_len) m->max_key_len = (uint32_t)len;
    return 1;
}

static int32_t omap_get(const omap* m, const uint8_t* key, size_t len) {
    if (len == 0 || len > m->max_key_len) return -1;
    const omap_slot* s = omap_find(m, key, len);
    return s->key ? s->value : -1;
}

/* struct MergeRule (bpe.h:11-16) in an open-addressing table keyed by (left_id, right_id) */
struct orule {
    int32_t left, right, merge, rank;
    int used;
};

/* struct ACNode (ac.h): children by byte, output_value = the byte the pattern stands for or -1 */
struct otrie {
    int32_t child[256];
    int out;
    uint32_t depth;
};

/* ------------------------------------------------------------------------------------------------ */
struct hut_oracle {
    omap vocab;
    size_t vocab_lines;       /* vocab_size_decode: counts every accepted line (lib.c:378) */
    uint8_t* special[256];    /* EncodeContext.special_chars (taskqueue.h:22) */
    uint32_t special_len[256];
    uint8_t* prefix;          /* NULL if none */
    uint32_t prefix_len;
    int is_byte_encoder;
    int limit_mode;
    /* decode side: vocab_decode[] (lib.c:423-433) and the special-chars trie (ac.c:27-50, lib.c:555-556) */
    const omap_slot** dec; /* [vocab_lines], NULL = no line carries the id */
    struct otrie* trie;
    size_t trie_n, trie_cap;
    /* merges-file mode (lib.c:573-663): (left_id, right_id) -> rule; NULL map = string path */
    struct orule* rules;
    size_t rules_cap; /* power of two; 0 = merges_map is NULL */
    size_t n_rules;
};

static void set_err(char* err, size_t errlen, const char* type, const char* msg) {
    if (err && errlen) snprintf(err, errlen, "%s: %s", type, msg);
}

/* hex_str_to_ascii (helper.c:82-128): every "0x" followed by two characters yields one byte
 * (strtol base 16 of those two characters); anything else is skipped one character at a time.
 * Returns the number of bytes produced, or -1 on overflow of the 2048-byte buffer (helper.c:103-113).
 * Where the reference would step past the terminating NUL (an incomplete trailing pair) we stop. */
static long hex_to_bytes(const char* hex, uint8_t* out, size_t out_size) {
    size_t n = 0;
    while (*hex) {
        if (hex[0] == '0' && hex[1] == 'x') {
            hex += 2;
            if (hex[0] != '\0' && hex[1] != '\0') {
                char pair[3] = {hex[0], hex[1], '\0'};
                unsigned int v = (unsigned int)strtol(pair, NULL, 16);
                if (n >= out_size - 1) return -1;
                out[n++] = (uint8_t)v;
                hex += 2;
            } else {
                break;
            }
        } else {
            hex++;
        }
    }
    return (long)n;
}

/* Vocab loader, lib.c:243-378.  Lines are assembled from 1 KiB fgets chunks; a trailing line without
 * '\n' is dropped (the EOF branch at lib.c:264-267 discards the partial line); split at the first
 * " == " (lib.c:295); id by strtol (lib.c:315); key is cut at its first 0x00 byte by strdup (lib.c:358)
 * and may not be empty (lib.c:345-355). */
static int load_vocab(hut_oracle* o, const char* path, char* err, size_t errlen) {
    FILE* f = fopen(path, "r");
    if (!f) {
        set_err(err, errlen, "FileNotFoundError", "Could not open vocab file.");
        return 0;
    }
    size_t cap = 4096, len = 0;
    char* line = malloc(cap);
    char chunk[1024];
    int ok = 1;
    for (;;) {
        len = 0;
        int at_eof = 0;
        for (;;) {
            if (!fgets(chunk, sizeof chunk, f)) {
                at_eof = 1;
                break;
            }
            size_t cl = strlen(chunk);
            if (len + cl + 1 > cap) {
                cap = (len + cl + 1) * 2;
                line = realloc(line, cap);
            }
            memcpy(line + len, chunk, cl);
            len += cl;
            line[len] = '\0';
            if (strchr(chunk, '\n')) break;
        }
        if (at_eof) break;
        if (len == 0) break;
        char* sep = strstr(line, " == ");
        if (!sep) {
            set_err(err, errlen, "ValueError", "Invalid format in vocab file.");
            ok = 0;
            break;
        }
        *sep = '\0';
        const char* vs = sep + 4;
        char* endp = NULL;
        errno = 0;
        long value = strtol(vs, &endp, 10);
        if (endp == vs) {
            set_err(err, errlen, "ValueError", "Invalid vocab format: could not parse integer value.");
            ok = 0;
            break;
        }
        if (errno == ERANGE || value > INT_MAX || value < INT_MIN) {
            set_err(err, errlen, "ValueError", "Integer value in vocab file is out of range.");
            ok = 0;
            break;
        }
        uint8_t key[2048];
        long kn = hex_to_bytes(line, key, sizeof key);
        size_t klen = 0;
        if (kn > 0) {
            const uint8_t* z = memchr(key, 0, (size_t)kn);
            klen = z ? (size_t)(z - key) : (size_t)kn;
        }
        if (klen == 0) {
            set_err(err, errlen, "ValueError", "Failed to convert hex string to ASCII.");
            ok = 0;
            break;
        }
        if (!omap_set(&o->vocab, key, klen, (int32_t)value)) {
            set_err(err, errlen, "MemoryError", "out of memory");
            ok = 0;
            break;
        }
        o->vocab_lines++;
    }
    free(line);
    fclose(f);
    if (ok && o->vocab_lines == 0) {
        set_err(err, errlen, "ValueError", "Vocab file is empty.");
        ok = 0;
    }
    return ok;
}

/* Special-chars loader, lib.c:460-566: 32-byte fgets buffer (lib.c:483), "<decimal> == <value>", index
 * 0..256 accepted by the reference (lib.c:516; 256 overflows its arrays -- we reject it), the LAST
 * character of the value is always dropped as the presumed newline (lib.c:528-531), an empty value
 * is an error (lib.c:533-543). */
/* ac_automaton_add_string (ac.c:27-50): a repeated pattern keeps the later output value */
static int32_t trie_new(hut_oracle* o, uint32_t depth) {
    if (o->trie_n == o->trie_cap) {
        o->trie_cap = o->trie_cap ? o->trie_cap * 2 : 64;
        o->trie = realloc(o->trie, o->trie_cap * sizeof *o->trie);
    }
    struct otrie* t = &o->trie[o->trie_n];
    memset(t->child, 0, sizeof t->child);
    t->out = -1;
    t->depth = depth;
    return (int32_t)o->trie_n++;
}

static void trie_add(hut_oracle* o, const uint8_t* pat, size_t len, int value) {
    if (!o->trie_n) trie_new(o, 0);
    int32_t cur = 0;
    for (size_t i = 0; i < len; ++i) {
        if (!o->trie[cur].child[pat[i]]) {
            int32_t c = trie_new(o, (uint32_t)i + 1);
            o->trie[cur].child[pat[i]] = c;
        }
        cur = o->trie[cur].child[pat[i]];
    }
    if (cur) {
        o->trie[cur].out = value;
        o->trie[cur].depth = (uint32_t)len;
    }
}

static int load_special(hut_oracle* o, const char* path, char* err, size_t errlen) {
    FILE* f = fopen(path, "r");
    if (!f) {
        set_err(err, errlen, "FileNotFoundError", "Could not open special characters file.");
        return 0;
    }
    char buf[32];
    int ok = 1;
    while (fgets(buf, sizeof buf, f)) {
        char* sep = strstr(buf, " == ");
        if (!sep) {
            set_err(err, errlen, "ValueError", "Invalid format in special character file.");
            ok = 0;
            break;
        }
        char* endp = NULL;
        errno = 0;
        long index = strtol(buf, &endp, 10);
        if (endp == buf) {
            set_err(err, errlen, "ValueError", "Invalid vocab format: could not parse integer value.");
            ok = 0;
            break;
        }
        if (errno == ERANGE || index > 255 || index < 0) {
            set_err(err, errlen, "ValueError", "Integer value in vocab file is out of range.");
            ok = 0;
            break;
        }
        const char* vs = sep + 4;
        size_t vl = strlen(vs);
        /* value = vs without its last character; vl == 0 underflows in the reference, treat as empty */
        size_t keep = vl ? vl - 1 : 0;
        if (keep == 0) {
            set_err(err, errlen, "ValueError", "Failed to convert hex string to ASCII.");
            ok = 0;
            break;
        }
        free(o->special[index]);
        o->special[index] = malloc(keep);
        memcpy(o->special[index], vs, keep);
        o->special_len[index] = (uint32_t)keep;
        trie_add(o, (const uint8_t*)vs, strnlen(vs, keep), (int)index); /* lib.c:555: the value as a C string */
    }
    fclose(f);
    return ok;
}

hut_oracle* hut_oracle_load(const char* vocab_path, const char* special_path, const char* prefix,
                            int is_byte_encoder, char* err, size_t errlen) {
    hut_oracle* o = calloc(1, sizeof *o);
    if (!o) return NULL;
    if (!omap_init(&o->vocab, 1024)) {
        free(o);
        return NULL;
    }
    o->is_byte_encoder = is_byte_encoder != 0;
    if (prefix && prefix[0]) { /* empty prefix is treated as no prefix (reference: UB at core.c:428-440) */
        o->prefix_len = (uint32_t)strlen(prefix);
        o->prefix = malloc(o->prefix_len);
        memcpy(o->prefix, prefix, o->prefix_len);
    }
    if (!load_vocab(o, vocab_path, err, errlen) || !load_special(o, special_path, err, errlen)) {
        hut_oracle_free(o);
        return NULL;
    }
    return o;
}

/* ------------------------------------------------------------------------------------------------
 * Merges file (lib.c:573-663).  Candidate lines = not starting with '#' and containing a space
 * (lib.c:587-591); none => merges_map stays NULL (string path).  Accepted rule = left, right and
 * left+right are vocab keys (lib.c:620-640); rank counts accepted rules (lib.c:643); hashmap_set
 * replaces an earlier rule with the same (left_id, right_id) (hashmap.c:207-214).
 * ---------------------------------------------------------------------------------------------- */
static uint64_t rule_hash(int32_t l, int32_t r) { /* pair_hash, bpe.c:32-41 (any hash would do) */
    uint64_t x = ((uint64_t)(uint32_t)l << 32) | (uint32_t)r;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}

static const struct orule* rule_find(const hut_oracle* o, int32_t l, int32_t r) {
    if (!o->rules_cap) return NULL;
    size_t i = (size_t)rule_hash(l, r) & (o->rules_cap - 1);
    while (o->rules[i].used) {
        if (o->rules[i].left == l && o->rules[i].right == r) return &o->rules[i];
        i = (i + 1) & (o->rules_cap - 1);
    }
    return NULL;
}

int hut_oracle_load_merges(hut_oracle* o, const char* path, char* err, size_t errlen) {
    FILE* f = fopen(path, "r");
    if (!f) {
        set_err(err, errlen, "FileNotFoundError", "Could not open merges file.");
        return 0;
    }
    enum { LINE = 10000 }; /* MAX_LINE_LENGTH, lib.c:71 */
    char* line = malloc(LINE);
    size_t line_count = 0;
    while (fgets(line, LINE, f))
        if (line[0] != '#' && strchr(line, ' ') != NULL) line_count++;
    if (line_count > 0) {
        o->rules_cap = 64;
        while (o->rules_cap < 2 * line_count + 8) o->rules_cap <<= 1;
        o->rules = calloc(o->rules_cap, sizeof *o->rules);
        fseek(f, 0L, SEEK_SET);
        size_t accepted = 0;
        int32_t rank = 0;
        uint8_t* cat = malloc(2 * LINE);
        while (fgets(line, LINE, f) && accepted < line_count) {
            if (line[0] == '#') continue;
            line[strcspn(line, "\r\n")] = 0;
            char* save = NULL;
            char* ls = strtok_r(line, " ", &save);
            char* rs = strtok_r(NULL, " ", &save);
            if (!ls || !rs) continue;
            size_t ll = strlen(ls), rl = strlen(rs);
            memcpy(cat, ls, ll);
            memcpy(cat + ll, rs, rl);
            const omap_slot* a = omap_find(&o->vocab, (const uint8_t*)ls, ll);
            const omap_slot* b = omap_find(&o->vocab, (const uint8_t*)rs, rl);
            const omap_slot* c = omap_find(&o->vocab, cat, ll + rl);
            if (!a->key || !b->key || !c->key) continue;
            size_t i = (size_t)rule_hash(a->value, b->value) & (o->rules_cap - 1);
            while (o->rules[i].used && !(o->rules[i].left == a->value && o->rules[i].right == b->value))
                i = (i + 1) & (o->rules_cap - 1);
            if (!o->rules[i].used) o->n_rules++;
            o->rules[i] = (struct orule){a->value, b->value, c->value, rank++, 1};
            accepted++;
        }
        free(cat);
    }
    free(line);
    fclose(f);
    return 1;
}

size_t hut_oracle_merge_rules(const hut_oracle* o) { return o->n_rules; }
int hut_oracle_merges_active(const hut_oracle* o) { return o->rules_cap != 0; }

void hut_oracle_free(hut_oracle* o) {
    if (!o) return;
    omap_free(&o->vocab);
    for (int i = 0; i < 256; ++i) free(o->special[i]);
    free(o->prefix);
    free(o->rules);
    free(o->dec);
    free(o->trie);
    free(o);
}

void hut_oracle_set_limit_mode(hut_oracle* o, int limit_mode) { o->limit_mode = limit_mode; }

int32_t hut_oracle_vocab_lookup(const hut_oracle* o, const uint8_t* key, size_t len) {
    return omap_get(&o->vocab, key, len);
}
size_t hut_oracle_vocab_lines(const hut_oracle* o) { return o->vocab_lines; }

/* ------------------------------------------------------------------------------------------------
 * Pretokeniser (parser.c:24-183).
 * ---------------------------------------------------------------------------------------------- */
enum { CL_END = 0, CL_ALPHA, CL_DIGIT, CL_OTHER, CL_NONE };

/* decode_utf8 (parser.c:144-183): structural check of continuation bits only; bytes past the end of the
 * text read as NUL.  Returns the char length, 0 if invalid / end; *cp receives the code point. */
static int dec_utf8(const uint8_t* t, size_t len, size_t pos, uint32_t* cp) {
#define B(i) ((pos + (i)) < len ? t[pos + (i)] : 0)
    uint8_t b0 = B(0);
    if (b0 == 0) return 0;
    if (b0 < 0x80) {
        *cp = b0;
        return 1;
    }
    if ((b0 & 0xE0) == 0xC0) {
        if ((B(1) & 0xC0) != 0x80) return 0;
        *cp = ((uint32_t)(b0 & 0x1F) << 6) | (B(1) & 0x3F);
        return 2;
    }
    if ((b0 & 0xF0) == 0xE0) {
        if ((B(1) & 0xC0) != 0x80 || (B(2) & 0xC0) != 0x80) return 0;
        *cp = ((uint32_t)(b0 & 0x0F) << 12) | ((uint32_t)(B(1) & 0x3F) << 6) | (B(2) & 0x3F);
        return 3;
    }
    if ((b0 & 0xF8) == 0xF0) {
        if ((B(1) & 0xC0) != 0x80 || (B(2) & 0xC0) != 0x80 || (B(3) & 0xC0) != 0x80) return 0;
        *cp = ((uint32_t)(b0 & 0x07) << 18) | ((uint32_t)(B(1) & 0x3F) << 12) |
              ((uint32_t)(B(2) & 0x3F) << 6) | (B(3) & 0x3F);
        return 4;
    }
    return 0;
#undef B
}

/* is_custom_alpha / is_digit / is_whitespace / is_other (parser.c:102-142).  isspace() is frozen to the
 * C / C.UTF-8 locale set {0x09..0x0D, 0x20}. */
static int cp_class(uint32_t cp) {
    if (cp == 0) return CL_NONE;
    if ((cp >= 'a' && cp <= 'z') || (cp >= 'A' && cp <= 'Z')) return CL_ALPHA;
    switch (cp) {
        case 0xE1: case 0xE9: case 0xED: case 0xF3: case 0xFA: case 0x151: case 0x171: case 0xFC: case 0xF6:
        case 0xC1: case 0xC9: case 0xCD: case 0xD3: case 0xDA: case 0x150: case 0x170: case 0xDC: case 0xD6:
            return CL_ALPHA;
        default: break;
    }
    if (cp >= '0' && cp <= '9') return CL_DIGIT;
    if (cp == 0x20 || (cp >= 0x09 && cp <= 0x0D)) return CL_NONE; /* whitespace: no run class */
    return CL_OTHER;
}

/* consume_while (parser.c:90-100) */
static size_t run_of(const uint8_t* t, size_t len, size_t pos, int cls) {
    while (pos < len && t[pos] != 0) {
        uint32_t cp = 0;
        int l = dec_utf8(t, len, pos, &cp);
        if (l == 0 || cp == 0 || cp_class(cp) != cls) break;
        pos += (size_t)l;
    }
    return pos;
}

/* parser_next_token (parser.c:24-88) */
size_t hut_oracle_next_pretoken(const uint8_t* t, size_t len, size_t pos) {
    if (pos >= len || t[pos] == 0) return 0;
    size_t s = pos + (t[pos] == ' ' ? 1 : 0);
    for (int cls = CL_ALPHA; cls <= CL_OTHER; ++cls) {
        size_t e = run_of(t, len, s, cls);
        if (e > s) return e - pos;
    }
    if (t[pos] == ' ') {
        size_t e = pos + 1;
        while (e < len && t[e] == ' ') e++;
        return e - pos;
    }
    return 1;
}

/* ------------------------------------------------------------------------------------------------
 * Byte / special-char encoding (pretokenizer.c:102-168) and initial symbols (core.c:35-55,483-488).
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    uint8_t* p;
    size_t n, cap;
} bytebuf;

static void bb_push(bytebuf* b, const uint8_t* src, size_t n) {
    if (b->n + n + 1 > b->cap) {
        b->cap = (b->n + n + 1) * 2 + 64;
        b->p = realloc(b->p, b->cap);
    }
    memcpy(b->p + b->n, src, n);
    b->n += n;
}

/* utf8_char_length (pretokenizer.c:14-28): by lead byte only */
static int lead_len(uint8_t c) {
    if ((c & 0x80) == 0x00) return 1;
    if ((c & 0xE0) == 0xC0) return 2;
    if ((c & 0xF0) == 0xE0) return 3;
    if ((c & 0xF8) == 0xF0) return 4;
    return 1;
}

/* pretokenizer_encode_arena (pretokenizer.c:102-168).  In plain mode a char whose lead-byte length runs
 * past the end of the word is clamped to the word (the reference reads past its NUL there -- undefined;
 * unreachable for valid UTF-8 input, which is all the Python boundary admits). */
static void byte_encode(const hut_oracle* o, const uint8_t* w, size_t wl, const uint8_t* prefix,
                        size_t prefix_len, bytebuf* out) {
    out->n = 0;
    if (prefix) bb_push(out, prefix, prefix_len);
    size_t i = 0;
    while (i < wl) {
        uint8_t b = w[i];
        size_t cl = o->is_byte_encoder ? 1 : (size_t)lead_len(b);
        if (cl > wl - i) cl = wl - i;
        if (o->special[b]) {
            bb_push(out, o->special[b], o->special_len[b]);
        } else if (o->is_byte_encoder && b >= 0x80) {
            uint8_t pair[2] = {(uint8_t)(0xC0 | (b >> 6)), (uint8_t)(0x80 | (b & 0x3F))};
            bb_push(out, pair, 2);
        } else {
            bb_push(out, w + i, cl);
        }
        i += cl;
    }
}

/* hex_token_length (core.c:35-47): "<0x" or "<0X", hex digits*, ">" */
static size_t hex_symbol_len(const uint8_t* p, size_t n) {
    if (n >= 4 && p[0] == '<' && p[1] == '0' && (p[2] == 'x' || p[2] == 'X')) {
        size_t k = 3;
        while (k < n && ((p[k] >= '0' && p[k] <= '9') || (p[k] >= 'a' && p[k] <= 'f') ||
                         (p[k] >= 'A' && p[k] <= 'F')))
            k++;
        if (k < n && p[k] == '>') return k + 1;
    }
    return 0;
}

typedef struct {
    uint32_t start, end; /* [start,end) in the encoded buffer */
} span;

/* next_token_length loop (core.c:483-488); with use_hex == 0 the prefix-only split of core.c:432-437 */
static size_t split_symbols(const uint8_t* e, size_t n, int use_hex, span* out) {
    size_t cnt = 0, i = 0;
    while (i < n) {
        size_t l = use_hex ? hex_symbol_len(e + i, n - i) : 0;
        if (l == 0) l = (size_t)lead_len(e[i]);
        if (l > n - i) l = n - i;
        out[cnt].start = (uint32_t)i;
        out[cnt].end = (uint32_t)(i + l);
        cnt++;
        i += l;
    }
    return cnt;
}

/* ------------------------------------------------------------------------------------------------
 * Merge loop (core.c:66-209) with the heap order of queue.c:152-199: min by (rank, left index).
 * Lazy deletion + re-validation exactly as core.c:119-137.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int32_t rank;
    uint32_t left, right;
} cand;

typedef struct {
    cand* d;
    size_t n, cap;
} heap;

static int cand_less(const cand* a, const cand* b) {
    return a->rank < b->rank || (a->rank == b->rank && a->left < b->left);
}

static void heap_push(heap* h, cand c) {
    if (h->n == h->cap) {
        h->cap = h->cap ? h->cap * 2 : 64;
        h->d = realloc(h->d, h->cap * sizeof(cand));
    }
    size_t i = h->n++;
    h->d[i] = c;
    while (i > 0) {
        size_t p = (i - 1) / 2;
        if (!cand_less(&h->d[i], &h->d[p])) break;
        cand t = h->d[i];
        h->d[i] = h->d[p];
        h->d[p] = t;
        i = p;
    }
}

static cand heap_pop(heap* h) {
    cand top = h->d[0];
    h->d[0] = h->d[--h->n];
    size_t i = 0;
    for (;;) {
        size_t l = 2 * i + 1, r = l + 1, s = i;
        if (l < h->n && cand_less(&h->d[l], &h->d[s])) s = l;
        if (r < h->n && cand_less(&h->d[r], &h->d[s])) s = r;
        if (s == i) break;
        cand t = h->d[i];
        h->d[i] = h->d[s];
        h->d[s] = t;
        i = s;
    }
    return top;
}

typedef struct {
    span* sp;
    int32_t* prev;
    int32_t* next;
    uint8_t* dead;
    size_t cap;
    heap h;
    bytebuf enc;
} scratch;

static void scratch_reserve(scratch* s, size_t n) {
    if (n <= s->cap) return;
    s->cap = n * 2 + 64;
    s->sp = realloc(s->sp, s->cap * sizeof(span));
    s->prev = realloc(s->prev, s->cap * sizeof(int32_t));
    s->next = realloc(s->next, s->cap * sizeof(int32_t));
    s->dead = realloc(s->dead, s->cap);
}

static void scratch_free(scratch* s) {
    free(s->sp);
    free(s->prev);
    free(s->next);
    free(s->dead);
    free(s->h.d);
    free(s->enc.p);
}

/* get_pair_rank_from_strings (core.c:700-722): rank = vocab id of the concatenated bytes, -1 if absent.
 * Adjacent live tokens are contiguous in the encoded buffer, so the concatenation is one range. */
static int32_t pair_rank(const hut_oracle* o, const uint8_t* e, const span* sp, int32_t l, int32_t r) {
    return omap_get(&o->vocab, e + sp[l].start, sp[r].end - sp[l].start);
}

static __thread hut_oracle_stats g_stats;

typedef struct {
    int32_t* d;
    size_t n, cap;
} idvec;

static void idvec_push(idvec* v, int32_t x) {
    if (v->n == v->cap) {
        v->cap = v->cap ? v->cap * 2 : 256;
        v->d = realloc(v->d, v->cap * sizeof(int32_t));
    }
    v->d[v->n++] = x;
}

/* bpe_encode_arena_string (core.c:66-209) over m initial symbols in s->sp; appends ids to out. */
static void merge_symbols(const hut_oracle* o, scratch* s, const uint8_t* e, size_t m, idvec* out) {
    if (m == 0) return;
    for (size_t i = 0; i < m; ++i) {
        s->prev[i] = (int32_t)i - 1;
        s->next[i] = (int32_t)i + 1;
        s->dead[i] = 0;
    }
    s->next[m - 1] = -1;
    s->h.n = 0;
    for (size_t i = 0; i + 1 < m; ++i) { /* core.c:97-109 */
        int32_t rk = pair_rank(o, e, s->sp, (int32_t)i, (int32_t)i + 1);
        if (rk != -1) heap_push(&s->h, (cand){rk, (uint32_t)i, (uint32_t)i + 1});
    }
    size_t live = m;
    while (s->h.n) { /* core.c:111-176 */
        cand c = heap_pop(&s->h);
        int32_t l = (int32_t)c.left, r = (int32_t)c.right;
        if (s->dead[l] || s->dead[r]) continue;
        if (s->next[l] != r) continue;
        if (pair_rank(o, e, s->sp, l, r) != c.rank) continue;
        s->sp[l].end = s->sp[r].end;
        s->dead[r] = 1;
        live--;
        int32_t p = s->prev[l], nx = s->next[r];
        s->next[l] = nx;
        if (nx != -1) s->prev[nx] = l;
        if (p != -1) {
            int32_t rk = pair_rank(o, e, s->sp, p, l);
            if (rk != -1) heap_push(&s->h, (cand){rk, (uint32_t)p, (uint32_t)l});
        }
        if (nx != -1) {
            int32_t rk = pair_rank(o, e, s->sp, l, nx);
            if (rk != -1) heap_push(&s->h, (cand){rk, (uint32_t)l, (uint32_t)nx});
        }
    }
    for (size_t i = 0; i < m; ++i) { /* core.c:178-208 */
        if (s->dead[i]) continue;
        idvec_push(out, omap_get(&o->vocab, e + s->sp[i].start, s->sp[i].end - s->sp[i].start));
    }
    g_stats.pretokens += 1;
    g_stats.symbols += m;
    g_stats.tokens += live;
    g_stats.probes += (m - 1) + 2 * (m - live) + live;
}

/* bpe_encode_arena_ids (core.c:211-337): same lazy-deletion loop keyed by (left_id, right_id) rules;
 * get_pair_rank_from_ids (core.c:724-736).  tok[] holds m ids (-1 = character not in the vocab, core.c:467-471). */
static void merge_ids(const hut_oracle* o, scratch* s, int32_t* tok, size_t m, idvec* out) {
    if (m == 0) return;
    for (size_t i = 0; i < m; ++i) {
        s->prev[i] = (int32_t)i - 1;
        s->next[i] = (int32_t)i + 1;
        s->dead[i] = 0;
    }
    s->next[m - 1] = -1;
    s->h.n = 0;
    for (size_t i = 0; i + 1 < m; ++i) { /* core.c:244-256 */
        const struct orule* r = rule_find(o, tok[i], tok[i + 1]);
        if (r) heap_push(&s->h, (cand){r->rank, (uint32_t)i, (uint32_t)i + 1});
    }
    size_t live = m;
    while (s->h.n) { /* core.c:258-325 */
        cand c = heap_pop(&s->h);
        int32_t l = (int32_t)c.left, r = (int32_t)c.right;
        if (s->dead[l] || s->dead[r]) continue;
        if (s->next[l] != r) continue;
        const struct orule* rule = rule_find(o, tok[l], tok[r]);
        if (!rule || rule->rank != c.rank) continue;
        tok[l] = rule->merge;
        s->dead[r] = 1;
        live--;
        int32_t p = s->prev[l], nx = s->next[r];
        s->next[l] = nx;
        if (nx != -1) s->prev[nx] = l;
        if (p != -1) {
            const struct orule* q = rule_find(o, tok[p], tok[l]);
            if (q) heap_push(&s->h, (cand){q->rank, (uint32_t)p, (uint32_t)l});
        }
        if (nx != -1) {
            const struct orule* q = rule_find(o, tok[l], tok[nx]);
            if (q) heap_push(&s->h, (cand){q->rank, (uint32_t)l, (uint32_t)nx});
        }
    }
    for (size_t i = 0; i < m; ++i) /* core.c:327-336 */
        if (!s->dead[i]) idvec_push(out, tok[i]);
    g_stats.pretokens += 1;
    g_stats.symbols += m;
    g_stats.tokens += live;
    g_stats.probes += (m - 1) + 2 * (m - live) + live;
}

/* encode() (core.c:339-511), default parser path (no regex pattern); string or merges-file ranks. */
static void encode_doc(const hut_oracle* o, scratch* s, const uint8_t* text, size_t len, idvec* out) {
    const uint8_t* z = memchr(text, 0, len);
    if (z) len = (size_t)(z - text);
    if (len == 0) return;
    int add_prefix = text[0] != ' ';          /* core.c:365 */
    int add_prefix_token = !add_prefix;       /* core.c:366 */
    size_t pos = 0;
    for (;;) {
        size_t wl = hut_oracle_next_pretoken(text, len, pos);
        if (wl == 0) break;
        if (o->limit_mode == 0 && wl > HUT_ORACLE_REF_WORD_LIMIT) break; /* core.c:402-407, 503 */
        if (add_prefix_token && o->prefix) { /* core.c:421-446 */
            byte_encode(o, o->prefix, o->prefix_len, NULL, 0, &s->enc);
            scratch_reserve(s, s->enc.n + 1);
            size_t m = split_symbols(s->enc.p, s->enc.n, 0, s->sp);
            merge_symbols(o, s, s->enc.p, m, out);
            add_prefix_token = 0;
        }
        byte_encode(o, text + pos, wl, (add_prefix && o->prefix) ? o->prefix : NULL, o->prefix_len, &s->enc);
        add_prefix = 0; /* core.c:451 */
        scratch_reserve(s, s->enc.n + 1);
        if (o->rules_cap) { /* core.c:456-474: one UTF-8 character per symbol, id or -1 */
            size_t m = split_symbols(s->enc.p, s->enc.n, 0, s->sp);
            int32_t* tok = malloc((m ? m : 1) * sizeof *tok);
            for (size_t i = 0; i < m; ++i)
                tok[i] = omap_get(&o->vocab, s->enc.p + s->sp[i].start, s->sp[i].end - s->sp[i].start);
            merge_ids(o, s, tok, m, out);
            free(tok);
        } else {
            size_t m = split_symbols(s->enc.p, s->enc.n, 1, s->sp);
            merge_symbols(o, s, s->enc.p, m, out);
        }
        pos += wl;
    }
}

int hut_oracle_encode(const hut_oracle* o, const uint8_t* text, size_t len, int32_t** ids, size_t* n_ids) {
    scratch s;
    memset(&s, 0, sizeof s);
    idvec v = {0};
    encode_doc(o, &s, text, len, &v);
    scratch_free(&s);
    *ids = v.d;
    *n_ids = v.n;
    return 0;
}

void hut_oracle_free_ids(int32_t* ids) { free(ids); }

int hut_oracle_encode_batch(const hut_oracle* o, const uint8_t* bytes, const uint64_t* offs, uint64_t ndocs,
                            int32_t** ids, uint64_t* tok_offs) {
    scratch s;
    memset(&s, 0, sizeof s);
    memset(&g_stats, 0, sizeof g_stats);
    idvec v = {0};
    for (uint64_t d = 0; d < ndocs; ++d) {
        tok_offs[d] = v.n;
        encode_doc(o, &s, bytes + offs[d], (size_t)(offs[d + 1] - offs[d]), &v);
    }
    tok_offs[ndocs] = v.n;
    scratch_free(&s);
    *ids = v.d;
    return 0;
}

void hut_oracle_get_stats(hut_oracle_stats* out) { *out = g_stats; }
/* ------------------------------------------------------------------------------------------------
 * decode() (core.c:513-581) + pretokenizer_decode (pretokenizer.c:197-296).
 * Returns 0 and the malloc'd text (*len bytes, may contain 0x00; the reference's binding cuts there), or -1 for an
 * id outside [0, vocab_lines) or carried by no line (core.c:522-530; the second case is an uninitialised pointer in
 * the reference).  Bytes the reference would read past the end of the raw text (a truncated UTF-8 character at its
 * very end) count as 0 / are not copied.
 * ---------------------------------------------------------------------------------------------- */
static void build_dec(hut_oracle* o) {
    o->dec = calloc(o->vocab_lines ? o->vocab_lines : 1, sizeof *o->dec);
    for (size_t i = 0; i < o->vocab.cap; ++i) {
        const omap_slot* s = &o->vocab.slots[i];
        if (!s->key || s->value < 0 || (size_t)s->value >= o->vocab_lines) continue;
        if (!o->dec[s->value] || o->dec[s->value]->seq < s->seq) o->dec[s->value] = s;
    }
}

int hut_oracle_decode(hut_oracle* o, const int32_t* ids, size_t n, uint8_t** text, size_t* len) {
    if (!o->dec) build_dec(o);
    size_t total = 0;
    for (size_t i = 0; i < n; ++i) {
        if (ids[i] < 0 || (size_t)ids[i] >= o->vocab_lines || !o->dec[ids[i]]) return -1;
        total += o->dec[ids[i]]->len;
    }
    uint8_t* raw = calloc(total + 8, 1);
    size_t at = 0;
    for (size_t i = 0; i < n; ++i) {
        memcpy(raw + at, o->dec[ids[i]]->key, o->dec[ids[i]]->len);
        at += o->dec[ids[i]]->len;
    }
    uint8_t* out = malloc(total + 8);
    size_t p = 0, w = 0;
    if (o->prefix && total >= o->prefix_len && memcmp(raw, o->prefix, o->prefix_len) == 0) p = o->prefix_len;
    while (p < total) {
        int best = -1;
        size_t best_len = 0;
        int32_t cur = 0;
        for (size_t q = p; q < total && o->trie_n; ++q) { /* pretokenizer.c:224-240 / 262-278 */
            int32_t c = o->trie[cur].child[raw[q]];
            if (!c) break;
            cur = c;
            if (o->trie[cur].out != -1) {
                best = o->trie[cur].out;
                best_len = o->trie[cur].depth;
            }
        }
        if (best >= 0) {
            out[w++] = (uint8_t)best;
            p += best_len;
            continue;
        }
        const uint8_t* c = raw + p;
        size_t cl = (size_t)lead_len(c[0]);
        if (o->is_byte_encoder) { /* utf8_to_codepoint, pretokenizer.c:176-195,243-255 */
            uint32_t cp;
            if (c[0] < 0x80) cp = c[0];
            else if ((c[0] & 0xE0) == 0xC0) cp = ((c[0] & 0x1Fu) << 6) | (c[1] & 0x3Fu);
            else if ((c[0] & 0xF0) == 0xE0) cp = ((c[0] & 0x0Fu) << 12) | ((c[1] & 0x3Fu) << 6) | (c[2] & 0x3Fu);
            else if ((c[0] & 0xF8) == 0xF0)
                cp = ((c[0] & 0x07u) << 18) | ((c[1] & 0x3Fu) << 12) | ((c[2] & 0x3Fu) << 6) | (c[3] & 0x3Fu);
            else cp = 0xFFFD;
            out[w++] = cp < 256 ? (uint8_t)cp : (uint8_t)'?';
        } else { /* pretokenizer.c:283-287 */
            size_t k = cl < total - p ? cl : total - p;
            memcpy(out + w, c, k);
            w += k;
        }
        p += cl;
    }
    free(raw);
    out[w] = 0;
    *text = out;
    *len = w;
    return 0;
}

void hut_oracle_free_text(uint8_t* t) { free(t); }

