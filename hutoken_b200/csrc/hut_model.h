/* Host-side model: file loaders + the tables the sm_100a kernels consume.
 *
 * Replaces (reference, matyasosvath/hutoken): the vocab / special-chars loaders of p_initialize
 * (lib.c:243-566, helper.c:82-128) and the string->id hash map (hashmap.c, hash.c, bpe.c:18-30) as the
 * source of pair ranks (core.c:700-722).  On the GPU there are no strings: every token that can exist
 * during merging is a "unit" with a dense id, and the rank of an adjacent pair is looked up by
 * (unit_left, unit_right) in an open-addressing table DERIVED from the string vocab at init
 * (SURVEY.md F3 / H1).
 */
#ifndef HUT_MODEL_H
#define HUT_MODEL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HUT_UNIT_BITS 21u
#define HUT_UNIT_MASK ((1u << HUT_UNIT_BITS) - 1u)
#define HUT_MAX_UNITS (HUT_UNIT_MASK - 2u) /* top ids reserved: INERT */
#define HUT_PAIR_EMPTY 0xFFFFFFFFFFFFFFFFull
#define HUT_CHAR_EMPTY 0xFFFFFFFFFFFFFFFFull
#define HUT_TRIE_EMPTY 0xFFFFFFFFFFFFFFFFull
#define HUT_DEC_MISSING 0xFFFFFFFFu

/* error classes -> Python exception types at the CPython boundary (lib.c error conventions) */
enum {
    HUT_OK = 0,
    HUT_ERR_FILE_NOT_FOUND = 1, /* FileNotFoundError */
    HUT_ERR_VALUE = 2,          /* ValueError */
    HUT_ERR_MEMORY = 3,         /* MemoryError */
    HUT_ERR_CUDA = 4,           /* RuntimeError */
    HUT_ERR_UNSUPPORTED = 5,    /* RuntimeError: valid for the reference, outside this engine's scope */
    HUT_ERR_ARG = 6,            /* TypeError / ValueError on bad arguments */
    HUT_ERR_STATE = 7           /* RuntimeError: not initialised */
};

typedef struct hut_model {
    /* ---- units ------------------------------------------------------------------------------
     * [0, n_vocab)            distinct vocab strings, sorted by (value, bytes)
     * [n_vocab, n_units)      virtual units: initial symbols absent from the vocab that occur as a part of
     *                         some vocab string (they can merge but print as -1, core.c:205-207)
     * inert (== n_units)      any other unknown symbol: never merges, prints -1                      */
    uint32_t n_vocab, n_units, inert;
    uint32_t vocab_lines;      /* accepted lines incl. duplicates = vocab_size_decode (lib.c:378) */
    uint8_t** unit_str;        /* [n_units] bytes of each unit */
    uint32_t* unit_len;
    int32_t* out_id;           /* [n_units + 1] id printed for a unit (-1 for virtual / inert) */
    int32_t* rank_of;          /* [n_units + 1] rank a pair gets when it merges INTO this unit (= vocab value) */
    int distinct_values;       /* 1: no two strings share a value => unit order == rank order */
    int identity_ids;          /* 1: out_id[u] == u for all vocab units */
    int32_t max_value;
    int monotone;              /* 1: every pair (L, R) -> M ranks above L and above R whenever those are themselves the
                                  product of a pair (true for any vocabulary grown by BPE training).  Then "smallest
                                  rank first, leftmost first" (core.c:111-176, queue.c:162-164) applies all pairs of
                                  the current minimum rank before anything else: huge_kernel's cooperative rounds. */

    /* ---- pair table: (L,R) -> M.  slot = L << 42 | R << 21 | M ; empty = all ones -------------- */
    uint64_t* pair_tab;
    uint32_t pair_cap; /* power of two */
    uint32_t pair_count;

    /* ---- initial symbols ------------------------------------------------------------------------
     * byte mode : input byte b -> units sym_pool[sym_off[b] .. sym_off[b+1])   (>= 1 each, b != 0)
     * plain mode: a char whose LEAD byte b has a replacement (pretokenizer.c:129-136) -> the same pool
     *             range; otherwise the raw char is one symbol: ASCII via ascii_unit[], multi-byte via
     *             char_tab (key = bytes packed little-endian, slot = key << 32 | unit).             */
    int is_byte_encoder;
    uint32_t sym_off[257];
    uint32_t* sym_pool;
    uint32_t ascii_unit[128];
    uint64_t* char_tab;
    uint32_t char_cap;
    uint32_t max_sym_per_char; /* Kmax: 1 for every practical table */

    /* ---- prefix (core.c:365-366, 421-451) -------------------------------------------------------- */
    int has_prefix;
    uint32_t n_prefix_syms;    /* symbols of the raw prefix string, prepended to the first pretoken */
    uint32_t* prefix_syms;
    uint32_t n_prefix_alone;   /* ids emitted before the first pretoken when the text starts with ' ' */
    int32_t* prefix_alone_ids;

    uint32_t max_unit_len;

    /* ---- decode side (core.c:513-581, pretokenizer.c:197-296, ac.c:27-50) ------------------------------
     * dec_off/dec_pool: bytes of the token printed for each id in [0, vocab_lines); dec_off[i] == HUT_DEC_MISSING
     *   marks an id no vocab line carries (the reference reads an uninitialised pointer there; we refuse it).
     * trie of the special-chars replacement strings, one node per distinct prefix, root = 0:
     *   trie_edge: open addressing, slot = (node << 8 | byte) << 32 | child, empty = all ones;
     *   trie_out[node] = byte the replacement stands for, or -1;  depth of a node = pattern length.          */
    uint32_t* dec_off;  /* [vocab_lines + 1] */
    uint32_t* dec_len;  /* [vocab_lines]; HUT_DEC_MISSING if absent */
    uint8_t* dec_pool;
    uint32_t dec_pool_len;
    uint64_t* trie_edge;
    uint32_t trie_cap; /* power of two */
    int16_t* trie_out;
    uint8_t* trie_depth;
    uint32_t trie_nodes;
    uint32_t trie_root_child[256]; /* child of the root per byte, 0 = none */
    uint8_t* prefix_bytes; /* raw prefix string (decode strips it once at the start, pretokenizer.c:209-215) */
    uint32_t prefix_len;

    /* ---- merges-file mode (core.c:211-337): units are ids, see apply_merges() in hut_model.c ---- */
    int merges_mode;
    uint32_t n_merge_rules;
} hut_model;

/* Loads the files and builds every table (merges_path may be NULL: string-keyed ranks, core.c:66-209).  Returns HUT_OK or an error class; msg receives the exact
 * message string the reference raises for the same failure where one exists. */
int hut_model_load(hut_model** out, const char* vocab_path, const char* special_path, const char* prefix,
                   int is_byte_encoder, const char* merges_path, char* msg, size_t msglen);
void hut_model_free(hut_model* m);

/* table hash shared by host builder and device probes */
static inline uint64_t hut_mix64(uint64_t x) {
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    return x;
}

/* slot of the pair (l, r) in the pair table: cheap enough for the merge loop (2 multiply-adds, shift, xor-and) */
#ifdef __CUDACC__
__host__ __device__
#endif
static inline uint32_t hut_pair_hash(uint32_t l, uint32_t r, uint32_t mask) {
    uint32_t h = l * 0x9E3779B1u + r * 0x85EBCA77u;
    return (h ^ (h >> 15)) & mask;
}

/* host-side lookups (used by the builder itself and by tests/host_harness.c) */
uint32_t hut_model_pair_lookup(const hut_model* m, uint32_t l, uint32_t r); /* unit or HUT_UNIT_MASK if none */
uint32_t hut_model_char_lookup(const hut_model* m, const uint8_t* p, uint32_t len);
uint32_t hut_model_trie_child(const hut_model* m, uint32_t node, uint8_t byte); /* 0 if none */

#ifdef __CUDACC__
__host__ __device__
#endif
static inline uint32_t hut_trie_hash(uint32_t key, uint32_t mask) { return (key * 0x9E3779B1u >> 7) & mask; }

#ifdef __cplusplus
}
#endif
#endif
