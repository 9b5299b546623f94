/* TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement ("oracle") of the hutoken encode path, used by tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg as the CHECKER for the CUDA engine.  Nothing in hutoken_b200/ may include,
 * link or call this.  Every function cites the reference file:line (under matyasosvath/hutoken) whose
 * behaviour it restates.  Parity of this restatement is PINNED against the unmodified reference compiled
 * into oracle/_ref (tests/test_oracle_vs_ref.py, tools/make_golden.py -> tests/golden/).
 */
#ifndef HUT_ORACLE_H
#define HUT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hut_oracle hut_oracle;

/* Pretoken limit of the reference: 16 MiB arena / 64 (core.c:27-28,402-407). */
#define HUT_ORACLE_REF_WORD_LIMIT 262144u

/* Load vocab + special-chars files exactly like p_initialize (lib.c:243-566).
 * prefix may be NULL.  On failure returns NULL and writes "<ExceptionType>: <message>" into err. */
hut_oracle* hut_oracle_load(const char* vocab_path, const char* special_path, const char* prefix,
                            int is_byte_encoder, char* err, size_t errlen);
void hut_oracle_free(hut_oracle* o);

/* Optional merges file (lib.c:573-663): switches encode to the id-keyed path (core.c:211-337) unless the file
 * holds no candidate line.  Returns 1, or 0 with err set when the file cannot be opened. */
int hut_oracle_load_merges(hut_oracle* o, const char* path, char* err, size_t errlen);
size_t hut_oracle_merge_rules(const hut_oracle* o); /* distinct (left_id, right_id) rules */
int hut_oracle_merges_active(const hut_oracle* o);

/* limit_mode 0: reference-compatible (document cut at the first pretoken > 262144 bytes, core.c:402-407,503)
 * limit_mode 1: no limit (the restatement has no arena / stack bounds). */
void hut_oracle_set_limit_mode(hut_oracle* o, int limit_mode);

/* encode() of core.c:339-511 on text[0..len) (also stops at the first NUL byte).
 * Appends ids to a malloc'd array returned in *ids (caller frees with hut_oracle_free_ids). */
int hut_oracle_encode(const hut_oracle* o, const uint8_t* text, size_t len, int32_t** ids, size_t* n_ids);
void hut_oracle_free_ids(int32_t* ids);

/* Batch form used for golden vectors and the CPU baseline: docs are bytes[offs[i]..offs[i+1]).
 * tok_offs must hold ndocs+1 entries.  ids returned malloc'd. */
int hut_oracle_encode_batch(const hut_oracle* o, const uint8_t* bytes, const uint64_t* offs, uint64_t ndocs,
                            int32_t** ids, uint64_t* tok_offs);

/* decode() (core.c:513-581) + pretokenizer_decode (pretokenizer.c:197-296): ids -> bytes.  Returns 0, or -1 for an
 * id that cannot be decoded ("Element must be non-negative and less than vocab size.").  *text is malloc'd
 * (hut_oracle_free_text), NUL-terminated after *len bytes and may contain 0x00 inside. */
int hut_oracle_decode(hut_oracle* o, const int32_t* ids, size_t n, uint8_t** text, size_t* len);
void hut_oracle_free_text(uint8_t* t);

/* parser_next_token (parser.c:24-88): length in bytes of the pretoken starting at text[pos]; 0 at end. */
size_t hut_oracle_next_pretoken(const uint8_t* text, size_t len, size_t pos);

/* Statistics of the last hut_oracle_encode_batch call on this thread (for bench.py's probe-inclusive
 * roofline figure): pretokens, initial symbols m, output tokens t, and
 * P = sum((m-1) + 2(m-t) + t). */
typedef struct {
    uint64_t pretokens, symbols, tokens, probes;
} hut_oracle_stats;
void hut_oracle_get_stats(hut_oracle_stats* out);

/* vocab introspection for tests */
int32_t hut_oracle_vocab_lookup(const hut_oracle* o, const uint8_t* key, size_t len); /* -1 if absent */
size_t hut_oracle_vocab_lines(const hut_oracle* o);

#ifdef __cplusplus
}
#endif
#endif
