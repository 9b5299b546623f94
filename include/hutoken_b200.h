/* hutoken_b200 -- C-ABI of the B200-native BPE encode engine (libhutoken_b200.so).
 *
 * Drop-in boundary: the seam between hutoken's CPython layer and its core,
 *     void encode(struct EncodeTask*)            reference include/hutoken/core.h:11
 *     struct EncodeTask / struct EncodeContext   reference include/hutoken/taskqueue.h:16-25,40-45
 * called per document from p_encode (src/lib.c:692-697) and from the batch_encode worker threads
 * (src/lib.c:48-57, 779-794).  The functions below are what a cgo/ctypes/CPython binding of that path
 * binds instead; INTEGRATION.md shows the binding.  Plain C types only: no torch, no C++ in signatures.
 *
 * All work runs in sm_100a CUDA kernels.  There is NO CPU fallback: without a CUDA device hut_create fails.
 */
#ifndef HUTOKEN_B200_H
#define HUTOKEN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hut_ctx hut_ctx;

/* status / error classes (mapped to Python exception types by the _hutoken extension) */
#define HUT_STATUS_OK 0
#define HUT_STATUS_FILE_NOT_FOUND 1 /* FileNotFoundError  (lib.c:248, 466) */
#define HUT_STATUS_VALUE 2          /* ValueError         (lib.c:277-355, 494-543) */
#define HUT_STATUS_MEMORY 3         /* MemoryError */
#define HUT_STATUS_CUDA 4           /* RuntimeError: CUDA failure / no device */
#define HUT_STATUS_UNSUPPORTED 5    /* RuntimeError: accepted by the reference, outside this engine's scope */
#define HUT_STATUS_ARG 6            /* bad argument */
#define HUT_STATUS_STATE 7          /* not initialised */
#define HUT_STATUS_CAPACITY 8       /* caller-provided output buffer too small */

/* Mirrors the arguments of hutoken.initialize / _hutoken.initialize (lib.c:185-232, format "ss|zpizz"). */
typedef struct hut_model_desc {
    const char* vocab_file_path;   /* "(0xHH)+ == <id>" lines, reference README.md:260-319 */
    const char* special_file_path; /* "<byte> == <replacement>" lines; mandatory like the reference (lib.c:203) */
    const char* prefix;            /* NULL or e.g. U+2581 for sentencepiece-style vocabularies */
    int is_byte_encoder;           /* GPT-2 byte-level mode (pretokenizer.c:138-146) */
    int special_token_id;          /* parsed and unused by the reference (lib.c:197); ignored */
    const char* pattern;           /* regex pretokenisation (core.c:352-399): must be NULL -> HUT_STATUS_UNSUPPORTED */
    const char* merges_file_path;  /* NULL, or a merges file (lib.c:573-663): ranks and merged ids then come from
                                      (left_id, right_id) rules, core.c:211-337 */
} hut_model_desc;

typedef struct hut_options {
    uint32_t struct_size;      /* sizeof(hut_options), for forward compatibility */
    uint32_t cache_log2_slots; /* unique-pretoken cache size; 0 = default (2^22 slots of 32 B per device) */
    uint64_t max_pretoken;     /* 0 = reference-compatible: a document is cut at its first pretoken longer than
                                  262144 bytes (core.c:27-28,402-407,503).  Otherwise: no cut; pretokens up to this
                                  many bytes are merged (block-cooperative path), longer ones are an error. */
    uint32_t big_slots;        /* working areas (= CTAs of huge_kernel) for pretokens of more than 4160 symbols; 0 = default
                                  (8; 2 when max_pretoken > 64 MiB).  Allocated when the first such pretoken shows up. */
    uint32_t chunk_mb;         /* hut_encode_batch streams the batch through the devices in chunks of about this many
                                  MiB so that upload, kernels and download overlap; 0 = default (32).  hut_decode_batch
                                  does the same with chunks of whole documents holding about this many MiB of ids. */
} hut_options;

/* Replaces p_initialize (lib.c:185-666) + initialize_context (lib.c:128-183): loads the files, derives the
 * device tables and uploads them to every device in `devices` (NULL/0 = device 0 only, or the device named by
 * LOCAL_RANK when set).  On failure returns a status and writes the reference's message text into err. */
int hut_create(const hut_model_desc* desc, const hut_options* opt, const int* devices, int ndev, hut_ctx** out,
               char* err, size_t errlen);
void hut_destroy(hut_ctx* ctx);
const char* hut_last_error(const hut_ctx* ctx);

typedef struct hut_result {
    const int32_t* ids;          /* n_ids token ids; may contain -1 (unknown symbol, core.c:207) */
    const uint64_t* tok_offsets; /* n_docs + 1: ids of document d are ids[tok_offsets[d] .. tok_offsets[d+1]) */
    uint64_t n_ids;
    uint64_t n_docs;
} hut_result;

/* Replaces the per-document encode() loop of p_batch_encode (lib.c:722-874) / p_encode (lib.c:668-720).
 * Host buffers: documents are bytes[doc_offsets[i] .. doc_offsets[i+1]) (UTF-8, packed, no separators).
 * A 0x00 byte inside a document is an error (HUT_STATUS_VALUE): the binding cuts documents at their first NUL
 * like the reference's strdup (lib.c:772).  The batch is cut into contiguous document ranges balanced by bytes
 * (hut_plan_shards) which are dealt round-robin to the context's devices and streamed through them (upload of one
 * chunk overlaps the kernels of the previous and the download of the one before); there is no inter-device
 * communication.  A small batch (<= 69 120 bytes, <= 2 048 documents: every single sentence) takes one kernel launch on
 * pinned host buffers instead (no copies).  The result lives in
 * pinned host memory owned by the context and stays valid until the next call on this context. */
int hut_encode_batch(hut_ctx* ctx, const uint8_t* bytes, const uint64_t* doc_offsets, uint64_t n_docs,
                     hut_result* result);

/* The sharding rule of hut_encode_batch as a pure host function (no CUDA call): cuts documents [0, n_docs) into
 * n_shards contiguous ranges balanced by bytes; shard s gets documents [bounds[s], bounds[s+1]).  bounds holds
 * n_shards + 1 entries.  Replaces the dynamic doc->thread assignment of taskqueue_get (taskqueue.c:16-35): shards
 * are independent, so ranks (or devices) only ever exchange their per-shard id counts. */
int hut_plan_shards(const uint64_t* doc_offsets, uint64_t n_docs, int n_shards, uint64_t* bounds);

/* Same path with every buffer already resident on device `dev_index` of the context (benchmark / pipeline use):
 * d_bytes must be 16-byte aligned; d_ids holds ids_capacity entries, d_tok_offsets n_docs+1 entries.
 * stream is a cudaStream_t (NULL = the context's own stream).  Returns after the ids are complete on device. */
int hut_encode_device(hut_ctx* ctx, int dev_index, const uint8_t* d_bytes, uint64_t n_bytes,
                      const uint64_t* d_doc_offsets, uint64_t n_docs, int32_t* d_ids, uint64_t ids_capacity,
                      uint64_t* d_tok_offsets, uint64_t* n_ids_out, void* stream);

/* ---- decode (core.c:513-581 decode(), pretokenizer.c:197-296 pretokenizer_decode; p_decode / p_batch_decode,
 * lib.c:876-1085) ------------------------------------------------------------------------------------------- */
typedef struct hut_text_result {
    const uint8_t* text;          /* decoded bytes of all documents, packed; NUL-terminated after n_bytes */
    const uint64_t* text_offsets; /* n_docs + 1: document d is text[text_offsets[d] .. text_offsets[d+1]) */
    uint64_t n_bytes;
    uint64_t n_docs;
} hut_text_result;

/* Host buffers: ids of document d are ids[tok_offsets[d] .. tok_offsets[d+1]) (the layout hut_encode_batch returns).
 * An id < 0, >= hut_vocab_lines() or carried by no vocab line fails the whole call with HUT_STATUS_VALUE and the
 * reference's message "Element must be non-negative and less than vocab size." (core.c:522-530).  The text may
 * contain 0x00 bytes (a replacement standing for byte 0); the reference's binding cuts each string there
 * (PyUnicode_FromString, lib.c:946,1066) and so does ours.  The result lives in pinned host memory owned by the
 * context until the next decode call.  A batch of up to one and a half chunks (hut_options.chunk_mb) runs on the
 * context's first device in one piece; a larger one is streamed in chunks of whole documents, dealt round-robin to
 * all devices of the context, ids going up while text comes down. */
int hut_decode_batch(hut_ctx* ctx, const int32_t* ids, const uint64_t* tok_offsets, uint64_t n_docs,
                     hut_text_result* result);

/* Same with ids / offsets resident on the context's first device; the outputs are device pointers owned by the
 * context (valid until the next decode call).  d_tok_offsets must start at 0.  stream: cudaStream_t or NULL. */
int hut_decode_device(hut_ctx* ctx, const int32_t* d_ids, uint64_t n_ids, const uint64_t* d_tok_offsets,
                      uint64_t n_docs, const uint8_t** d_text_out, const uint64_t** d_text_offsets_out,
                      uint64_t* n_bytes_out, void* stream);

/* device time (CUDA events on the launch stream) and kernel count of the last decode call */
int hut_decode_stats(const hut_ctx* ctx, double* kernel_ms, uint64_t* kernel_launches);

/* Which kernels the last decode call ran: 1 = token-wise (every token of the batch decodes independently of its
 * neighbours: lengths, scan, one copy pass), 2 = general (greedy scan of the concatenated token strings, tile by
 * tile; taken when a token ends inside a special-chars pattern or inside a character), 3 = token-wise in one kernel
 * launch on pinned host buffers (hut_decode_batch of at most 2048 ids in at most 256 lists), 0 = no call yet.  All
 * produce the bytes of pretokenizer_decode (pretokenizer.c:197-296).  HUTOKEN_B200_DECODE=general in the environment
 * keeps a context to the general kernels. */
int hut_decode_path(const hut_ctx* ctx);

/* hut_decode_device for callers that must not allocate (code under CUDA graph capture set-up rules, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same token-wise kernels and the same bytes as hut_decode_device, but
 *   - the caller owns the outputs: d_text (text_cap bytes) and d_text_offsets (n_docs + 1 entries), on the device;
 *   - the scratch (lengths per 256 ids, their scan, first document per sub-block, flags) lives in a file-scope
 *     device array of HUT_NOALLOC_SCRATCH_BYTES inside the library, which bounds a call to HUT_NOALLOC_MAX_IDS ids;
 *     more ids -> HUT_STATUS_ARG ("... at most 16777216 ids ...");
 *   - nothing is allocated or freed (no cudaMalloc / cudaFree / cudaMallocHost); the call synchronises `stream`
 *     once to read the text length and the flags.
 * The decode tables must exist: call any other hut_decode_* function once per context first (that call allocates);
 * otherwise HUT_STATUS_STATE.  A batch the token-wise kernels cannot take (a token that ends inside a character or
 * inside a special-chars pattern completed by its neighbours) is refused with HUT_STATUS_VALUE: use
 * hut_decode_device for it.  text_cap too small -> HUT_STATUS_CAPACITY, *n_bytes_out = bytes needed, no text written.
 * The array is shared by every call of the process on that GPU and keeps the last call's contents (what has to start
 * from zero is cleared by each call): calls that overlap on the GPU -- two streams, two contexts, two threads --
 * race on it; serialise them.  d_tok_offsets must start at 0 and end at n_ids. */
#define HUT_NOALLOC_MAX_IDS (1ull << 24)
#define HUT_NOALLOC_SCRATCH_BYTES (1026u << 10)
int hut_decode_device_noalloc(hut_ctx* ctx, const int32_t* d_ids, uint64_t n_ids, const uint64_t* d_tok_offsets,
                              uint64_t n_docs, uint8_t* d_text, uint64_t text_cap, uint64_t* d_text_offsets,
                              uint64_t* n_bytes_out, void* stream);

/* Drops the unique-pretoken cache of every device (it otherwise persists across calls). */
int hut_cache_reset(hut_ctx* ctx);

/* pinned host memory helpers for callers that want zero-staging H2D copies */
void* hut_host_alloc(size_t bytes);
void hut_host_free(void* p);

typedef struct hut_stats {
    uint64_t pretokens;     /* pretokens seen by the last call */
    uint64_t cache_hits;    /* answered by the unique-pretoken cache */
    uint64_t warp_merges;   /* merged by the warp-per-pretoken path */
    uint64_t block_merges;  /* merged by the block-cooperative path */
    uint64_t big_merges;    /* ... of which outgrew a tile (global working area) */
    uint64_t cut_documents; /* documents cut at an over-long pretoken (reference-compatible mode) */
    uint64_t kernel_launches; /* kernels launched by the last call */
    double kernel_ms;       /* device time of the last call's kernels (CUDA events on the launch stream) */
    double main_kernel_ms;  /* ... of which the fused tile kernel */
} hut_stats;
int hut_get_stats(const hut_ctx* ctx, int dev_index, hut_stats* out);

/* model introspection (used by decode-side callers and tests) */
uint32_t hut_vocab_lines(const hut_ctx* ctx); /* vocab_size_decode of the reference (lib.c:378) */
int hut_device_count(const hut_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif
