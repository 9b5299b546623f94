/* libhutoken_b200.so -- C-ABI implementation (include/hutoken_b200.h) and host orchestration.
 *
 * Replaces, for the encode path, hutoken's p_initialize / p_encode / p_batch_encode plumbing (lib.c) and its
 * pthread task queue (taskqueue.c, lib.c:779-794): documents are sharded by bytes over the context's devices,
 * each shard runs the kernel pipeline of hut_pipeline.cuh on its own stream; there is no collective.
 * No CPU fallback: every id is produced by those kernels.
 */
#include <cuda_runtime.h>
#include <stdarg.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/hutoken_b200.h"
#include "hut_model.h"
#include "hut_decode.cuh"
#include "hut_pipeline.cuh"

using namespace hut;

#define MAX_DEV 16

/* one launched shard whose counters have not been read yet.  Each flight owns the device-side state its kernels
 * write (counters, look-back words, deferred-pretoken list), so that the next chunk can be queued behind it without
 * touching anything the host may still want to look at (patch phase, cuts). */
struct hut_flight {
    Job job;
    Counters* ctr;       // device counters
    Counters* h_ctr;     // pinned mirror of the device counters, copied at the end of the launch
    uint32_t* tilebuf;   // cnt | tot | first_doc | def_first (ntiles + 1 uint32 each) | slot_off (uint64 each)
    size_t tilebuf_cap;
    unsigned long long* bsum;
    size_t bsum_cap;
    DefEntry* def;
    size_t def_cap;
    cudaEvent_t ev[4];   // start, before / after tile_kernel, end
    cudaEvent_t done;    // counters are in h_ctr
};

/* device-side buffers of one chunk of the host-buffer API (two per device: chunk i+1 uploads while chunk i runs
 * and chunk i-1 downloads) */
struct hut_chunkbuf {
    uint8_t* d_text;
    size_t d_text_cap;
    unsigned long long* d_off;
    size_t d_off_cap;
    int32_t* d_ids;
    size_t d_ids_cap;
    unsigned long long* d_tok;
    size_t d_tok_cap;
    unsigned long long* h_off;  // pinned: this chunk's rebased document offsets
    size_t h_off_cap;
    cudaEvent_t ev_in, ev_out;
    int out_pending;
};

struct hut_dev {
    int device;
    cudaStream_t stream;        // kernels
    cudaStream_t s_in, s_out;   // H2D / D2H copies of the host-buffer API
    DevModel dm;
    void* allocs[64];
    int nalloc;
    CacheSlot* cache;
    uint32_t cache_bits;
    CacheSlot2* cache2;
    uint32_t cache2_bits;
    CutRanges* cut;
    uint32_t* cta_scratch;
    unsigned long long cta_scratch_words;
    uint32_t med_cap;
    uint32_t* big_pool;
    unsigned long long big_slot_words;
    uint32_t big_cap, big_slots;
    int grid, long_grid, sms;
    int mid_grid[2];  /* resident one-warp CTAs of mid_kernel<1024> / <4096> */
    uint32_t stride, slot_extra;
    // grow-only temporaries shared by the device's flights (they run one after the other on the device's stream)
    int32_t* stage;
    size_t stage_cap;
    int32_t* long_ids;
    size_t long_cap;
    int32_t* tmp_ids;
    size_t tmp_ids_cap;
    hut_flight fl[2];
    hut_chunkbuf cb[2];
    // small-batch path (small_kernel): pinned buffers the device reads / writes directly
    uint8_t* sm_text;
    unsigned long long* sm_off;
    Counters* sm_ctr;            // only the head (before events) is used
    volatile unsigned int* sm_done;
    unsigned int sm_seq;
    hut_stats stats;
};

struct hut_decstate;
static void dec_free(hut_decstate* s);

struct hut_ctx {
    hut_model* model;
    hut_decstate* dec;  // decode-side state of the first device, created by the first hut_decode_* call
    hut_decstate* dec_x[MAX_DEV];  // ... of the other devices (index = device index; [0] unused): streamed host batches only
    int ndev;
    hut_dev dev[MAX_DEV];
    hut_options opt;
    char err[512];
    // pinned result of the host-buffer API
    int32_t* h_ids;
    size_t h_ids_cap;
    unsigned long long* h_tok;
    size_t h_tok_cap;
};

static int set_err(hut_ctx* ctx, int status, const char* fmt, const char* a) {
    if (ctx) snprintf(ctx->err, sizeof ctx->err, fmt, a);
    return status;
}
static int set_errf(hut_ctx* ctx, int status, const char* fmt, ...) __attribute__((format(printf, 3, 4)));
static int set_errf(hut_ctx* ctx, int status, const char* fmt, ...) {
    if (ctx) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(ctx->err, sizeof ctx->err, fmt, ap);
        va_end(ap);
    }
    return status;
}

#define CU(ctx, call)                                                                      \
    do {                                                                                   \
        cudaError_t e_ = (call);                                                           \
        if (e_ != cudaSuccess) return set_err(ctx, HUT_STATUS_CUDA, "CUDA error: %s", cudaGetErrorString(e_)); \
    } while (0)

template <class T>
static cudaError_t upload(hut_dev* d, const T* src, size_t n, const T** out) {
    T* p = nullptr;
    if (d->nalloc >= (int)(sizeof d->allocs / sizeof d->allocs[0])) return cudaErrorMemoryAllocation;  // list full
    cudaError_t e = cudaMalloc(&p, (n ? n : 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    d->allocs[d->nalloc++] = p;
    if (n) e = cudaMemcpy(p, src, n * sizeof(T), cudaMemcpyHostToDevice);
    *out = p;
    return e;
}

/* uninitialised device memory on the device's allocation list */
static cudaError_t dev_alloc(hut_dev* d, void** out, size_t bytes) {
    if (d->nalloc >= (int)(sizeof d->allocs / sizeof d->allocs[0])) return cudaErrorMemoryAllocation;  // list full
    cudaError_t e = cudaMalloc(out, bytes ? bytes : 1);
    if (e == cudaSuccess) d->allocs[d->nalloc++] = *out;
    return e;
}

static int dev_init(hut_ctx* ctx, hut_dev* d, int device) {
    const hut_model* m = ctx->model;
    memset(d, 0, sizeof *d);
    d->device = device;
    CU(ctx, cudaSetDevice(device));
    CU(ctx, cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking));
    CU(ctx, cudaStreamCreateWithFlags(&d->s_in, cudaStreamNonBlocking));
    CU(ctx, cudaStreamCreateWithFlags(&d->s_out, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        for (int i = 0; i < 4; ++i) CU(ctx, cudaEventCreate(&d->fl[b].ev[i]));
        CU(ctx, cudaEventCreateWithFlags(&d->fl[b].done, cudaEventDisableTiming));
        CU(ctx, cudaMallocHost(&d->fl[b].h_ctr, sizeof(Counters)));
        CU(ctx, cudaMalloc(&d->fl[b].ctr, sizeof(Counters)));
        CU(ctx, cudaMemset(d->fl[b].ctr, 0, sizeof(Counters)));
        CU(ctx, cudaEventCreateWithFlags(&d->cb[b].ev_in, cudaEventDisableTiming));
        CU(ctx, cudaEventCreateWithFlags(&d->cb[b].ev_out, cudaEventDisableTiming));
    }
    DevModel& dm = d->dm;
    CU(ctx, upload(d, (const unsigned long long*)m->pair_tab, m->pair_cap, &dm.pair_tab));
    dm.pair_mask = m->pair_cap - 1;
    CU(ctx, upload(d, m->out_id, m->n_units + 1, &dm.out_id));
    CU(ctx, upload(d, m->rank_of, m->n_units + 1, &dm.rank_of));
    dm.n_vocab = m->n_vocab;
    dm.n_units = m->n_units;
    dm.inert = m->inert;
    dm.distinct_values = m->distinct_values;
    dm.identity_ids = m->identity_ids;
    dm.byte_mode = m->is_byte_encoder;
    dm.has_prefix = m->has_prefix;
    dm.monotone = m->monotone;
    dm.cache_ok = (m->max_value < (1 << 20)) ? 1 : 0;  // ids are packed as signed 21-bit fields
    dm.kmax = m->max_sym_per_char ? m->max_sym_per_char : 1;
    dm.miss_len_max = MAXW;
    // a queued miss packs its position in the tile's staging slot into 18 bits (hut_tile.cuh q_pack): a model whose slots
    // are larger leaves every miss to the long-path kernels
    if (992ull * dm.kmax + 960ull * (m->n_prefix_syms + m->n_prefix_alone + 1) >= (1ull << 18)) dm.miss_len_max = 0;
    dm.first_extra = m->has_prefix ? (m->n_prefix_syms > m->n_prefix_alone ? m->n_prefix_syms : m->n_prefix_alone) : 0u;
    CU(ctx, upload(d, m->sym_off, 257, &dm.sym_off));
    CU(ctx, upload(d, m->sym_pool, m->sym_off[256], &dm.sym_pool));
    uint32_t byte1[256];
    for (int b = 0; b < 256; ++b) {
        uint32_t n = m->sym_off[b + 1] - m->sym_off[b];
        if (n == 1) byte1[b] = m->sym_pool[m->sym_off[b]];
        else if (n == 0) byte1[b] = m->is_byte_encoder ? m->inert : RAWCHAR;  // byte mode: only b == 0 has none
        else byte1[b] = NONE;
    }
    CU(ctx, upload(d, byte1, 256, &dm.byte1));
    CU(ctx, upload(d, m->ascii_unit, 128, &dm.ascii_unit));
    if (m->char_cap) {
        CU(ctx, upload(d, (const unsigned long long*)m->char_tab, m->char_cap, &dm.char_tab));
        dm.char_mask = m->char_cap - 1;
    } else {
        dm.char_tab = nullptr;
        dm.char_mask = 0;
    }
    CU(ctx, upload(d, m->prefix_syms, m->n_prefix_syms, &dm.prefix_syms));
    dm.n_prefix_syms = m->n_prefix_syms;
    CU(ctx, upload(d, m->prefix_alone_ids, m->n_prefix_alone, &dm.prefix_alone));
    dm.n_prefix_alone = m->n_prefix_alone;
    dm.compat_cut = ctx->opt.max_pretoken == 0;
    dm.limit = ctx->opt.max_pretoken ? ctx->opt.max_pretoken : 262144ull;  // core.c:27-28

    d->cache_bits = ctx->opt.cache_log2_slots ? ctx->opt.cache_log2_slots : 22;
    if (!ctx->opt.cache_log2_slots && getenv("HUTOKEN_B200_CACHE_LOG2")) d->cache_bits = (uint32_t)atoi(getenv("HUTOKEN_B200_CACHE_LOG2"));  // tuning sweeps
    if (d->cache_bits < 10) d->cache_bits = 10;
    if (d->cache_bits > 26) d->cache_bits = 26;
    CU(ctx, cudaMalloc(&d->cache, sizeof(CacheSlot) << d->cache_bits));
    CU(ctx, cudaMemset(d->cache, 0, sizeof(CacheSlot) << d->cache_bits));
    d->cache2_bits = d->cache_bits > 4 ? d->cache_bits - 4 : 6;  // long pretokens are rare
    CU(ctx, cudaMalloc(&d->cache2, sizeof(CacheSlot2) << d->cache2_bits));
    CU(ctx, cudaMemset(d->cache2, 0, sizeof(CacheSlot2) << d->cache2_bits));
    CU(ctx, cudaMalloc(&d->cut, sizeof(CutRanges)));

    if (getenv("HUTOKEN_B200_CARVEOUT")) {  // experiment: shared-memory carve-out of tile_kernel in percent of the maximum
        const int pc = atoi(getenv("HUTOKEN_B200_CARVEOUT"));
        CU(ctx, cudaFuncSetAttribute(tile_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, pc));
        CU(ctx, cudaFuncSetAttribute(tile_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, pc));
    }
    int per_sm = 0, sms = 0;
    if (m->has_prefix) CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tile_kernel<true>, NT, 0));
    else CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tile_kernel<false>, NT, 0));
    CU(ctx, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (per_sm < 1) per_sm = 1;
    d->sms = sms;
    d->grid = per_sm * sms;
    {   // long_kernel walks its pretokens one merge at a time: many small CTAs in flight hide that latency
        int per = 0;
        CU(ctx, cudaFuncSetAttribute(long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(4 * LSM_CAP * sizeof(uint32_t))));
        CU(ctx, cudaFuncSetAttribute(mid_kernel<4096>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(4 * 4096 * sizeof(uint32_t))));
        int per1 = 0, per4 = 0;
        CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per1, mid_kernel<1024>, 32, 4 * 1024 * sizeof(uint32_t)));
        CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per4, mid_kernel<4096>, 32, 4 * 4096 * sizeof(uint32_t)));
        d->mid_grid[0] = (per1 > 0 ? per1 : 1) * sms;
        d->mid_grid[1] = (per4 > 0 ? per4 : 1) * sms;
        CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, long_kernel, LNT, 4 * LSM_CAP * sizeof(uint32_t)));
        if (per < 1) per = 1;
        if (per > 8) per = 8;
        d->long_grid = per * sms;
    }
    // ids a 960-byte tile can produce: one per symbol of the pretokens STARTING in it (the last one may run up to
    // MAXW - 1 bytes past the tile) + the prefix ids of every document starting in it
    d->stride = (uint32_t)(OWN + MAXW) * dm.kmax;
    d->slot_extra = m->has_prefix ? m->n_prefix_syms + m->n_prefix_alone : 0u;
    d->med_cap = 4096u * dm.kmax + m->n_prefix_syms + 64;
    d->cta_scratch_words = 5ull * d->med_cap + 8;  // U, MG, NX, PV + tournament-tree levels
    CU(ctx, cudaMalloc(&d->cta_scratch, d->cta_scratch_words * sizeof(uint32_t) * d->long_grid));
    // working areas for pretokens of more than med_cap symbols (huge_kernel): allocated when the first one shows up
    d->big_slots = 0;
    d->big_pool = nullptr;
    unsigned long long bc = dm.limit * dm.kmax + m->n_prefix_syms + 64;
    if (bc > 0xFFFFFFF0ull) return set_err(ctx, HUT_STATUS_ARG, "%s", "max_pretoken too large");
    d->big_cap = (uint32_t)bc;
    d->big_slot_words = 7ull * d->big_cap + d->big_cap / 4 + 4096;  // U, MG, NX, PV, tree, lists of the cooperative rounds
    return HUT_STATUS_OK;
}

static void dev_free(hut_dev* d) {
    cudaSetDevice(d->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < d->nalloc; ++i) cudaFree(d->allocs[i]);
    cudaFree(d->cache);
    cudaFree(d->cache2);
    cudaFree(d->cut);
    cudaFree(d->cta_scratch);
    cudaFree(d->big_pool);
    cudaFree(d->stage);
    cudaFree(d->long_ids);
    cudaFree(d->tmp_ids);
    for (int b = 0; b < 2; ++b) {
        hut_chunkbuf* c = &d->cb[b];
        cudaFree(c->d_text);
        cudaFree(c->d_off);
        cudaFree(c->d_ids);
        cudaFree(c->d_tok);
        cudaFreeHost(c->h_off);
        if (c->ev_in) cudaEventDestroy(c->ev_in);
        if (c->ev_out) cudaEventDestroy(c->ev_out);
        cudaFreeHost(d->fl[b].h_ctr);
        cudaFree(d->fl[b].ctr);
        cudaFree(d->fl[b].tilebuf);
        cudaFree(d->fl[b].bsum);
        cudaFree(d->fl[b].def);
        for (int i = 0; i < 4; ++i)
            if (d->fl[b].ev[i]) cudaEventDestroy(d->fl[b].ev[i]);
        if (d->fl[b].done) cudaEventDestroy(d->fl[b].done);
    }
    cudaFreeHost(d->sm_text);
    cudaFreeHost(d->sm_off);
    cudaFreeHost(d->sm_ctr);
    cudaFreeHost((void*)d->sm_done);
    if (d->stream) cudaStreamDestroy(d->stream);
    if (d->s_in) cudaStreamDestroy(d->s_in);
    if (d->s_out) cudaStreamDestroy(d->s_out);
}

extern "C" int hut_create(const hut_model_desc* desc, const hut_options* opt, const int* devices, int ndev,
                          hut_ctx** out, char* err, size_t errlen) {
    auto fail = [&](int st, const char* msg) {
        if (err && errlen) snprintf(err, errlen, "%s", msg);
        return st;
    };
    if (!desc || !out || !desc->vocab_file_path || !desc->special_file_path)
        return fail(HUT_STATUS_ARG, "vocab_file_path and special_file_path are required");
    if (desc->pattern) return fail(HUT_STATUS_UNSUPPORTED, "regex pattern pretokenisation is not supported by the B200 engine");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail(HUT_STATUS_CUDA, "no CUDA device: the hutoken_b200 engine has no CPU fallback");
    hut_ctx* ctx = (hut_ctx*)calloc(1, sizeof *ctx);
    if (!ctx) return fail(HUT_STATUS_MEMORY, "out of memory");
    if (opt) memcpy(&ctx->opt, opt, opt->struct_size < sizeof ctx->opt ? opt->struct_size : sizeof ctx->opt);
    char msg[256] = "";
    int rc = hut_model_load(&ctx->model, desc->vocab_file_path, desc->special_file_path, desc->prefix,
                            desc->is_byte_encoder, desc->merges_file_path, msg, sizeof msg);
    if (rc != HUT_OK) {
        free(ctx);
        return fail(rc, msg);
    }
    int devs[MAX_DEV];
    if (!devices || ndev <= 0) {
        const char* lr = getenv("LOCAL_RANK");
        devs[0] = lr ? atoi(lr) % count : 0;
        ndev = 1;
    } else {
        if (ndev > MAX_DEV) ndev = MAX_DEV;
        for (int i = 0; i < ndev; ++i) devs[i] = devices[i];
    }
    for (int i = 0; i < ndev; ++i) {
        if (devs[i] < 0 || devs[i] >= count) {
            hut_destroy(ctx);
            return fail(HUT_STATUS_ARG, "device index out of range");
        }
        rc = dev_init(ctx, &ctx->dev[i], devs[i]);
        ctx->ndev = i + 1;
        if (rc != HUT_STATUS_OK) {
            fail(rc, ctx->err);
            hut_destroy(ctx);
            return rc;
        }
    }
    *out = ctx;
    return HUT_STATUS_OK;
}

extern "C" void hut_destroy(hut_ctx* ctx) {
    if (!ctx) return;
    if (ctx->ndev) cudaSetDevice(ctx->dev[0].device);
    dec_free(ctx->dec);
    for (int i = 1; i < MAX_DEV; ++i) dec_free(ctx->dec_x[i]);
    for (int i = 0; i < ctx->ndev; ++i) dev_free(&ctx->dev[i]);
    cudaFreeHost(ctx->h_ids);
    cudaFreeHost(ctx->h_tok);
    hut_model_free(ctx->model);
    free(ctx);
}

extern "C" const char* hut_last_error(const hut_ctx* ctx) { return ctx ? ctx->err : "no context"; }
extern "C" uint32_t hut_vocab_lines(const hut_ctx* ctx) { return ctx ? ctx->model->vocab_lines : 0; }
extern "C" int hut_device_count(const hut_ctx* ctx) { return ctx ? ctx->ndev : 0; }

extern "C" void* hut_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void hut_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

extern "C" int hut_cache_reset(hut_ctx* ctx) {
    if (!ctx) return HUT_STATUS_STATE;
    for (int i = 0; i < ctx->ndev; ++i) {
        hut_dev* d = &ctx->dev[i];
        CU(ctx, cudaSetDevice(d->device));
        CU(ctx, cudaMemsetAsync(d->cache, 0, sizeof(CacheSlot) << d->cache_bits, d->stream));
        CU(ctx, cudaMemsetAsync(d->cache2, 0, sizeof(CacheSlot2) << d->cache2_bits, d->stream));
        CU(ctx, cudaStreamSynchronize(d->stream));
    }
    return HUT_STATUS_OK;
}

extern "C" int hut_get_stats(const hut_ctx* ctx, int dev_index, hut_stats* out) {
    if (!ctx || dev_index < 0 || dev_index >= ctx->ndev || !out) return HUT_STATUS_ARG;
    *out = ctx->dev[dev_index].stats;
    return HUT_STATUS_OK;
}

template <class T>
static cudaError_t grow(T** p, size_t* cap, size_t need) {
    if (need <= *cap) return cudaSuccess;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    size_t n = need + need / 8 + 1024;
    cudaError_t e = cudaMalloc(p, n * sizeof(T));
    if (e == cudaSuccess) *cap = n;
    return e;
}

/* the pieces of a flight's tile buffer */
/* launches the kernel pipeline for one resident shard on `st`; the device must be current */
static int launch_shard(hut_ctx* ctx, hut_dev* d, hut_flight* f, const uint8_t* d_bytes, uint64_t n_bytes,
                        const uint64_t* d_doc_offsets, uint64_t n_docs, int32_t* d_ids, uint64_t ids_capacity,
                        uint64_t* d_tok_offsets, cudaStream_t st, size_t def_want, size_t long_want) {
    if (n_docs >= 0xFFFFFFF0ull) return set_err(ctx, HUT_STATUS_ARG, "%s", "too many documents in one shard");
    if ((uintptr_t)d_bytes & 15) return set_err(ctx, HUT_STATUS_ARG, "%s", "device text buffer must be 16-byte aligned");
    const uint64_t ntiles64 = (n_bytes + OWN - 1) / OWN;
    if (ntiles64 >= 0x7FFFFFF0ull || n_bytes >= (1ull << 38))  // byte positions of queued misses are packed into 38 bits
        return set_err(ctx, HUT_STATUS_ARG, "%s", "shard too large (max 256 GiB per device and call)");
    const uint32_t ntiles = (uint32_t)ntiles64;
    if (ntiles == 0) {
        CU(ctx, cudaMemsetAsync(d_tok_offsets, 0, (n_docs + 1) * sizeof(uint64_t), st));
        CU(ctx, cudaMemsetAsync(f->ctr, 0, offsetof(Counters, events), st));
        for (int i = 0; i < 4; ++i) CU(ctx, cudaEventRecord(f->ev[i], st));
        memset(&f->job, 0, sizeof f->job);
        CU(ctx, cudaMemcpyAsync(f->h_ctr, f->ctr, offsetof(Counters, events), cudaMemcpyDeviceToHost, st));
        CU(ctx, cudaEventRecord(f->done, st));
        return HUT_STATUS_OK;
    }
    const uint32_t nblocks = (ntiles + SCAN_BLK - 1) / SCAN_BLK;
    if (def_want < (size_t)ntiles / 8 + 65536) def_want = (size_t)ntiles / 8 + 65536;
    if (long_want < (size_t)n_bytes / 16 + (1u << 20)) long_want = (size_t)n_bytes / 16 + (1u << 20);
    const bool pfx = d->dm.has_prefix != 0;
    CU(ctx, grow(&d->stage, &d->stage_cap, (size_t)ntiles * d->stride + (size_t)n_docs * d->slot_extra + 64));
    CU(ctx, grow(&f->tilebuf, &f->tilebuf_cap, 6 * ((size_t)ntiles + 1) + 4));
    CU(ctx, grow(&f->bsum, &f->bsum_cap, (size_t)nblocks));
    CU(ctx, grow(&f->def, &f->def_cap, def_want));
    CU(ctx, grow(&d->long_ids, &d->long_cap, long_want));
    Job J;
    memset(&J, 0, sizeof J);
    J.text = d_bytes;
    J.nbytes = n_bytes;
    J.doc_off = (const unsigned long long*)d_doc_offsets;
    J.ndocs = n_docs;
    J.out = d_ids;
    J.out_cap = ids_capacity;
    J.tok_off = (unsigned long long*)d_tok_offsets;
    J.ntiles = ntiles;
    J.stride = d->stride;
    J.slot_extra = d->slot_extra;
    J.slot_off = reinterpret_cast<unsigned long long*>(f->tilebuf + 4 * ((size_t)ntiles + 1));  // 16-byte aligned
    J.stage = d->stage;
    J.cnt = f->tilebuf;
    J.tot = J.cnt + ((size_t)ntiles + 1);
    J.first_doc = J.tot + ((size_t)ntiles + 1);
    J.def_first = J.first_doc + ((size_t)ntiles + 1);
    J.bsum = f->bsum;
    J.nblocks = nblocks;
    J.def = f->def;
    J.def_cap = f->def_cap > 0xFFFFFFF0ull ? 0xFFFFFFF0u : (uint32_t)f->def_cap;
    J.long_ids = d->long_ids;
    J.long_cap = d->long_cap;
    J.ctr = f->ctr;
    J.cache = d->cache;
    J.cache_bits = d->cache_bits;
    J.cache2 = d->cache2;
    J.cache2_bits = d->cache2_bits;
    J.cta_scratch = d->cta_scratch;
    J.cta_scratch_words = d->cta_scratch_words;
    J.med_cap = d->med_cap;
    J.big_pool = d->big_pool;
    J.big_slot_words = d->big_slot_words;
    J.big_cap = d->big_cap;
    J.big_slots = d->big_slots;
    CU(ctx, cudaEventRecord(f->ev[0], st));
    CU(ctx, cudaMemsetAsync(J.first_doc, 0xFF, ((size_t)ntiles + 1) * sizeof(uint32_t), st));
    init_kernel<<<1, 32, 0, st>>>(J);
    if (n_docs) doc_kernel<<<(unsigned)((n_docs + 255) / 256), 256, 0, st>>>(J);
    if (pfx)
        slot_docs_kernel<<<(unsigned)(((size_t)ntiles + 1 + 255) / 256), 256, 0, st>>>(
            J, reinterpret_cast<unsigned long long*>(f->tilebuf + 4 * ((size_t)ntiles + 1)));
    CU(ctx, cudaEventRecord(f->ev[1], st));
    const uint32_t want = (ntiles + WPB - 1) / WPB;
    const uint32_t tile_grid = want < (uint32_t)d->grid ? want : (uint32_t)d->grid;
    if (pfx) tile_kernel<true><<<tile_grid, NT, 0, st>>>(d->dm, J);
    else tile_kernel<false><<<tile_grid, NT, 0, st>>>(d->dm, J);
    CU(ctx, cudaEventRecord(f->ev[2], st));
    mid_kernel<1024><<<d->mid_grid[0], 32, 4 * 1024 * sizeof(uint32_t), st>>>(d->dm, J);
    if (d->dm.inert < (1u << 20)) /* keys of rank << 12 | position in 32 bits */
        mid_kernel<4096><<<d->mid_grid[1], 32, 4 * 4096 * sizeof(uint32_t), st>>>(d->dm, J);
    long_kernel<<<d->long_grid, LNT, 4 * LSM_CAP * sizeof(uint32_t), st>>>(d->dm, J);
    if (d->big_slots) huge_kernel<<<d->big_slots, HNT, 0, st>>>(d->dm, J);
    scan1_kernel<<<nblocks, 1024, 0, st>>>(J);
    scan2_kernel<<<1, 1024, 0, st>>>(J);
    const uint32_t cgrid = (ntiles + 7) / 8;
    compact_kernel<<<cgrid < (uint32_t)d->sms * 8u ? cgrid : (uint32_t)d->sms * 8u, 256, 0, st>>>(J);
    fixdoc_kernel<<<(unsigned)((n_docs + 1 + 255) / 256), 256, 0, st>>>(J);
    CU(ctx, cudaEventRecord(f->ev[3], st));
    CU(ctx, cudaGetLastError());
    d->stats.kernel_launches += 7 + (d->dm.inert < (1u << 20) ? 1 : 0) + (n_docs ? 1 : 0) + (pfx ? 1 : 0);
    f->job = J;
    CU(ctx, cudaMemcpyAsync(f->h_ctr, f->ctr, offsetof(Counters, events), cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaEventRecord(f->done, st));
    return HUT_STATUS_OK;
}

static void stats_begin(hut_dev* d) { memset(&d->stats, 0, sizeof d->stats); }

/* waits for the flight's counters: re-runs with larger side buffers if they overflowed, applies document cuts,
 * accumulates stats */
static int finish_shard(hut_ctx* ctx, hut_dev* d, hut_flight* f, uint64_t n_docs, int32_t* d_ids,
                        uint64_t* d_tok_offsets, uint64_t* n_ids_out, cudaStream_t st) {
    const Counters* c = f->h_ctr;
    for (int attempt = 0;; ++attempt) {
        CU(ctx, cudaEventSynchronize(f->done));
        if (!(c->flags & (FLAG_DEF_CAP | FLAG_LONG_CAP | FLAG_BIG_NEED)) || attempt == 4) break;
        // rare: many / very long deferred pretokens.  The counters kept counting, so they state the need.  The staging
        // area and the id buffer for long pretokens are shared by the device's flights: drain the stream first.
        CU(ctx, cudaStreamSynchronize(st));
        if ((c->flags & FLAG_BIG_NEED) && !d->big_pool) {
            uint32_t slots = ctx->opt.big_slots ? ctx->opt.big_slots : (ctx->opt.max_pretoken > (64ull << 20) ? 2 : 8);
            if (slots > 32) slots = 32;
            CU(ctx, cudaMalloc(&d->big_pool, d->big_slot_words * sizeof(uint32_t) * slots));
            d->big_slots = slots;
        }
        const Job J = f->job;
        const size_t def_want = (size_t)c->n_def + (size_t)c->n_def / 8 + 1024;
        const size_t long_want = (size_t)c->long_used + (size_t)c->long_used / 8 + 4096;
        int rc = launch_shard(ctx, d, f, J.text, J.nbytes, (const uint64_t*)J.doc_off, J.ndocs, J.out, J.out_cap,
                              (uint64_t*)J.tok_off, st, def_want, long_want);
        if (rc != HUT_STATUS_OK) return rc;
    }
#ifdef HUT_LONGSTAT
    if (c->st_big)
        fprintf(stderr, "hut longstat: rounds at once %llu (prefix only %llu), merges applied at once %llu, stretches %llu\n",
                c->lp[0], c->lp[3], c->lp[2], c->lp[1]);
#endif
    float ms_all = 0, ms_main = 0;
    cudaEventElapsedTime(&ms_all, f->ev[0], f->ev[3]);
    cudaEventElapsedTime(&ms_main, f->ev[1], f->ev[2]);
    d->stats.pretokens += c->st_pretokens;
    d->stats.cache_hits += c->st_hits;
    d->stats.warp_merges += c->st_warp;
    d->stats.block_merges += c->st_block;
    d->stats.big_merges += c->st_big;
    d->stats.kernel_ms += ms_all;
    d->stats.main_kernel_ms += ms_main;
    uint64_t total = c->total_ids;
    if (c->flags & FLAG_OUT_CAP) return set_err(ctx, HUT_STATUS_CAPACITY, "%s", "output id buffer too small");
    if (c->flags & FLAG_BIG_CAP)
        return set_err(ctx, HUT_STATUS_UNSUPPORTED, "%s", "a pretoken exceeds the configured max_pretoken working area");
    if (c->flags & (FLAG_DEF_CAP | FLAG_LONG_CAP | FLAG_BIG_NEED))
        return set_errf(ctx, HUT_STATUS_MEMORY,
                       "side buffers for long pretokens still too small after 4 re-runs: %u deferred pretokens, %llu of their ids%s",
                       c->n_def, (unsigned long long)c->long_used,
                       (c->flags & FLAG_BIG_NEED) ? ", working area for pretokens beyond 4096 symbols missing" : "");
    if (c->flags & FLAG_EVENTS)
        return set_err(ctx, HUT_STATUS_UNSUPPORTED, "%s", "too many over-long pretokens in one call");
    if (c->flags & FLAG_NUL)
        return set_err(ctx, HUT_STATUS_VALUE, "%s", "document contains a 0x00 byte (cut documents at NUL before calling)");
    if (c->n_events) {
        // reference-compatible cut (core.c:402-407,503): drop the rest of each affected document.  The events sit in
        // this flight's own counters, so chunks queued behind it have not touched them.
        CU(ctx, grow(&d->tmp_ids, &d->tmp_ids_cap, (size_t)total + 1));
        CU(ctx, cudaMemcpyAsync(d->tmp_ids, d_ids, total * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
        cut_prepare_kernel<<<1, 1, 0, st>>>(f->job, d->cut);
        cut_apply_ids_kernel<<<592, 256, 0, st>>>(d->tmp_ids, d_ids, total, d->cut);
        cut_apply_offs_kernel<<<148, 256, 0, st>>>((unsigned long long*)d_tok_offsets, n_docs + 1, d->cut);
        CU(ctx, cudaGetLastError());
        CutRanges head;
        CU(ctx, cudaMemcpyAsync(&head, d->cut, offsetof(CutRanges, x), cudaMemcpyDeviceToHost, st));
        CU(ctx, cudaStreamSynchronize(st));
        total = head.new_total;
        d->stats.cut_documents += head.n;
        d->stats.kernel_launches += 3;
    }
    *n_ids_out = total;
    return HUT_STATUS_OK;
}

extern "C" int hut_encode_device(hut_ctx* ctx, int dev_index, const uint8_t* d_bytes, uint64_t n_bytes,
                                 const uint64_t* d_doc_offsets, uint64_t n_docs, int32_t* d_ids,
                                 uint64_t ids_capacity, uint64_t* d_tok_offsets, uint64_t* n_ids_out, void* stream) {
    if (!ctx) return HUT_STATUS_STATE;
    if (dev_index < 0 || dev_index >= ctx->ndev || !n_ids_out || !d_doc_offsets || !d_tok_offsets)
        return set_err(ctx, HUT_STATUS_ARG, "%s", "bad argument");
    hut_dev* d = &ctx->dev[dev_index];
    CU(ctx, cudaSetDevice(d->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : d->stream;
    stats_begin(d);
    hut_flight* f = &d->fl[0];
    int rc = launch_shard(ctx, d, f, d_bytes, n_bytes, d_doc_offsets, n_docs, d_ids, ids_capacity, d_tok_offsets, st, 0, 0);
    if (rc != HUT_STATUS_OK) return rc;
    return finish_shard(ctx, d, f, n_docs, d_ids, d_tok_offsets, n_ids_out, st);
}

template <class T>
static cudaError_t grow_host(T** p, size_t* cap, size_t need, bool keep) {
    if (need <= *cap) return cudaSuccess;
    size_t n = need + need / 4 + 4096;
    T* q = nullptr;
    cudaError_t e = cudaMallocHost(&q, n * sizeof(T));
    if (e != cudaSuccess) return e;
    if (*p) {
        if (keep) memcpy(q, *p, *cap * sizeof(T));
        cudaFreeHost(*p);
    }
    *p = q;
    *cap = n;
    return cudaSuccess;
}

extern "C" int hut_plan_shards(const uint64_t* doc_offsets, uint64_t n_docs, int n_shards, uint64_t* bounds) {
    if (!doc_offsets || !bounds || n_shards <= 0) return HUT_STATUS_ARG;
    const uint64_t b0 = doc_offsets[0], total = doc_offsets[n_docs] - b0;
    uint64_t d = 0;
    bounds[0] = 0;
    for (int s = 0; s < n_shards; ++s) {
        if (s == n_shards - 1) d = n_docs;
        else {
            const uint64_t target = b0 + (uint64_t)(((unsigned __int128)total * (uint64_t)(s + 1)) / (uint64_t)n_shards);
            uint64_t lo = d, hi = n_docs;  // first doc boundary >= target
            while (lo < hi) {
                const uint64_t mid = (lo + hi) >> 1;
                if (doc_offsets[mid] < target) lo = mid + 1;
                else hi = mid;
            }
            d = lo;
        }
        bounds[s + 1] = d;
    }
    return HUT_STATUS_OK;
}

/* One chunk of a host-buffer batch: documents [doc_lo, doc_hi) on device `dev`, buffers cb[buf] / fl[buf]. */
struct chunk_plan {
    uint64_t doc_lo, doc_hi, byte_lo, byte_hi, n_ids, id_base;
    int dev, buf;
    cudaEvent_t tr[6];  // HUTOKEN_B200_TRACE: upload, kernels, download (start / end each); null otherwise
};
#define TRACE(p, i, stream)                                      \
    do {                                                         \
        if ((p)->tr[i]) CU(ctx, cudaEventRecord((p)->tr[i], stream)); \
    } while (0)

static int drain_outputs(hut_ctx* ctx) {
    for (int i = 0; i < ctx->ndev; ++i) {
        CU(ctx, cudaSetDevice(ctx->dev[i].device));
        CU(ctx, cudaStreamSynchronize(ctx->dev[i].s_out));
    }
    return HUT_STATUS_OK;
}

/* upload + kernels of one chunk (asynchronous) */
static int chunk_issue(hut_ctx* ctx, const uint8_t* bytes, const uint64_t* doc_offsets, chunk_plan* p) {
    hut_dev* d = &ctx->dev[p->dev];
    hut_chunkbuf* c = &d->cb[p->buf];
    const uint64_t nb = p->byte_hi - p->byte_lo, nd = p->doc_hi - p->doc_lo;
    CU(ctx, cudaSetDevice(d->device));
    if (c->out_pending) {  // the ids of the chunk that used these buffers before are still going to the host
        CU(ctx, cudaStreamWaitEvent(d->stream, c->ev_out, 0));
        c->out_pending = 0;
    }
    // worst case: every symbol survives as its own id (+ prefix ids per document)
    const uint64_t worst = nb * d->dm.kmax + nd * (uint64_t)(d->dm.n_prefix_syms + d->dm.n_prefix_alone) + 64;
    if (nb + 64 > c->d_text_cap || nd + 1 > c->d_off_cap || nd + 1 > c->d_tok_cap || worst > c->d_ids_cap) {
        CU(ctx, cudaDeviceSynchronize());  // growing frees buffers that queued work may still use
        CU(ctx, grow(&c->d_text, &c->d_text_cap, (size_t)nb + 64));
        CU(ctx, grow(&c->d_off, &c->d_off_cap, (size_t)nd + 1));
        CU(ctx, grow(&c->d_tok, &c->d_tok_cap, (size_t)nd + 1));
        CU(ctx, grow(&c->d_ids, &c->d_ids_cap, (size_t)worst));
    }
    if (nd + 1 > c->h_off_cap) {
        CU(ctx, cudaStreamSynchronize(d->s_in));
        if (c->h_off) cudaFreeHost(c->h_off);
        c->h_off = nullptr;
        c->h_off_cap = 0;
        const size_t n = (size_t)nd + 1 + nd / 4 + 1024;
        CU(ctx, cudaMallocHost(&c->h_off, n * sizeof(uint64_t)));
        c->h_off_cap = n;
    }
    for (uint64_t i = 0; i <= nd; ++i) c->h_off[i] = doc_offsets[p->doc_lo + i] - p->byte_lo;
    TRACE(p, 0, d->s_in);
    if (nb) CU(ctx, cudaMemcpyAsync(c->d_text, bytes + p->byte_lo, nb, cudaMemcpyHostToDevice, d->s_in));
    CU(ctx, cudaMemcpyAsync(c->d_off, c->h_off, (nd + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, d->s_in));
    CU(ctx, cudaEventRecord(c->ev_in, d->s_in));
    TRACE(p, 1, d->s_in);
    CU(ctx, cudaStreamWaitEvent(d->stream, c->ev_in, 0));
    TRACE(p, 2, d->stream);
    const int rc = launch_shard(ctx, d, &d->fl[p->buf], c->d_text, nb, (const uint64_t*)c->d_off, nd, c->d_ids,
                                c->d_ids_cap, (uint64_t*)c->d_tok, d->stream, 0, 0);
    if (rc == HUT_STATUS_OK) TRACE(p, 3, d->stream);
    return rc;
}

/* waits for the chunk's kernels, then starts the download of its ids and offsets into the final pinned buffers */
static int chunk_complete(hut_ctx* ctx, chunk_plan* p, uint64_t id_base, uint64_t bytes_done, uint64_t bytes_total) {
    hut_dev* d = &ctx->dev[p->dev];
    hut_chunkbuf* c = &d->cb[p->buf];
    const uint64_t nd = p->doc_hi - p->doc_lo;
    CU(ctx, cudaSetDevice(d->device));
    int rc = finish_shard(ctx, d, &d->fl[p->buf], nd, c->d_ids, (uint64_t*)c->d_tok, &p->n_ids, d->stream);
    if (rc != HUT_STATUS_OK) return rc;
    p->id_base = id_base;
    if (id_base + p->n_ids + 1 > ctx->h_ids_cap) {  // grow the pinned result, keeping what has arrived
        rc = drain_outputs(ctx);
        if (rc != HUT_STATUS_OK) return rc;
        const double per_byte = (double)(id_base + p->n_ids) / (double)(bytes_done ? bytes_done : 1);
        size_t need = (size_t)(per_byte * 1.05 * (double)bytes_total) + 4096;
        if (need < id_base + p->n_ids + 1) need = id_base + p->n_ids + 1;
        int32_t* q = nullptr;
        CU(ctx, cudaMallocHost(&q, need * sizeof(int32_t)));
        if (ctx->h_ids) {
            if (id_base) memcpy(q, ctx->h_ids, id_base * sizeof(int32_t));
            cudaFreeHost(ctx->h_ids);
        }
        ctx->h_ids = q;
        ctx->h_ids_cap = need;
        CU(ctx, cudaSetDevice(d->device));
    }
    TRACE(p, 4, d->s_out);
    if (p->n_ids)
        CU(ctx, cudaMemcpyAsync(ctx->h_ids + id_base, c->d_ids, p->n_ids * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                d->s_out));
    // nd entries, not nd + 1: entry doc_hi belongs to the next chunk (two streams must not write one word)
    if (nd)
        CU(ctx, cudaMemcpyAsync(ctx->h_tok + p->doc_lo, c->d_tok, nd * sizeof(uint64_t), cudaMemcpyDeviceToHost,
                                d->s_out));
    CU(ctx, cudaEventRecord(c->ev_out, d->s_out));
    TRACE(p, 5, d->s_out);
    c->out_pending = 1;
    return HUT_STATUS_OK;
}

/* ---- small batches: one launch of one CTA, no copies (small_kernel in hut_pipeline.cuh) -------------------------- */
#define SMALL_MAX_BYTES ((uint64_t)SMALL_MAX_TILES * OWN)
#define SMALL_MAX_DOCS 2048u
#define HUT_SMALL_RETRY (-1) /* the general path has to do it */

static int encode_small(hut_ctx* ctx, const uint8_t* bytes, const uint64_t* doc_offsets, uint64_t n_docs,
                        hut_result* result) {
    hut_dev* d = &ctx->dev[0];
    hut_flight* f = &d->fl[0];
    const uint64_t b0 = doc_offsets[0], nb = doc_offsets[n_docs] - b0;
    CU(ctx, cudaSetDevice(d->device));
    if (!d->sm_text) {
        CU(ctx, cudaMallocHost(&d->sm_text, SMALL_MAX_BYTES + 256));
        CU(ctx, cudaMallocHost(&d->sm_off, (SMALL_MAX_DOCS + 2) * sizeof(unsigned long long)));
        CU(ctx, cudaMallocHost(&d->sm_ctr, offsetof(Counters, events)));
        CU(ctx, cudaMallocHost((void**)&d->sm_done, 64));
        *d->sm_done = 0;
        d->sm_seq = 0;
    }
    const uint32_t ntiles = (uint32_t)((nb + OWN - 1) / OWN);
    const uint64_t worst = nb * d->dm.kmax + n_docs * (uint64_t)(d->dm.n_prefix_syms + d->dm.n_prefix_alone) + 64;
    {
        cudaError_t e = grow_host(&ctx->h_ids, &ctx->h_ids_cap, (size_t)worst, false);
        if (e == cudaSuccess) e = grow_host(&ctx->h_tok, &ctx->h_tok_cap, (size_t)n_docs + 1, false);
        if (e != cudaSuccess) return set_err(ctx, HUT_STATUS_CUDA, "CUDA error: %s", cudaGetErrorString(e));
    }
    CU(ctx, grow(&d->stage, &d->stage_cap, (size_t)ntiles * d->stride + (size_t)n_docs * d->slot_extra + 64));
    CU(ctx, grow(&f->tilebuf, &f->tilebuf_cap, 6 * ((size_t)ntiles + 1) + 4));
    CU(ctx, grow(&f->bsum, &f->bsum_cap, 1));
    CU(ctx, grow(&f->def, &f->def_cap, 1024));
    memcpy(d->sm_text, bytes + b0, nb);
    memset(d->sm_text + nb, 0, 128);
    for (uint64_t i = 0; i <= n_docs; ++i) d->sm_off[i] = doc_offsets[i] - b0;
    Job J;
    memset(&J, 0, sizeof J);
    J.text = d->sm_text;
    J.nbytes = nb;
    J.doc_off = d->sm_off;
    J.ndocs = n_docs;
    J.out = ctx->h_ids;
    J.out_cap = ctx->h_ids_cap;
    J.tok_off = ctx->h_tok;
    J.ntiles = ntiles;
    J.stride = d->stride;
    J.slot_extra = d->slot_extra;
    J.slot_off = reinterpret_cast<unsigned long long*>(f->tilebuf + 4 * ((size_t)ntiles + 1));
    J.stage = d->stage;
    J.cnt = f->tilebuf;
    J.tot = J.cnt + ((size_t)ntiles + 1);
    J.first_doc = J.tot + ((size_t)ntiles + 1);
    J.def_first = J.first_doc + ((size_t)ntiles + 1);
    J.bsum = f->bsum;
    J.nblocks = 1;
    J.def = f->def;
    J.def_cap = (uint32_t)f->def_cap;
    J.ctr = f->ctr;
    J.cache = d->cache;
    J.cache_bits = d->cache_bits;
    J.cache2 = d->cache2;
    J.cache2_bits = d->cache2_bits;
    const unsigned int seq = ++d->sm_seq;
    if (d->dm.has_prefix) small_kernel<true><<<1, NT, 0, d->stream>>>(d->dm, J, d->sm_ctr, d->sm_done, seq);
    else small_kernel<false><<<1, NT, 0, d->stream>>>(d->dm, J, d->sm_ctr, d->sm_done, seq);
    CU(ctx, cudaGetLastError());
    // the kernel writes `done` last (after a system-wide fence): spinning on it is cheaper than a stream synchronize
    for (unsigned spins = 0; *d->sm_done != seq; ++spins) {
        if ((spins & 0xFFFFu) == 0xFFFFu) {
            cudaError_t e = cudaStreamQuery(d->stream);
            if (e != cudaSuccess && e != cudaErrorNotReady)
                return set_err(ctx, HUT_STATUS_CUDA, "CUDA error: %s", cudaGetErrorString(e));
            if (e == cudaSuccess && *d->sm_done != seq)
                return set_err(ctx, HUT_STATUS_CUDA, "%s", "small_kernel ended without reporting");
        }
    }
    const Counters* c = d->sm_ctr;
    if (c->flags & SMALL_FALLBACK) return HUT_SMALL_RETRY;
    stats_begin(d);
    d->stats.pretokens = c->st_pretokens;
    d->stats.cache_hits = c->st_hits;
    d->stats.warp_merges = c->st_warp;
    d->stats.kernel_launches = 1;
    if (c->flags & FLAG_NUL)
        return set_err(ctx, HUT_STATUS_VALUE, "%s", "document contains a 0x00 byte (cut documents at NUL before calling)");
    result->ids = ctx->h_ids;
    result->tok_offsets = (const uint64_t*)ctx->h_tok;
    result->n_ids = c->total_ids;
    result->n_docs = n_docs;
    return HUT_STATUS_OK;
}

extern "C" int hut_encode_batch(hut_ctx* ctx, const uint8_t* bytes, const uint64_t* doc_offsets, uint64_t n_docs,
                                hut_result* result) {
    if (!ctx) return HUT_STATUS_STATE;
    if (!result || !doc_offsets || (n_docs && !bytes && doc_offsets[n_docs] != doc_offsets[0]))
        return set_err(ctx, HUT_STATUS_ARG, "%s", "bad argument");
    memset(result, 0, sizeof *result);
    for (uint64_t i = 0; i < n_docs; ++i)  // decreasing offsets would underflow every byte count below
        if (doc_offsets[i] > doc_offsets[i + 1]) return set_err(ctx, HUT_STATUS_ARG, "%s", "doc_offsets must be non-decreasing");
    const uint64_t total_bytes = doc_offsets[n_docs] - doc_offsets[0];
    if (total_bytes && total_bytes <= SMALL_MAX_BYTES && n_docs <= SMALL_MAX_DOCS && !getenv("HUTOKEN_B200_NO_SMALL")) {
        const int rc = encode_small(ctx, bytes, doc_offsets, n_docs, result);
        if (rc != HUT_SMALL_RETRY) return rc;
    }
    // ---- plan: contiguous document ranges balanced by bytes (hut_plan_shards), dealt round-robin to the devices;
    //      every device works on two chunks at a time: upload of chunk k+1 | kernels of chunk k | download of k-1 ----
    const int ndev = ctx->ndev;
    uint64_t chunk_bytes = (uint64_t)(ctx->opt.chunk_mb ? ctx->opt.chunk_mb : 32u) << 20;
    uint64_t nch64 = (total_bytes + chunk_bytes - 1) / chunk_bytes;
    if (nch64 < (uint64_t)ndev) nch64 = (uint64_t)ndev;
    if (nch64 > n_docs) nch64 = n_docs ? n_docs : 1;
    if (nch64 > (1u << 20)) nch64 = 1u << 20;
    const int nch = (int)nch64;
    uint64_t* bounds = (uint64_t*)malloc(((size_t)nch + 1) * sizeof(uint64_t));
    chunk_plan* plan = (chunk_plan*)calloc((size_t)nch, sizeof(chunk_plan));
    if (!bounds || !plan) {
        free(bounds);
        free(plan);
        return set_err(ctx, HUT_STATUS_MEMORY, "%s", "out of memory");
    }
    hut_plan_shards(doc_offsets, n_docs, nch, bounds);
    for (int k = 0; k < nch; ++k) {
        plan[k].doc_lo = bounds[k];
        plan[k].doc_hi = bounds[k + 1];
        plan[k].byte_lo = doc_offsets[plan[k].doc_lo];
        plan[k].byte_hi = doc_offsets[plan[k].doc_hi];
        plan[k].dev = k % ndev;
        plan[k].buf = (k / ndev) & 1;
    }
    free(bounds);
    const bool trace = getenv("HUTOKEN_B200_TRACE") != nullptr;  // per-chunk timeline on stderr (profiles/README.md)
    if (trace)
        for (int k = 0; k < nch; ++k) {
            cudaSetDevice(ctx->dev[plan[k].dev].device);
            for (int i = 0; i < 6; ++i) cudaEventCreate(&plan[k].tr[i]);
        }
    int rc = HUT_STATUS_OK;
    auto bail = [&](int st) {
        for (int i = 0; i < ctx->ndev; ++i) {
            cudaSetDevice(ctx->dev[i].device);
            cudaDeviceSynchronize();
            ctx->dev[i].cb[0].out_pending = ctx->dev[i].cb[1].out_pending = 0;
        }
        if (trace)
            for (int k = 0; k < nch; ++k)
                for (int i = 0; i < 6; ++i)
                    if (plan[k].tr[i]) cudaEventDestroy(plan[k].tr[i]);
        free(plan);
        return st;
    };
    {
        cudaError_t e = grow_host(&ctx->h_tok, &ctx->h_tok_cap, (size_t)n_docs + 1, false);
        if (e != cudaSuccess) return bail(set_err(ctx, HUT_STATUS_CUDA, "CUDA error: %s", cudaGetErrorString(e)));
    }
    for (int i = 0; i < ndev; ++i) stats_begin(&ctx->dev[i]);
    // device buffers of every (device, buffer) pair sized once for the largest chunk it will see, before anything is
    // queued: growing inside the loop would have to drain the device first (chunk_issue keeps that path as a net)
    for (int i = 0; i < ndev; ++i)
        for (int b = 0; b < 2; ++b) {
            hut_dev* d = &ctx->dev[i];
            hut_chunkbuf* c = &d->cb[b];
            uint64_t nb = 0, nd = 0, worst = 0;
            for (int k = 0; k < nch; ++k) {
                if (plan[k].dev != i || plan[k].buf != b) continue;
                const uint64_t kb = plan[k].byte_hi - plan[k].byte_lo, kd = plan[k].doc_hi - plan[k].doc_lo;
                const uint64_t kw = kb * d->dm.kmax + kd * (uint64_t)(d->dm.n_prefix_syms + d->dm.n_prefix_alone) + 64;
                nb = kb > nb ? kb : nb;
                nd = kd > nd ? kd : nd;
                worst = kw > worst ? kw : worst;
            }
            if (!worst) continue;
            if (nb + 64 > c->d_text_cap || nd + 1 > c->d_off_cap || nd + 1 > c->d_tok_cap || worst > c->d_ids_cap) {
                cudaError_t e = cudaSetDevice(d->device);
                if (e == cudaSuccess) e = cudaDeviceSynchronize();
                if (e == cudaSuccess) e = grow(&c->d_text, &c->d_text_cap, (size_t)nb + 64);
                if (e == cudaSuccess) e = grow(&c->d_off, &c->d_off_cap, (size_t)nd + 1);
                if (e == cudaSuccess) e = grow(&c->d_tok, &c->d_tok_cap, (size_t)nd + 1);
                if (e == cudaSuccess) e = grow(&c->d_ids, &c->d_ids_cap, (size_t)worst);
                if (e != cudaSuccess) return bail(set_errf(ctx, e == cudaErrorMemoryAllocation ? HUT_STATUS_MEMORY : HUT_STATUS_CUDA,
                                                          "chunk buffers of %llu text bytes / %llu ids: %s",
                                                          (unsigned long long)nb, (unsigned long long)worst, cudaGetErrorString(e)));
            }
        }
    const int depth = 2 * ndev;
    uint64_t id_base = 0, bytes_done = 0;
    for (int k = 0; k < nch + depth; ++k) {
        if (k >= depth) {
            chunk_plan* p = &plan[k - depth];
            bytes_done += p->byte_hi - p->byte_lo;
            rc = chunk_complete(ctx, p, id_base, bytes_done, total_bytes);
            if (rc != HUT_STATUS_OK) return bail(rc);
            id_base += p->n_ids;
        }
        if (k < nch) {
            rc = chunk_issue(ctx, bytes, doc_offsets, &plan[k]);
            if (rc != HUT_STATUS_OK) return bail(rc);
        }
    }
    rc = drain_outputs(ctx);
    if (rc != HUT_STATUS_OK) return bail(rc);
    for (int i = 0; i < ndev; ++i) ctx->dev[i].cb[0].out_pending = ctx->dev[i].cb[1].out_pending = 0;
    if (trace) {
        for (int k = 0; k < nch; ++k) {
            float t[6] = {0, 0, 0, 0, 0, 0};
            // times are relative to the device's first chunk: chunk k runs on device k % ndev, so chunk `dev` is it
            for (int i = 0; i < 6; ++i) cudaEventElapsedTime(&t[i], plan[plan[k].dev].tr[0], plan[k].tr[i]);
            fprintf(stderr, "hut trace chunk %d dev %d MB %.1f ids %llu | up %.3f-%.3f | kernels %.3f-%.3f | down %.3f-%.3f ms\n",
                    k, plan[k].dev, (double)(plan[k].byte_hi - plan[k].byte_lo) / 1048576.0,
                    (unsigned long long)plan[k].n_ids, t[0], t[1], t[2], t[3], t[4], t[5]);
        }
        for (int k = 0; k < nch; ++k)
            for (int i = 0; i < 6; ++i) cudaEventDestroy(plan[k].tr[i]);
    }
    // per-chunk offsets are local: add each chunk's base (host side, one add per document -- 8(e))
    for (int k = 0; k < nch; ++k)
        if (plan[k].id_base)
            for (uint64_t i = plan[k].doc_lo; i < plan[k].doc_hi; ++i) ctx->h_tok[i] += plan[k].id_base;
    ctx->h_tok[n_docs] = id_base;
    if (!ctx->h_ids) {  // nothing was produced: still hand out a valid pointer
        cudaError_t e = cudaMallocHost(&ctx->h_ids, 64);
        if (e != cudaSuccess) return bail(set_err(ctx, HUT_STATUS_CUDA, "CUDA error: %s", cudaGetErrorString(e)));
        ctx->h_ids_cap = 16;
    }
    free(plan);
    result->ids = ctx->h_ids;
    result->tok_offsets = (const uint64_t*)ctx->h_tok;
    result->n_ids = id_base;
    result->n_docs = n_docs;
    return HUT_STATUS_OK;
}

/* ---- scratch of hut_decode_device_noalloc (below): a file-scope device array instead of cudaMalloc'ed buffers.
 * Every GPU that loads the library holds these HUT_NOALLOC_SCRATCH_BYTES (about 1 MiB) until the process exits.
 * In host code `g_noalloc_scratch` is only a stand-in for the symbol: the address a kernel may use comes from
 * cudaGetSymbolAddress, asked once per context with the context's decode device current (noalloc_setup). ---- */
constexpr unsigned long long NA_MAX_IDS = HUT_NOALLOC_MAX_IDS;
constexpr size_t NA_MAX_SUB = (size_t)(NA_MAX_IDS / DEC_SUB);
constexpr size_t na_align(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t NA_OFF_BAD = 0;                                                             /* 4 x unsigned int */
constexpr size_t NA_OFF_SUB_LEN = na_align(NA_OFF_BAD + 16);                                 /* uint32[nsub + 1] */
constexpr size_t NA_OFF_SUB_OFF = na_align(NA_OFF_SUB_LEN + 4 * (NA_MAX_SUB + 1));           /* uint64[nsub + 2] */
constexpr size_t NA_OFF_FIRST_DOC = na_align(NA_OFF_SUB_OFF + 8 * (NA_MAX_SUB + 2));         /* uint32[n_ids / 256 + 1] */
constexpr size_t NA_OFF_BSUM = na_align(NA_OFF_FIRST_DOC + 4 * (NA_MAX_SUB + 1));            /* uint64[scan blocks] */
constexpr size_t NA_BYTES = na_align(NA_OFF_BSUM + 8 * (NA_MAX_SUB / SCAN_BLOCK + 2));
static_assert(NA_BYTES <= HUT_NOALLOC_SCRATCH_BYTES, "scratch layout larger than the size the header states");

static __device__ __align__(256) unsigned char g_noalloc_scratch[HUT_NOALLOC_SCRATCH_BYTES];

/* decode set-up (dec_init_tables), device of the context current: device address of the array on THAT device */
static int noalloc_setup(hut_ctx* ctx, unsigned char** out) {
    void* p = nullptr;
    CU(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    *out = (unsigned char*)p;
    return HUT_STATUS_OK;
}

#include "hut_decode_api.inl"
