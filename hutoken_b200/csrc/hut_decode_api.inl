/* hut_decode_batch / hut_decode_device: host orchestration of the decode kernels (hut_decode.cuh).
 * Included at the end of hut_engine.cu (needs hut_ctx / hut_dev).  Device 0 of the context does the work of the
 * device-resident calls and of small host batches; a large host batch is streamed in document chunks dealt round-robin
 * to all devices of the context (decode_chunked), each of which holds its own decode tables. */

struct hut_decstate {
    DecModel dm;
    int ready;
    int dev;  // index into ctx->dev of the device this state lives on
    int32_t* d_ids;
    size_t d_ids_cap;
    unsigned long long* d_tok;
    size_t d_tok_cap;
    uint32_t* d_sub_len;          // token bytes per sub-block of DEC_SUB ids
    size_t d_sub_len_cap;
    unsigned long long* d_sub_off;  // their exclusive scan
    size_t d_sub_off_cap;
    unsigned long long* d_doc_raw;  // raw offset of every document
    size_t d_doc_raw_cap;
    unsigned long long* d_bsum;
    size_t d_bsum_cap;
    uint8_t* d_raw;
    size_t d_raw_cap;
    uint8_t* d_dec;
    size_t d_dec_cap;
    uint32_t* d_doc_len;
    size_t d_doc_len_cap;
    unsigned long long* d_text_off;
    size_t d_text_off_cap;
    uint8_t* d_text;
    size_t d_text_cap;
    unsigned int* d_bad;
    uint32_t* d_first_doc;  /* token-wise path: first document that starts in each sub-block */
    size_t d_first_doc_cap;
    uint8_t* h_text;
    size_t h_text_cap;
    unsigned long long* h_text_off;
    size_t h_text_off_cap;
    cudaEvent_t ev[2];
    float last_ms;
    unsigned last_launches;
    /* one-launch path (dec_small_kernel): pinned host buffers the kernel reads and writes directly */
    int32_t* ds_ids;
    unsigned long long* ds_tok;
    uint8_t* ds_text;
    unsigned long long* ds_toff;
    DecSmallHdr* ds_hdr;
    volatile unsigned int* ds_done;
    unsigned int ds_seq;
    /* large host batches: document chunks streamed through two sets of input buffers (decode_chunked) */
    int32_t* c_ids[2];
    size_t c_ids_cap[2];
    unsigned long long* c_tok[2];
    size_t c_tok_cap[2];
    unsigned long long* h_ctok[2]; /* pinned staging of a chunk's rebased offsets */
    size_t h_ctok_cap[2];
    cudaEvent_t ev_up[2], ev_k[2], ev_dl[2]; /* upload done / kernels done / download done, per buffer set */
    uint8_t* x_text;           /* second set of output buffers: odd chunks decode into these while the even chunk's */
    size_t x_text_cap;         /* text is still on its way to the host */
    unsigned long long* x_text_off;
    size_t x_text_off_cap;
    unsigned char* na_scratch; /* device address of g_noalloc_scratch on this context's decode device */
    int token_wise;  /* the per-token tables exist and HUTOKEN_B200_DECODE=general is not set */
    int last_path;   /* 1: token-wise kernels, 2: general kernels */
};

static void dec_free(hut_decstate* s) {
    if (!s) return;
    cudaFree(s->d_ids);
    cudaFree(s->d_tok);
    cudaFree(s->d_sub_len);
    cudaFree(s->d_sub_off);
    cudaFree(s->d_doc_raw);
    cudaFree(s->d_bsum);
    cudaFree(s->d_raw);
    cudaFree(s->d_dec);
    cudaFree(s->d_doc_len);
    cudaFree(s->d_text_off);
    cudaFree(s->d_text);
    cudaFree(s->d_bad);
    cudaFree(s->d_first_doc);
    cudaFreeHost(s->h_text);
    cudaFreeHost(s->ds_ids);
    cudaFreeHost(s->ds_tok);
    cudaFreeHost(s->ds_text);
    cudaFreeHost(s->ds_toff);
    cudaFreeHost(s->ds_hdr);
    cudaFreeHost((void*)s->ds_done);
    cudaFreeHost(s->h_text_off);
    for (int b = 0; b < 2; ++b) {
        cudaFree(s->c_ids[b]);
        cudaFree(s->c_tok[b]);
        cudaFreeHost(s->h_ctok[b]);
        if (s->ev_up[b]) cudaEventDestroy(s->ev_up[b]);
        if (s->ev_k[b]) cudaEventDestroy(s->ev_k[b]);
        if (s->ev_dl[b]) cudaEventDestroy(s->ev_dl[b]);
    }
    cudaFree(s->x_text);
    cudaFree(s->x_text_off);
    if (s->ev[0]) cudaEventDestroy(s->ev[0]);
    if (s->ev[1]) cudaEventDestroy(s->ev[1]);
    free(s);
}

static int dec_init_tables(hut_ctx* ctx, hut_decstate* s);
static int dec_init_token_tables(hut_ctx* ctx, hut_decstate* s);

/* decode-side device state, created by the first hut_decode_* call.  A failure half way leaves nothing behind: the
 * tables uploaded so far are released (they are the newest entries of the device's allocation list) and the next call
 * starts again from scratch. */
static int dec_init_on(hut_ctx* ctx, hut_decstate** slot, int dev) {
    if (*slot && (*slot)->ready) return HUT_STATUS_OK;
    hut_dev* d = &ctx->dev[dev];
    CU(ctx, cudaSetDevice(d->device));
    if (!*slot) *slot = (hut_decstate*)calloc(1, sizeof(hut_decstate));
    if (!*slot) return set_err(ctx, HUT_STATUS_MEMORY, "%s", "out of memory");
    hut_decstate* s = *slot;
    s->dev = dev;
    const int first = d->nalloc;
    const int rc = dec_init_tables(ctx, s);
    if (rc != HUT_STATUS_OK) {
        while (d->nalloc > first) cudaFree(d->allocs[--d->nalloc]);
        cudaFree(s->d_bad);
        if (s->ev[0]) cudaEventDestroy(s->ev[0]);
        if (s->ev[1]) cudaEventDestroy(s->ev[1]);
        memset(s, 0, sizeof *s);
        s->dev = dev;
    }
    return rc;
}
static int dec_init(hut_ctx* ctx) { return dec_init_on(ctx, &ctx->dec, 0); }

/* Decoded bytes of every token (hut_decode.cuh, token-wise path): lengths and flags first, offsets by a host scan,
 * then the bytes; the 16-byte slots and the one-byte lengths are put together on the host. */
static int dec_init_token_tables(hut_ctx* ctx, hut_decstate* s) {  // on device s->dev (current)
    hut_dev* d = &ctx->dev[s->dev];
    DecModel& D = s->dm;
    const uint32_t V = D.vocab_lines;
    uint8_t* kids = nullptr;
    CU(ctx, dev_alloc(d, (void**)&kids, (size_t)D.trie_nodes + 1));
    CU(ctx, cudaMemset(kids, 0, (size_t)D.trie_nodes + 1));
    if (D.trie_mask != 0xFFFFFFFFu) dec2_kids_kernel<<<64, 256, 0, d->stream>>>(D.trie_edge, D.trie_mask + 1, kids);
    D.trie_kids = kids;
    uint32_t *meta = nullptr, *off = nullptr;
    CU(ctx, dev_alloc(d, (void**)&meta, ((size_t)V + 1) * sizeof(uint32_t)));
    CU(ctx, dev_alloc(d, (void**)&off, ((size_t)V + 1) * sizeof(uint32_t)));
    const unsigned grid = (V + 127) / 128;
    if (V) dec2_token_kernel<false><<<grid, 128, 0, d->stream>>>(D, kids, meta, nullptr, nullptr);
    std::vector<uint32_t> h_meta((size_t)V + 1), h_off((size_t)V + 1);
    CU(ctx, cudaMemcpyAsync(h_meta.data(), meta, (size_t)V * sizeof(uint32_t), cudaMemcpyDeviceToHost, d->stream));
    CU(ctx, cudaStreamSynchronize(d->stream));
    unsigned long long total = 0;
    for (uint32_t t = 0; t < V; ++t) {
        h_off[t] = (uint32_t)total;
        total += h_meta[t] & T_LEN_MASK;
    }
    h_off[V] = (uint32_t)total;
    if (total > 0xFFFFFF00ull) return set_err(ctx, HUT_STATUS_MEMORY, "%s", "decoded vocabulary too large");
    uint8_t* pool = nullptr;
    CU(ctx, dev_alloc(d, (void**)&pool, (size_t)total + 16));
    CU(ctx, cudaMemsetAsync(pool, 0, (size_t)total + 16, d->stream));
    CU(ctx, cudaMemcpyAsync(off, h_off.data(), ((size_t)V + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, d->stream));
    if (V) dec2_token_kernel<true><<<grid, 128, 0, d->stream>>>(D, kids, nullptr, off, pool);
    std::vector<uint8_t> h_pool((size_t)total + 16), h_len8((size_t)V + 1), h_slot(((size_t)V + 1) * 16);
    CU(ctx, cudaMemcpyAsync(h_pool.data(), pool, (size_t)total + 16, cudaMemcpyDeviceToHost, d->stream));
    CU(ctx, cudaStreamSynchronize(d->stream));
    CU(ctx, cudaGetLastError());
    for (uint32_t t = 0; t < V; ++t) {
        const uint32_t m = h_meta[t], len = m & T_LEN_MASK;
        h_len8[t] = (uint8_t)((m & (T_MISSING | T_UNSAFE)) || len >= 255u ? 255u : len);
        uint8_t* slot = &h_slot[(size_t)t * 16];
        for (uint32_t k = 0; k < 15 && k < len; ++k) slot[k] = h_pool[h_off[t] + k];
        slot[15] = (uint8_t)(len <= 15u ? len : 255u);
    }
    CU(ctx, upload(d, (const uint8_t*)h_len8.data(), (size_t)V + 1, &D.t_len8));
    CU(ctx, upload(d, reinterpret_cast<const uint4*>(h_slot.data()), (size_t)V + 1, &D.t_slot));
    D.t_meta = meta;
    D.t_off = off;
    D.t_pool = pool;
    return HUT_STATUS_OK;
}

static int dec_init_tables(hut_ctx* ctx, hut_decstate* s) {
    hut_dev* d = &ctx->dev[s->dev];
    const hut_model* m = ctx->model;
    DecModel& D = s->dm;
    CU(ctx, upload(d, m->dec_off, (size_t)m->vocab_lines + 1, &D.dec_off));
    CU(ctx, upload(d, m->dec_len, (size_t)m->vocab_lines, &D.dec_len));
    CU(ctx, upload(d, m->dec_pool, (size_t)m->dec_pool_len + 16, &D.dec_pool));
    D.vocab_lines = m->vocab_lines;
    CU(ctx, upload(d, (const unsigned long long*)m->trie_edge, m->trie_cap, &D.trie_edge));
    D.trie_mask = m->trie_cap - 1;
    D.trie_nodes = m->trie_nodes;
    CU(ctx, upload(d, m->trie_out, m->trie_nodes, &D.trie_out));
    CU(ctx, upload(d, m->trie_depth, m->trie_nodes, &D.trie_depth));
    CU(ctx, upload(d, m->trie_root_child, 256, &D.root_child));
    CU(ctx, upload(d, m->prefix_bytes, m->prefix_len, &D.prefix));
    D.prefix_len = m->prefix_len;
    D.byte_mode = m->is_byte_encoder;
    CU(ctx, cudaMalloc(&s->d_bad, 4 * sizeof(unsigned int)));
    CU(ctx, cudaEventCreate(&s->ev[0]));
    CU(ctx, cudaEventCreate(&s->ev[1]));
    const int rc = dec_init_token_tables(ctx, s);
    if (rc != HUT_STATUS_OK) return rc;
    const int rc_na = noalloc_setup(ctx, &s->na_scratch);
    if (rc_na != HUT_STATUS_OK) return rc_na;
    const char* mode = getenv("HUTOKEN_B200_DECODE"); /* "general": keep to the general kernels (tests compare the two) */
    s->token_wise = !(mode && !strcmp(mode, "general"));
    s->ready = 1;
    return HUT_STATUS_OK;
}

/* exclusive scan uint32[n] -> uint64[n + 1] on `st` */
static int dec_scan(hut_ctx* ctx, hut_decstate* s, const uint32_t* in, unsigned long long n, unsigned long long* out,
                    cudaStream_t st) {
    const unsigned long long nb = n / SCAN_BLOCK + 1;
    if (nb > 0x7FFFFFFFull) return set_err(ctx, HUT_STATUS_ARG, "%s", "too many items");
    CU(ctx, grow(&s->d_bsum, &s->d_bsum_cap, (size_t)nb));
    scan_sums_kernel<<<(unsigned)nb, SCAN_NT, 0, st>>>(in, n, s->d_bsum);
    scan_top_kernel<<<1, 1024, 0, st>>>(s->d_bsum, (uint32_t)nb);
    scan_apply_kernel<<<(unsigned)nb, SCAN_NT, 0, st>>>(in, n, s->d_bsum, out);
    s->last_launches += 3;
    return HUT_STATUS_OK;
}

/* ids + tok_off resident on device 0 -> packed text + text_off on the device (buffers owned by the context).
 * *n_text receives the text length.  Synchronises the stream twice (raw length, text length). */
static int dec_run(hut_ctx* ctx, hut_decstate* s, const int32_t* d_ids, unsigned long long n_ids,
                   const unsigned long long* d_tok, unsigned long long n_docs, unsigned long long* n_text, cudaStream_t st) {
    hut_dev* d = &ctx->dev[s->dev];
    const DecModel& D = s->dm;
    s->last_launches = 0;
    const unsigned long long nsub = (n_ids + DEC_SUB - 1) / DEC_SUB;
    if (n_ids / DEC_SUB + 1 > 0x7FFFFFFFull) return set_err(ctx, HUT_STATUS_ARG, "%s", "too many ids in one call");
    CU(ctx, grow(&s->d_sub_len, &s->d_sub_len_cap, (size_t)nsub + 1));
    CU(ctx, grow(&s->d_sub_off, &s->d_sub_off_cap, (size_t)nsub + 2));
    CU(ctx, grow(&s->d_doc_raw, &s->d_doc_raw_cap, (size_t)n_docs + 1));
    CU(ctx, grow(&s->d_doc_len, &s->d_doc_len_cap, (size_t)n_docs + 1));
    CU(ctx, grow(&s->d_text_off, &s->d_text_off_cap, (size_t)n_docs + 2));
    CU(ctx, cudaEventRecord(s->ev[0], st));
    CU(ctx, cudaMemsetAsync(s->d_bad, 0, 4 * sizeof(unsigned int), st));
    if (s->token_wise && n_ids && n_docs && n_docs < 0xFFFFFFF0ull) {
        /* token-wise kernels: lengths -> scan -> bytes and text_off in one pass; one synchronisation */
        const bool pfx = D.prefix_len != 0;
        const unsigned g1 = (unsigned)((nsub + 7) / 8);
        const size_t nfirst = (size_t)(n_ids / DEC_SUB + 1);
        CU(ctx, grow(&s->d_first_doc, &s->d_first_doc_cap, nfirst));
        CU(ctx, cudaMemsetAsync(s->d_first_doc, 0xFF, nfirst * sizeof(uint32_t), st));
        dec2_first_doc_kernel<<<(unsigned)(n_docs / 256 + 1), 256, 0, st>>>(d_tok, n_docs, s->d_first_doc);
        if (pfx)
            dec2_sub_kernel<true><<<g1, 256, 0, st>>>(D, d_ids, n_ids, d_tok, n_docs, s->d_first_doc, s->d_sub_len, s->d_bad);
        else
            dec2_sub_kernel<false><<<g1, 256, 0, st>>>(D, d_ids, n_ids, d_tok, n_docs, s->d_first_doc, s->d_sub_len, s->d_bad);
        s->last_launches += 2;
        int rc2 = dec_scan(ctx, s, s->d_sub_len, nsub, s->d_sub_off, st);
        if (rc2 != HUT_STATUS_OK) return rc2;
        unsigned long long total = 0;
        unsigned int bad2[3] = {0, 0, 0};
        CU(ctx, cudaMemcpyAsync(&total, s->d_sub_off + nsub, sizeof total, cudaMemcpyDeviceToHost, st));
        CU(ctx, cudaMemcpyAsync(bad2, s->d_bad, sizeof bad2, cudaMemcpyDeviceToHost, st));
        CU(ctx, cudaStreamSynchronize(st));
        if (bad2[0]) /* core.c:522-530 */
            return set_err(ctx, HUT_STATUS_VALUE, "%s", "Element must be non-negative and less than vocab size.");
        if (!bad2[1] && bad2[2]) { /* tokens that end like the beginning of a pattern: looked at in their context */
            dec2_soft_kernel<<<(unsigned)d->sms * 8, 256, 0, st>>>(D, d_ids, n_ids, d_tok, n_docs, s->d_bad);
            s->last_launches++;
            CU(ctx, cudaMemcpyAsync(bad2, s->d_bad, sizeof bad2, cudaMemcpyDeviceToHost, st));
            CU(ctx, cudaStreamSynchronize(st));
        }
        if (!bad2[1]) {
            CU(ctx, grow(&s->d_text, &s->d_text_cap, (size_t)total + 16));
            const unsigned g2 = (unsigned)((n_ids / DEC_SUB + 1 + G2_WARPS - 1) / G2_WARPS);
            if (pfx)
                dec2_gather_kernel<true><<<g2, G2_WARPS * 32, 0, st>>>(D, d_ids, n_ids, d_tok, n_docs, s->d_first_doc,
                                                                       s->d_sub_off, s->d_text, s->d_text_off);
            else
                dec2_gather_kernel<false><<<g2, G2_WARPS * 32, 0, st>>>(D, d_ids, n_ids, d_tok, n_docs, s->d_first_doc,
                                                                        s->d_sub_off, s->d_text, s->d_text_off);
            s->last_launches++;
            CU(ctx, cudaEventRecord(s->ev[1], st));
            CU(ctx, cudaGetLastError());
            s->last_path = 1;
            *n_text = total;
            return HUT_STATUS_OK;
        }
        CU(ctx, cudaMemsetAsync(s->d_bad, 0, 4 * sizeof(unsigned int), st));
    }
    s->last_path = 2;
    if (nsub) {
        dec_sub_kernel<<<(unsigned)((nsub + 7) / 8), 256, 0, st>>>(D, d_ids, n_ids, s->d_sub_len, s->d_bad);
        s->last_launches++;
    }
    int rc = dec_scan(ctx, s, s->d_sub_len, nsub, s->d_sub_off, st);
    if (rc != HUT_STATUS_OK) return rc;
    unsigned long long raw_total = 0;
    unsigned int bad = 0;
    CU(ctx, cudaMemcpyAsync(&raw_total, s->d_sub_off + nsub, sizeof raw_total, cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaMemcpyAsync(&bad, s->d_bad, sizeof bad, cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaStreamSynchronize(st));
    if (bad)  /* core.c:522-530 */
        return set_err(ctx, HUT_STATUS_VALUE, "%s", "Element must be non-negative and less than vocab size.");
    CU(ctx, grow(&s->d_raw, &s->d_raw_cap, (size_t)raw_total + 128));
    CU(ctx, grow(&s->d_dec, &s->d_dec_cap, (size_t)raw_total + 128));
    dec_gather_kernel<<<(unsigned)((n_ids / DEC_SUB + 1 + GATHER_WARPS - 1) / GATHER_WARPS), GATHER_WARPS * 32, 0, st>>>(
        D, d_ids, n_ids, s->d_sub_off, s->d_raw);
    s->last_launches++;
    if (n_docs) {
        const unsigned long long want = n_docs < (unsigned long long)d->sms * 12 ? n_docs : (unsigned long long)d->sms * 12;
        dec_doc_kernel<<<(unsigned)want, DEC_NT, 0, st>>>(D, d_ids, d_tok, n_docs, s->d_sub_off, s->d_raw, s->d_dec,
                                                          s->d_doc_len, s->d_doc_raw);
        s->last_launches++;
    }
    rc = dec_scan(ctx, s, s->d_doc_len, n_docs, s->d_text_off, st);
    if (rc != HUT_STATUS_OK) return rc;
    unsigned long long text_total = 0;
    CU(ctx, cudaMemcpyAsync(&text_total, s->d_text_off + n_docs, sizeof text_total, cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaStreamSynchronize(st));
    CU(ctx, grow(&s->d_text, &s->d_text_cap, (size_t)text_total + 16));
    if (n_docs) {
        const unsigned long long blocks = (n_docs + 7) / 8;
        const unsigned long long cap = (unsigned long long)d->sms * 16;
        dec_pack_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, st>>>(n_docs, s->d_doc_raw, s->d_dec,
                                                                                 s->d_text_off, s->d_text);
        s->last_launches++;
    }
    CU(ctx, cudaEventRecord(s->ev[1], st));
    CU(ctx, cudaGetLastError());
    *n_text = text_total;
    return HUT_STATUS_OK;
}

/* A few ids: one kernel launch on pinned host buffers, the host spins on the flag the kernel raises last.
 * HUT_SMALL_RETRY: dec_run has to do it (a token that does not decode independently, or more text than fits). */
static int decode_small(hut_ctx* ctx, const int32_t* ids, const uint64_t* tok_offsets, uint64_t n_docs,
                        hut_text_result* result) {
    hut_decstate* s = ctx->dec;
    hut_dev* d = &ctx->dev[0];
    const uint64_t base = tok_offsets[0], n = tok_offsets[n_docs] - base;
    if (!s->ds_ids) {
        CU(ctx, cudaMallocHost(&s->ds_ids, DS_MAX_IDS * sizeof(int32_t)));
        CU(ctx, cudaMallocHost(&s->ds_tok, (DS_MAX_DOCS + 1) * sizeof(unsigned long long)));
        CU(ctx, cudaMallocHost(&s->ds_text, DS_TEXT + 64));
        CU(ctx, cudaMallocHost(&s->ds_toff, (DS_MAX_DOCS + 1) * sizeof(unsigned long long)));
        CU(ctx, cudaMallocHost(&s->ds_hdr, sizeof(DecSmallHdr)));
        CU(ctx, cudaMallocHost((void**)&s->ds_done, 64));
        *s->ds_done = 0;
        s->ds_seq = 0;
    }
    memcpy(s->ds_ids, ids + base, (size_t)n * sizeof(int32_t));
    for (uint64_t i = 0; i <= n_docs; ++i) s->ds_tok[i] = tok_offsets[i] - base;
    const unsigned int seq = ++s->ds_seq;
    const DecModel& D = s->dm;
    if (D.prefix_len)
        dec_small_kernel<true><<<1, DS_NT, 0, d->stream>>>(D, s->ds_ids, (uint32_t)n, s->ds_tok, (uint32_t)n_docs, s->ds_text,
                                                           s->ds_toff, s->ds_hdr, s->ds_done, seq);
    else
        dec_small_kernel<false><<<1, DS_NT, 0, d->stream>>>(D, s->ds_ids, (uint32_t)n, s->ds_tok, (uint32_t)n_docs, s->ds_text,
                                                            s->ds_toff, s->ds_hdr, s->ds_done, seq);
    CU(ctx, cudaGetLastError());
    for (unsigned spins = 0; *s->ds_done != seq; ++spins) {
        if ((spins & 0xFFFFu) == 0xFFFFu) {
            cudaError_t e = cudaStreamQuery(d->stream);
            if (e != cudaSuccess && e != cudaErrorNotReady)
                return set_err(ctx, HUT_STATUS_CUDA, "CUDA error: %s", cudaGetErrorString(e));
            if (e == cudaSuccess && *s->ds_done != seq)
                return set_err(ctx, HUT_STATUS_CUDA, "%s", "dec_small_kernel ended without reporting");
        }
    }
    if (s->ds_hdr->flags & DS_BAD_ID) /* core.c:522-530 */
        return set_err(ctx, HUT_STATUS_VALUE, "%s", "Element must be non-negative and less than vocab size.");
    if (s->ds_hdr->flags & DS_FALLBACK) return HUT_SMALL_RETRY;
    const unsigned long long n_text = s->ds_hdr->total;
    s->ds_text[n_text] = 0;
    s->last_launches = 1;
    s->last_ms = 0.0f;
    s->last_path = 3;
    result->text = s->ds_text;
    result->text_offsets = (const uint64_t*)s->ds_toff;
    result->n_bytes = n_text;
    result->n_docs = n_docs;
    return HUT_STATUS_OK;
}

/* A large host batch goes through the devices in chunks of whole documents, chunk c on device c mod ndev: the ids of a
 * device's next chunk are uploaded (its copy-in stream, second set of input buffers) while its current chunk is decoded
 * (main stream) and the text of the one before leaves (copy-out stream, second set of output buffers), so the link is
 * busy in both directions at once and the kernels (2 ms per GiB of text) disappear behind the copies.  Documents are
 * decoded independently of each other (core.c:513-581 is called per document), so cutting between them changes nothing.
 * The host visits the chunks in order (a chunk's place in the text is known once the chunks before it have reported their
 * lengths); what it has queued on the other devices runs meanwhile. */
static int decode_chunked(hut_ctx* ctx, const int32_t* ids, const uint64_t* tok_offsets, uint64_t n_docs,
                          unsigned long long chunk_ids, hut_text_result* result) {
    const int ndev = ctx->ndev;
    hut_decstate* S[MAX_DEV];
    S[0] = ctx->dec;
    for (int i = 1; i < ndev; ++i) {
        int rc = dec_init_on(ctx, &ctx->dec_x[i], i);
        if (rc != HUT_STATUS_OK) return rc;
        S[i] = ctx->dec_x[i];
    }
    hut_decstate* const s0 = ctx->dec;  /* owner of the pinned result */
    std::vector<uint64_t> cut;  /* chunk c = documents [cut[c], cut[c + 1]) */
    cut.push_back(0);
    for (uint64_t i = 0; i < n_docs;) {
        uint64_t j = i + 1;
        while (j < n_docs && tok_offsets[j + 1] - tok_offsets[i] <= chunk_ids) ++j;
        cut.push_back(j);
        i = j;
    }
    const size_t nch = cut.size() - 1;
    size_t max_ids = 0, max_docs = 0;
    for (size_t c = 0; c < nch; ++c) {
        const size_t ni = (size_t)(tok_offsets[cut[c + 1]] - tok_offsets[cut[c]]), nd = (size_t)(cut[c + 1] - cut[c]);
        max_ids = ni > max_ids ? ni : max_ids;
        max_docs = nd > max_docs ? nd : max_docs;
    }
    for (int i = 0; i < ndev && (size_t)i < nch; ++i) {
        hut_decstate* s = S[i];
        CU(ctx, cudaSetDevice(ctx->dev[i].device));
        for (int b = 0; b < 2; ++b) {
            CU(ctx, grow(&s->c_ids[b], &s->c_ids_cap[b], max_ids + 1));
            CU(ctx, grow(&s->c_tok[b], &s->c_tok_cap[b], max_docs + 1));
            CU(ctx, grow_host(&s->h_ctok[b], &s->h_ctok_cap[b], max_docs + 1, false));
            if (!s->ev_up[b]) CU(ctx, cudaEventCreateWithFlags(&s->ev_up[b], cudaEventDisableTiming));
            if (!s->ev_k[b]) CU(ctx, cudaEventCreateWithFlags(&s->ev_k[b], cudaEventDisableTiming));
            if (!s->ev_dl[b]) CU(ctx, cudaEventCreateWithFlags(&s->ev_dl[b], cudaEventDisableTiming));
        }
    }
    CU(ctx, grow_host(&s0->h_text_off, &s0->h_text_off_cap, (size_t)n_docs + 1, false));
    auto sync_all = [&](bool out_only) -> int {
        for (int i = 0; i < ndev; ++i) {
            CU(ctx, cudaSetDevice(ctx->dev[i].device));
            CU(ctx, cudaStreamSynchronize(ctx->dev[i].s_out));
            if (!out_only) {
                CU(ctx, cudaStreamSynchronize(ctx->dev[i].s_in));
                CU(ctx, cudaStreamSynchronize(ctx->dev[i].stream));
            }
        }
        return HUT_STATUS_OK;
    };
    /* chunk c uses buffer set (c / ndev) & 1 of device c % ndev; the set's previous user is chunk c - 2 * ndev, whose
     * kernels the host has waited for (end of its visit) */
    auto upload = [&](size_t c) -> int {
        const int dv = (int)(c % (size_t)ndev), b = (int)((c / (size_t)ndev) & 1);
        hut_decstate* s = S[dv];
        CU(ctx, cudaSetDevice(ctx->dev[dv].device));
        cudaStream_t sin = ctx->dev[dv].s_in;
        const uint64_t lo = cut[c], hi = cut[c + 1], base = tok_offsets[lo];
        for (uint64_t i = lo; i <= hi; ++i) s->h_ctok[b][i - lo] = tok_offsets[i] - base;
        CU(ctx, cudaMemcpyAsync(s->c_tok[b], s->h_ctok[b], (size_t)(hi - lo + 1) * sizeof(unsigned long long),
                                cudaMemcpyHostToDevice, sin));
        if (tok_offsets[hi] > base)
            CU(ctx, cudaMemcpyAsync(s->c_ids[b], ids + base, (size_t)(tok_offsets[hi] - base) * sizeof(int32_t),
                                    cudaMemcpyHostToDevice, sin));
        CU(ctx, cudaEventRecord(s->ev_up[b], sin));
        return HUT_STATUS_OK;
    };
    int rc = HUT_STATUS_OK;
    for (size_t c = 0; c < nch && c < (size_t)ndev; ++c) {
        rc = upload(c);
        if (rc != HUT_STATUS_OK) return rc;
    }
    std::vector<unsigned long long> tbase(nch + 1, 0);
    float ms_sum = 0;
    unsigned launches = 0;
    int path = 1;
    for (size_t c = 0; c < nch; ++c) {
        const int dv = (int)(c % (size_t)ndev), b = (int)((c / (size_t)ndev) & 1);
        hut_decstate* s = S[dv];
        hut_dev* d = &ctx->dev[dv];
        cudaStream_t st = d->stream, sout = d->s_out;
        const uint64_t lo = cut[c], hi = cut[c + 1];
        if (c + (size_t)ndev < nch) {  /* this device's next chunk: on its way while this one is decoded */
            rc = upload(c + (size_t)ndev);
            if (rc != HUT_STATUS_OK) break;
        }
        CU(ctx, cudaSetDevice(d->device));
        CU(ctx, cudaStreamWaitEvent(st, s->ev_up[b], 0));
        if (c >= 2 * (size_t)ndev) CU(ctx, cudaStreamWaitEvent(st, s->ev_dl[b], 0)); /* this output set's last text has left */
        auto swap_out = [&]() {  /* odd visits of a device use its second set of output buffers */
            std::swap(s->d_text, s->x_text);
            std::swap(s->d_text_cap, s->x_text_cap);
            std::swap(s->d_text_off, s->x_text_off);
            std::swap(s->d_text_off_cap, s->x_text_off_cap);
        };
        unsigned long long n_text = 0;
        if (b) swap_out();
        rc = dec_run(ctx, s, s->c_ids[b], tok_offsets[hi] - tok_offsets[lo], s->c_tok[b], hi - lo, &n_text, st);
        const uint8_t* const o_text = s->d_text;
        const unsigned long long* const o_off = s->d_text_off;
        if (b) swap_out();
        if (rc != HUT_STATUS_OK) break;
        CU(ctx, cudaEventRecord(s->ev_k[b], st));
        launches += s->last_launches;
        if (s->last_path == 2) path = 2;
        const unsigned long long need = tbase[c] + n_text + 1;
        if (need > s0->h_text_cap) {
            /* room for the whole batch from the bytes per id seen so far (+ 1/8), at least what this chunk needs */
            const unsigned long long done_ids = tok_offsets[hi] - tok_offsets[0], all_ids = tok_offsets[n_docs] - tok_offsets[0];
            unsigned long long est = done_ids ? (unsigned long long)((double)(tbase[c] + n_text) / (double)done_ids * (double)all_ids) : need;
            est += est / 8 + 4096;
            rc = sync_all(true); /* earlier chunks' text is still on its way into the old buffer */
            if (rc != HUT_STATUS_OK) break;
            CU(ctx, cudaSetDevice(d->device));
            CU(ctx, grow_host(&s0->h_text, &s0->h_text_cap, (size_t)(est > need ? est : need), true));
        }
        /* download on the device's copy-out stream: its next chunk's kernels (other output set) do not wait for it */
        CU(ctx, cudaStreamWaitEvent(sout, s->ev_k[b], 0));
        if (n_text) CU(ctx, cudaMemcpyAsync(s0->h_text + tbase[c], o_text, (size_t)n_text, cudaMemcpyDeviceToHost, sout));
        /* chunk-local text offsets of the chunk's documents (the chunk's end is the next one's start) */
        CU(ctx, cudaMemcpyAsync(s0->h_text_off + lo, o_off, (size_t)(hi - lo) * sizeof(unsigned long long),
                                cudaMemcpyDeviceToHost, sout));
        CU(ctx, cudaEventRecord(s->ev_dl[b], sout));
        tbase[c + 1] = tbase[c] + n_text;
        /* the events of dec_run are re-recorded by the device's next chunk: read this chunk's kernel time once it is done
         * (the copies of both directions, and the other devices, carry on meanwhile) */
        CU(ctx, cudaEventSynchronize(s->ev[1]));
        float ms = 0;
        cudaEventElapsedTime(&ms, s->ev[0], s->ev[1]);
        ms_sum += ms;
    }
    {
        const int rc2 = sync_all(false);
        if (rc == HUT_STATUS_OK) rc = rc2;
    }
    CU(ctx, cudaSetDevice(ctx->dev[0].device));
    if (rc != HUT_STATUS_OK) return rc;
    for (size_t c = 1; c < nch; ++c)
        for (uint64_t i = cut[c]; i < cut[c + 1]; ++i) s0->h_text_off[i] += tbase[c];
    s0->h_text_off[n_docs] = tbase[nch];
    if (!s0->h_text) CU(ctx, grow_host(&s0->h_text, &s0->h_text_cap, 16, false));
    s0->h_text[tbase[nch]] = 0;
    s0->last_ms = ms_sum;
    s0->last_launches = launches;
    s0->last_path = path;
    result->text = s0->h_text;
    result->text_offsets = (const uint64_t*)s0->h_text_off;
    result->n_bytes = tbase[nch];
    result->n_docs = n_docs;
    return HUT_STATUS_OK;
}

extern "C" int hut_decode_batch(hut_ctx* ctx, const int32_t* ids, const uint64_t* tok_offsets, uint64_t n_docs,
                                hut_text_result* result) {
    if (!ctx) return HUT_STATUS_STATE;
    if (!result || (!tok_offsets && n_docs)) return set_err(ctx, HUT_STATUS_ARG, "%s", "null argument");
    int rc = dec_init(ctx);
    if (rc != HUT_STATUS_OK) return rc;
    hut_decstate* s = ctx->dec;
    hut_dev* d = &ctx->dev[0];
    CU(ctx, cudaSetDevice(d->device));
    const unsigned long long base = n_docs ? tok_offsets[0] : 0, n_ids = n_docs ? tok_offsets[n_docs] - base : 0;
    for (uint64_t i = 0; i < n_docs; ++i)
        if (tok_offsets[i + 1] < tok_offsets[i]) return set_err(ctx, HUT_STATUS_ARG, "%s", "tok_offsets must not decrease");
    if (n_ids && !ids) return set_err(ctx, HUT_STATUS_ARG, "%s", "null argument");
    if (s->token_wise && n_ids && n_ids <= (unsigned long long)DS_MAX_IDS && n_docs <= (uint64_t)DS_MAX_DOCS &&
        !getenv("HUTOKEN_B200_NO_SMALL")) {
        rc = decode_small(ctx, ids, tok_offsets, n_docs, result);
        if (rc != HUT_SMALL_RETRY) return rc;
    }
    {
        /* ids per chunk: hut_options.chunk_mb of int32 ids (default 32 MiB); a batch of up to one and a half chunks goes in one piece */
        unsigned long long chunk_ids = ((unsigned long long)(ctx->opt.chunk_mb ? ctx->opt.chunk_mb : 32u) << 20) / sizeof(int32_t);
        if (getenv("HUTOKEN_B200_DECODE_CHUNK_IDS")) chunk_ids = strtoull(getenv("HUTOKEN_B200_DECODE_CHUNK_IDS"), nullptr, 10);  /* tests */
        if (chunk_ids && n_docs > 1 && n_ids > chunk_ids + chunk_ids / 2) {
            return decode_chunked(ctx, ids, tok_offsets, n_docs, chunk_ids, result);
        }
    }
    cudaStream_t st = d->stream;
    CU(ctx, grow(&s->d_ids, &s->d_ids_cap, (size_t)n_ids + 1));
    CU(ctx, grow(&s->d_tok, &s->d_tok_cap, (size_t)n_docs + 1));
    CU(ctx, grow_host(&s->h_text_off, &s->h_text_off_cap, (size_t)n_docs + 1, false));
    /* offsets rebased to the first id of the batch */
    for (uint64_t i = 0; i <= n_docs; ++i) s->h_text_off[i] = n_docs ? tok_offsets[i] - base : 0;
    CU(ctx, cudaMemcpyAsync(s->d_tok, s->h_text_off, ((size_t)n_docs + 1) * sizeof(unsigned long long),
                            cudaMemcpyHostToDevice, st));
    if (n_ids) CU(ctx, cudaMemcpyAsync(s->d_ids, ids + base, (size_t)n_ids * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    CU(ctx, cudaStreamSynchronize(st)); /* h_text_off is reused below */
    unsigned long long n_text = 0;
    rc = dec_run(ctx, s, s->d_ids, n_ids, s->d_tok, n_docs, &n_text, st);
    if (rc != HUT_STATUS_OK) return rc;
    CU(ctx, grow_host(&s->h_text, &s->h_text_cap, (size_t)n_text + 1, false));
    if (n_text) CU(ctx, cudaMemcpyAsync(s->h_text, s->d_text, (size_t)n_text, cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaMemcpyAsync(s->h_text_off, s->d_text_off, ((size_t)n_docs + 1) * sizeof(unsigned long long),
                            cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaStreamSynchronize(st));
    s->h_text[n_text] = 0;
    cudaEventElapsedTime(&s->last_ms, s->ev[0], s->ev[1]);
    result->text = s->h_text;
    result->text_offsets = (const uint64_t*)s->h_text_off;
    result->n_bytes = n_text;
    result->n_docs = n_docs;
    return HUT_STATUS_OK;
}

extern "C" int hut_decode_device(hut_ctx* ctx, const int32_t* d_ids, uint64_t n_ids, const uint64_t* d_tok_offsets,
                                 uint64_t n_docs, const uint8_t** d_text_out, const uint64_t** d_text_offsets_out,
                                 uint64_t* n_bytes_out, void* stream) {
    if (!ctx) return HUT_STATUS_STATE;
    int rc = dec_init(ctx);
    if (rc != HUT_STATUS_OK) return rc;
    hut_dev* d = &ctx->dev[0];
    CU(ctx, cudaSetDevice(d->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : d->stream;
    unsigned long long n_text = 0;
    rc = dec_run(ctx, ctx->dec, d_ids, n_ids, (const unsigned long long*)d_tok_offsets, n_docs, &n_text, st);
    if (rc != HUT_STATUS_OK) return rc;
    CU(ctx, cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->dec->last_ms, ctx->dec->ev[0], ctx->dec->ev[1]);
    if (d_text_out) *d_text_out = ctx->dec->d_text;
    if (d_text_offsets_out) *d_text_offsets_out = (const uint64_t*)ctx->dec->d_text_off;
    if (n_bytes_out) *n_bytes_out = n_text;
    return HUT_STATUS_OK;
}

extern "C" int hut_decode_stats(const hut_ctx* ctx, double* kernel_ms, uint64_t* kernel_launches) {
    if (!ctx || !ctx->dec) return HUT_STATUS_STATE;
    if (kernel_ms) *kernel_ms = ctx->dec->last_ms;
    if (kernel_launches) *kernel_launches = ctx->dec->last_launches;
    return HUT_STATUS_OK;
}

extern "C" int hut_decode_path(const hut_ctx* ctx) { return ctx && ctx->dec ? ctx->dec->last_path : 0; }

/* See include/hutoken_b200.h.  The token-wise kernels of dec_run with their scratch in g_noalloc_scratch; no
 * cudaMalloc / cudaFree / grow() on this path. */
extern "C" int hut_decode_device_noalloc(hut_ctx* ctx, const int32_t* d_ids, uint64_t n_ids, const uint64_t* d_tok_offsets,
                                         uint64_t n_docs, uint8_t* d_text, uint64_t text_cap, uint64_t* d_text_offsets,
                                         uint64_t* n_bytes_out, void* stream) {
    if (!ctx) return HUT_STATUS_STATE;
    if (n_bytes_out) *n_bytes_out = 0;
    hut_decstate* s = ctx->dec;
    if (!s || !s->ready || !s->na_scratch)
        return set_err(ctx, HUT_STATUS_STATE, "%s",
                       "hut_decode_device_noalloc: decode tables not built yet; call another hut_decode_* function once first");
    if (n_ids > NA_MAX_IDS)
        return set_err(ctx, HUT_STATUS_ARG, "%s",
                       "hut_decode_device_noalloc: at most 16777216 ids (HUT_NOALLOC_MAX_IDS) per call; split the batch or use "
                       "hut_decode_device");
    if (n_docs >= 0xFFFFFFF0ull) return set_err(ctx, HUT_STATUS_ARG, "%s", "hut_decode_device_noalloc: too many documents");
    if (!d_tok_offsets || !d_text_offsets || (n_ids && (!d_ids || !d_text)))
        return set_err(ctx, HUT_STATUS_ARG, "%s", "null argument");
    hut_dev* d = &ctx->dev[0];
    CU(ctx, cudaSetDevice(d->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : d->stream;
    const DecModel& D = s->dm;
    unsigned char* const base = s->na_scratch;
    unsigned int* const bad = reinterpret_cast<unsigned int*>(base + NA_OFF_BAD);
    uint32_t* const sub_len = reinterpret_cast<uint32_t*>(base + NA_OFF_SUB_LEN);
    unsigned long long* const sub_off = reinterpret_cast<unsigned long long*>(base + NA_OFF_SUB_OFF);
    uint32_t* const first_doc = reinterpret_cast<uint32_t*>(base + NA_OFF_FIRST_DOC);
    unsigned long long* const bsum = reinterpret_cast<unsigned long long*>(base + NA_OFF_BSUM);
    const unsigned long long* const tok = (const unsigned long long*)d_tok_offsets;
    unsigned long long* const toff = (unsigned long long*)d_text_offsets;
    if (!n_ids || !n_docs) { /* nothing to print: every document is empty */
        if (!n_docs) {
            CU(ctx, cudaMemsetAsync(toff, 0, sizeof(unsigned long long), st));
        } else {
            unsigned long long ends[2] = {0, 0};
            CU(ctx, cudaMemcpyAsync(ends, tok + (n_docs - 1), 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            CU(ctx, cudaStreamSynchronize(st));
            if (ends[1] != 0) return set_err(ctx, HUT_STATUS_ARG, "%s", "d_tok_offsets must end at n_ids");
            CU(ctx, cudaMemsetAsync(toff, 0, ((size_t)n_docs + 1) * sizeof(unsigned long long), st));
        }
        CU(ctx, cudaStreamSynchronize(st));
        return HUT_STATUS_OK;
    }
    const bool pfx = D.prefix_len != 0;
    const unsigned long long nsub = (n_ids + DEC_SUB - 1) / DEC_SUB;
    const size_t nfirst = (size_t)(n_ids / DEC_SUB + 1);
    /* the array keeps the previous call's contents: flags and the first-document table start from their empty values;
     * sub_len / sub_off / bsum are written in full before they are read */
    CU(ctx, cudaMemsetAsync(bad, 0, 4 * sizeof(unsigned int), st));
    CU(ctx, cudaMemsetAsync(first_doc, 0xFF, nfirst * sizeof(uint32_t), st));
    dec2_first_doc_kernel<<<(unsigned)(n_docs / 256 + 1), 256, 0, st>>>(tok, n_docs, first_doc);
    const unsigned g1 = (unsigned)((nsub + 7) / 8);
    if (pfx) dec2_sub_kernel<true><<<g1, 256, 0, st>>>(D, d_ids, n_ids, tok, n_docs, first_doc, sub_len, bad);
    else dec2_sub_kernel<false><<<g1, 256, 0, st>>>(D, d_ids, n_ids, tok, n_docs, first_doc, sub_len, bad);
    const unsigned nb = (unsigned)(nsub / SCAN_BLOCK + 1);
    scan_sums_kernel<<<nb, SCAN_NT, 0, st>>>(sub_len, nsub, bsum);
    scan_top_kernel<<<1, 1024, 0, st>>>(bsum, nb);
    scan_apply_kernel<<<nb, SCAN_NT, 0, st>>>(sub_len, nsub, bsum, sub_off);
    unsigned long long total = 0;
    unsigned int flags[3] = {0, 0, 0};
    CU(ctx, cudaMemcpyAsync(&total, sub_off + nsub, sizeof total, cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaMemcpyAsync(flags, bad, sizeof flags, cudaMemcpyDeviceToHost, st));
    CU(ctx, cudaStreamSynchronize(st));
    if (flags[0]) /* core.c:522-530 */
        return set_err(ctx, HUT_STATUS_VALUE, "%s", "Element must be non-negative and less than vocab size.");
    if (!flags[1] && flags[2]) {
        dec2_soft_kernel<<<(unsigned)d->sms * 8, 256, 0, st>>>(D, d_ids, n_ids, tok, n_docs, bad);
        CU(ctx, cudaMemcpyAsync(flags, bad, sizeof flags, cudaMemcpyDeviceToHost, st));
        CU(ctx, cudaStreamSynchronize(st));
    }
    if (flags[1])
        return set_err(ctx, HUT_STATUS_VALUE, "%s",
                       "hut_decode_device_noalloc: this batch needs the general decode kernels (a token ends inside a character or "
                       "a pattern, or d_tok_offsets does not span [0, n_ids]); use hut_decode_device");
    if (n_bytes_out) *n_bytes_out = total;
    if (total > text_cap) return set_err(ctx, HUT_STATUS_CAPACITY, "%s", "hut_decode_device_noalloc: text_cap too small");
    const unsigned g2 = (unsigned)((n_ids / DEC_SUB + 1 + G2_WARPS - 1) / G2_WARPS);
    if (pfx) dec2_gather_kernel<true><<<g2, G2_WARPS * 32, 0, st>>>(D, d_ids, n_ids, tok, n_docs, first_doc, sub_off, d_text, toff);
    else dec2_gather_kernel<false><<<g2, G2_WARPS * 32, 0, st>>>(D, d_ids, n_ids, tok, n_docs, first_doc, sub_off, d_text, toff);
    CU(ctx, cudaGetLastError());
    CU(ctx, cudaStreamSynchronize(st));
    return HUT_STATUS_OK;
}
