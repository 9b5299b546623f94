/* `_hutoken` CPython extension for the B200 engine -- same module name, method table, argument formats,
 * defaults and error strings as the reference extension (src/lib.c:185-232, 668-874, 1128-1153), so that
 * hutoken.py's surface (initialize / encode / batch_encode) works unchanged on top of it.
 * All tokenisation happens in libhutoken_b200.so (include/hutoken_b200.h); this file only converts
 * between Python objects and packed buffers.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <pthread.h>
#include <string.h>

#include "../../include/hutoken_b200.h"

static hut_ctx* g_ctx = NULL; /* one live model per process, like lib.c:73-74 */
static uint8_t* g_pack = NULL; /* pinned packing buffer (grow-only) */
static size_t g_pack_cap = 0;
static uint64_t* g_offs = NULL;
static size_t g_offs_cap = 0;

/* One engine call at a time.  g_ctx, the packing buffers above, the decode buffers below and the context's pinned
 * result are process-global (the reference, lib.c:782, uses per-call buffers instead), and include/hutoken_b200.h
 * allows one host thread per context: every method takes this mutex before it touches them and drops it after the
 * result has been converted to Python objects.  The GIL is released while waiting, so a thread that holds the
 * mutex and wants the GIL back can always get it. */
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
static void engine_lock(void) {
    Py_BEGIN_ALLOW_THREADS
    pthread_mutex_lock(&g_mu);
    Py_END_ALLOW_THREADS
}
static void engine_unlock(void) { pthread_mutex_unlock(&g_mu); }

static const char* NOT_INIT_ENC =
    "Vocabulary is not initialized for encoding. Call 'initialize_encode' function first."; /* lib.c:671-676 */
static const char* NOT_INIT_DEC =
    "Vocabulary is not initialized for decoding. Call 'initialize_decode' function first.";

static PyObject* raise_status(int status, const char* msg) {
    PyObject* type = PyExc_RuntimeError;
    switch (status) {
        case HUT_STATUS_FILE_NOT_FOUND: type = PyExc_FileNotFoundError; break;
        case HUT_STATUS_VALUE: type = PyExc_ValueError; break;
        case HUT_STATUS_MEMORY: type = PyExc_MemoryError; break;
        case HUT_STATUS_ARG: type = PyExc_ValueError; break;
        default: break;
    }
    PyErr_SetString(type, msg && msg[0] ? msg : "hutoken_b200 engine error");
    return NULL;
}

static int ensure_pack(size_t bytes, size_t ndocs) {
    if (bytes + 64 > g_pack_cap) {
        size_t cap = bytes + bytes / 4 + (1u << 16);
        uint8_t* p = hut_host_alloc(cap);
        if (!p) return 0;
        hut_host_free(g_pack);
        g_pack = p;
        g_pack_cap = cap;
    }
    if (ndocs + 2 > g_offs_cap) {
        size_t cap = ndocs + ndocs / 4 + 1024;
        uint64_t* p = hut_host_alloc(cap * sizeof(uint64_t));
        if (!p) return 0;
        hut_host_free(g_offs);
        g_offs = p;
        g_offs_cap = cap;
    }
    return 1;
}

/* initialize(vocab_file_path, special_file_path, prefix=None, is_byte_encoder=False, special_token_id=-1,
 *            pattern=None, merges_file_path=None)   -- lib.c:185-232, format "ss|zpizz" */
static PyObject* p_initialize(PyObject* self, PyObject* args, PyObject* kwargs) {
    static char* kwlist[] = {"vocab_file_path", "special_file_path", "prefix", "is_byte_encoder",
                             "special_token_id", "pattern", "merges_file_path", NULL};
    char *vocab = NULL, *special = NULL, *prefix = NULL, *pattern = NULL, *merges = NULL;
    int is_byte_encoder = 0, special_token_id = -1;
    if (!PyArg_ParseTupleAndKeywords(args, kwargs, "ss|zpizz", kwlist, &vocab, &special, &prefix, &is_byte_encoder,
                                     &special_token_id, &pattern, &merges)) {
        PyErr_SetString(PyExc_TypeError,
                        "Invalid arguments. Expected a string "
                        "(vocab_file_path), a string (special_file_path), "
                        "a string or None (prefix) a bool an"
                        "optional integer (special_token_id), "
                        " an optional string (regex_pattern) and"
                        "a string or None (merges_file_path)");
        return NULL;
    }
    hut_model_desc desc = {vocab, special, prefix, is_byte_encoder, special_token_id, pattern, merges};
    hut_options opt;
    memset(&opt, 0, sizeof opt);
    opt.struct_size = sizeof opt;
    const char* nl = getenv("HUTOKEN_B200_MAX_PRETOKEN"); /* opt-in "no-limit" mode, SURVEY.md 8(c) */
    if (nl) opt.max_pretoken = strtoull(nl, NULL, 10);
    /* HUTOKEN_B200_DEVICES="0,1,2,3": shard batch_encode over several GPUs of the box */
    int devs[16], nd = 0;
    const char* dv = getenv("HUTOKEN_B200_DEVICES");
    if (dv) {
        const char* p = dv;
        while (*p && nd < 16) {
            char* e = NULL;
            long v = strtol(p, &e, 10);
            if (e == p) break;
            devs[nd++] = (int)v;
            p = (*e == ',') ? e + 1 : e;
        }
    }
    char err[512] = "";
    hut_ctx* ctx = NULL;
    int rc;
    Py_BEGIN_ALLOW_THREADS
    rc = hut_create(&desc, &opt, nd ? devs : NULL, nd, &ctx, err, sizeof err);
    Py_END_ALLOW_THREADS
    if (rc != HUT_STATUS_OK) return raise_status(rc, err);
    engine_lock(); /* no encode / decode call is inside the old context now */
    hut_ctx* old = g_ctx;
    g_ctx = ctx;
    engine_unlock();
    if (old) hut_destroy(old); /* re-initialise replaces the model (the reference leaks it, lib.c:129-145) */
    Py_RETURN_NONE;
}

/* The reference builds one PyLong per id (lib.c:840-868); a token id is one of at most 2^21 values and most of a
 * result repeats a few thousand of them, so each value's int object is created once (on first use, under the GIL)
 * and handed out again with a new reference afterwards: the lists hold ordinary ints equal to the reference's, built
 * several times faster (bench.py `latency.batch_encode_list_api`). */
#define ID_OBJ_MAX (1u << 21)
static PyObject** g_id_obj; /* [ID_OBJ_MAX], NULL until an id is first seen; lives as long as the module */

static inline PyObject* id_object(int32_t id) {
    if ((uint32_t)id < ID_OBJ_MAX && g_id_obj) {
        PyObject* v = g_id_obj[id];
        if (!v) {
            v = PyLong_FromLong(id);
            if (!v) return NULL;
            g_id_obj[id] = v; /* the table keeps its own reference */
        }
        Py_INCREF(v);
        return v;
    }
    return PyLong_FromLong(id); /* -1 (unknown symbol), ids beyond the table, or no table */
}

static PyObject* ids_to_list(const int32_t* ids, uint64_t n) {
    PyObject* list = PyList_New((Py_ssize_t)n);
    if (!list) return PyErr_NoMemory();
    if (!g_id_obj) g_id_obj = (PyObject**)calloc(ID_OBJ_MAX, sizeof(PyObject*)); /* pages touched only where ids occur */
    for (uint64_t i = 0; i < n; ++i) {
        PyObject* v = id_object(ids[i]);
        if (!v) {
            Py_DECREF(list);
            return PyErr_NoMemory();
        }
        PyList_SET_ITEM(list, (Py_ssize_t)i, v);
    }
    return list;
}

/* encode(text) -- lib.c:668-720, format "s" (rejects embedded NUL like the reference) */
static PyObject* p_encode(PyObject* self, PyObject* args) {
    if (!g_ctx) {
        PyErr_SetString(PyExc_RuntimeError, NOT_INIT_ENC);
        return NULL;
    }
    char* text = NULL;
    if (!PyArg_ParseTuple(args, "s", &text)) return NULL;
    size_t len = strlen(text);
    engine_lock();
    PyObject* out = NULL;
    if (!ensure_pack(len, 1)) {
        PyErr_NoMemory();
    } else {
        memcpy(g_pack, text, len);
        g_offs[0] = 0;
        g_offs[1] = len;
        hut_result res;
        int rc;
        Py_BEGIN_ALLOW_THREADS
        rc = hut_encode_batch(g_ctx, g_pack, g_offs, 1, &res);
        Py_END_ALLOW_THREADS
        if (rc != HUT_STATUS_OK) raise_status(rc, hut_last_error(g_ctx));
        else out = ids_to_list(res.ids, res.n_ids);
    }
    engine_unlock();
    return out;
}

/* batch_encode(texts, num_threads=1) -- lib.c:722-874, format "O|i".  num_threads only keeps its one
 * observable effect: <= 0 workers encode nothing (lib.c:779-794). */
static PyObject* p_batch_encode(PyObject* self, PyObject* args) {
    if (!g_ctx) {
        PyErr_SetString(PyExc_RuntimeError, NOT_INIT_ENC);
        return NULL;
    }
    PyObject* texts = NULL;
    int num_threads = 1;
    if (!PyArg_ParseTuple(args, "O|i", &texts, &num_threads) || !PyList_Check(texts)) {
        PyErr_SetString(PyExc_TypeError, "Invalid arguments. Expected a list of strings.");
        return NULL;
    }
    Py_ssize_t n = PyList_GET_SIZE(texts);
    if (num_threads <= 0) {
        PyObject* result = PyList_New(n);
        if (!result) return PyErr_NoMemory();
        for (Py_ssize_t i = 0; i < n; ++i) PyList_SET_ITEM(result, i, PyList_New(0));
        return result;
    }
    /* sizing and packing both happen under the mutex with the GIL held throughout, so the list cannot change in
     * between.  pass 1: sizes (each text is cut at its first NUL, like strdup at lib.c:772) */
    engine_lock();
    PyObject* result = NULL;
    size_t total = 0;
    for (Py_ssize_t i = 0; i < n; ++i) {
        const char* s = PyUnicode_AsUTF8(PyList_GET_ITEM(texts, i));
        if (!s) {
            PyErr_SetString(PyExc_TypeError, "Invalid arguments. Expected a list of strings.");
            goto out;
        }
        total += strlen(s);
    }
    if (!ensure_pack(total, (size_t)n)) {
        PyErr_NoMemory();
        goto out;
    }
    {
        size_t at = 0;
        for (Py_ssize_t i = 0; i < n; ++i) {
            const char* s = PyUnicode_AsUTF8(PyList_GET_ITEM(texts, i));
            size_t l = strlen(s);
            g_offs[i] = at;
            memcpy(g_pack + at, s, l);
            at += l;
        }
        g_offs[n] = at;
    }
    {
        hut_result res;
        int rc;
        Py_BEGIN_ALLOW_THREADS
        rc = hut_encode_batch(g_ctx, g_pack, g_offs, (uint64_t)n, &res);
        Py_END_ALLOW_THREADS
        if (rc != HUT_STATUS_OK) {
            raise_status(rc, hut_last_error(g_ctx));
            goto out;
        }
        result = PyList_New(n);
        if (!result) {
            PyErr_NoMemory();
            goto out;
        }
        for (Py_ssize_t i = 0; i < n; ++i) {
            PyObject* sub = ids_to_list(res.ids + res.tok_offsets[i], res.tok_offsets[i + 1] - res.tok_offsets[i]);
            if (!sub) {
                Py_CLEAR(result);
                goto out;
            }
            PyList_SET_ITEM(result, i, sub);
        }
    }
out:
    engine_unlock();
    return result;
}

/* batch_encode_packed(texts) -> (ids: bytes of int32, offsets: bytes of uint64): zero-object output for
 * callers that do not want one PyLong per token. */
static PyObject* p_batch_encode_packed(PyObject* self, PyObject* args) {
    if (!g_ctx) {
        PyErr_SetString(PyExc_RuntimeError, NOT_INIT_ENC);
        return NULL;
    }
    Py_buffer data, offs;
    if (!PyArg_ParseTuple(args, "y*y*", &data, &offs)) return NULL;
    PyObject* out = NULL;
    if (offs.len < 8 || (offs.len % 8) != 0) {
        PyErr_SetString(PyExc_ValueError, "offsets must be a uint64 buffer with n_docs + 1 entries");
    } else {
        uint64_t nd = (uint64_t)offs.len / 8 - 1;
        const uint64_t* o = (const uint64_t*)offs.buf;
        int sorted = 1; /* decreasing offsets would underflow the shard planner's byte counts */
        for (uint64_t i = 0; i < nd && sorted; ++i) sorted = o[i] <= o[i + 1];
        if (!sorted) {
            PyErr_SetString(PyExc_ValueError, "offsets must be non-decreasing");
        } else if (o[nd] > (uint64_t)data.len) {
            PyErr_SetString(PyExc_ValueError, "offsets exceed the data buffer");
        } else {
            hut_result res;
            int rc;
            engine_lock();
            Py_BEGIN_ALLOW_THREADS
            rc = hut_encode_batch(g_ctx, (const uint8_t*)data.buf, o, nd, &res);
            Py_END_ALLOW_THREADS
            if (rc != HUT_STATUS_OK) raise_status(rc, hut_last_error(g_ctx));
            else
                out = Py_BuildValue("(y#y#)", (const char*)res.ids, (Py_ssize_t)(res.n_ids * 4),
                                    (const char*)res.tok_offsets, (Py_ssize_t)((nd + 1) * 8));
            engine_unlock();
        }
    }
    PyBuffer_Release(&data);
    PyBuffer_Release(&offs);
    return out;
}

/* ---- decode side ------------------------------------------------------------------------------------------ */
static int32_t* g_dec_ids = NULL;
static uint64_t* g_dec_offs = NULL;
static size_t g_dec_ids_cap = 0, g_dec_offs_cap = 0;

static int ensure_dec(size_t n_ids, size_t n_docs) {
    if (n_ids + 1 > g_dec_ids_cap) {
        size_t cap = (n_ids + 1) * 2;
        int32_t* p = realloc(g_dec_ids, cap * sizeof *p);
        if (!p) return 0;
        g_dec_ids = p;
        g_dec_ids_cap = cap;
    }
    if (n_docs + 2 > g_dec_offs_cap) {
        size_t cap = (n_docs + 2) * 2;
        uint64_t* p = realloc(g_dec_offs, cap * sizeof *p);
        if (!p) return 0;
        g_dec_offs = p;
        g_dec_offs_cap = cap;
    }
    return 1;
}

/* (int)PyLong_AsLong(item) as lib.c:923,1015: a non-integer item leaves an exception set, which we surface */
static int list_to_ids(PyObject* list, int32_t* out) {
    Py_ssize_t n = PyList_GET_SIZE(list);
    for (Py_ssize_t i = 0; i < n; ++i) {
        long v = PyLong_AsLong(PyList_GET_ITEM(list, i));
        if (v == -1 && PyErr_Occurred()) return 0;
        out[i] = (v < INT32_MIN || v > INT32_MAX) ? -1 : (int32_t)v; /* out of range either way (core.c:522) */
    }
    return 1;
}

/* PyUnicode_FromString (lib.c:946,1066): the C string ends at the first 0x00 */
static PyObject* text_to_str(const uint8_t* p, uint64_t n) {
    const void* z = memchr(p, 0, (size_t)n);
    return PyUnicode_DecodeUTF8((const char*)p, z ? (Py_ssize_t)((const uint8_t*)z - p) : (Py_ssize_t)n, NULL);
}

/* decode(tokens) -- lib.c:876-955, format "O" */
static PyObject* p_decode(PyObject* self, PyObject* args) {
    if (!g_ctx) {
        PyErr_SetString(PyExc_RuntimeError, NOT_INIT_DEC);
        return NULL;
    }
    PyObject* tokens = NULL;
    if (!PyArg_ParseTuple(args, "O", &tokens)) {
        PyErr_SetString(PyExc_TypeError, "Failed to parse arguments. Expected a single list of tokens.");
        return NULL;
    }
    if (!PyList_Check(tokens)) {
        PyErr_SetString(PyExc_TypeError, "Argument must be a list of integers");
        return NULL;
    }
    engine_lock();
    PyObject* out = NULL;
    Py_ssize_t n = PyList_GET_SIZE(tokens);
    if (!ensure_dec((size_t)n, 1)) {
        PyErr_NoMemory();
    } else if (list_to_ids(tokens, g_dec_ids)) {
        g_dec_offs[0] = 0;
        g_dec_offs[1] = (uint64_t)n;
        hut_text_result res;
        int rc;
        Py_BEGIN_ALLOW_THREADS
        rc = hut_decode_batch(g_ctx, g_dec_ids, g_dec_offs, 1, &res);
        Py_END_ALLOW_THREADS
        if (rc != HUT_STATUS_OK) raise_status(rc, hut_last_error(g_ctx));
        else out = text_to_str(res.text, res.n_bytes);
    }
    engine_unlock();
    return out;
}

/* batch_decode(tokens, num_threads=1) -- lib.c:957-1085, format "O|i"; num_threads is accepted and unused */
static PyObject* p_batch_decode(PyObject* self, PyObject* args) {
    if (!g_ctx) {
        PyErr_SetString(PyExc_RuntimeError, NOT_INIT_DEC);
        return NULL;
    }
    PyObject* tokens = NULL;
    int num_threads = 1;
    if (!PyArg_ParseTuple(args, "O|i", &tokens, &num_threads)) {
        PyErr_SetString(PyExc_TypeError, "Failed to parse arguments. Expected a single list of tokens.");
        return NULL;
    }
    Py_ssize_t nd = PyList_Check(tokens) ? PyList_GET_SIZE(tokens) : -1;
    if (nd <= 0) { /* PyList_Size of a non-list is -1 in the reference too (lib.c:979-983) */
        PyErr_Clear();
        PyErr_SetString(PyExc_ValueError, "No tokens provided.");
        return NULL;
    }
    engine_lock(); /* sizes are taken under the mutex: nothing runs between sizing and filling the buffers */
    PyObject* result = NULL;
    size_t total = 0;
    for (Py_ssize_t i = 0; i < nd; ++i) {
        PyObject* item = PyList_GET_ITEM(tokens, i);
        if (!PyList_Check(item)) {
            PyErr_SetString(PyExc_TypeError, "Each item must be a list of integers.");
            goto out;
        }
        total += (size_t)PyList_GET_SIZE(item);
    }
    if (!ensure_dec(total, (size_t)nd)) {
        PyErr_NoMemory();
        goto out;
    }
    {
        size_t at = 0;
        for (Py_ssize_t i = 0; i < nd; ++i) {
            PyObject* item = PyList_GET_ITEM(tokens, i);
            g_dec_offs[i] = at;
            if (!list_to_ids(item, g_dec_ids + at)) goto out;
            at += (size_t)PyList_GET_SIZE(item);
        }
        g_dec_offs[nd] = at;
    }
    {
        hut_text_result res;
        int rc;
        Py_BEGIN_ALLOW_THREADS
        rc = hut_decode_batch(g_ctx, g_dec_ids, g_dec_offs, (uint64_t)nd, &res);
        Py_END_ALLOW_THREADS
        if (rc != HUT_STATUS_OK) {
            raise_status(rc, hut_last_error(g_ctx));
            goto out;
        }
        result = PyList_New(nd);
        if (!result) {
            PyErr_NoMemory();
            goto out;
        }
        for (Py_ssize_t i = 0; i < nd; ++i) {
            PyObject* str = text_to_str(res.text + res.text_offsets[i], res.text_offsets[i + 1] - res.text_offsets[i]);
            if (!str) {
                Py_CLEAR(result);
                goto out;
            }
            PyList_SET_ITEM(result, i, str);
        }
    }
out:
    engine_unlock();
    return result;
}

static PyObject* p_out_of_scope(PyObject* self, PyObject* args) {
    PyErr_SetString(PyExc_RuntimeError, "offline trainers are outside the B200 encode engine's scope (SURVEY.md 2)");
    return NULL;
}

static PyMethodDef methods[] = {
    {"bpe_train", p_out_of_scope, METH_VARARGS, "BPE training (not provided by the B200 engine)"},
    {"bbpe_train", p_out_of_scope, METH_VARARGS, "BBPE training (not provided by the B200 engine)"},
    {"initialize", (PyCFunction)p_initialize, METH_VARARGS | METH_KEYWORDS, "Initalize tokenizer"},
    {"encode", p_encode, METH_VARARGS, "Encodes string"},
    {"batch_encode", p_batch_encode, METH_VARARGS, "Encodes list of strings"},
    {"batch_encode_packed", p_batch_encode_packed, METH_VARARGS, "Encodes packed documents into packed ids"},
    {"decode", p_decode, METH_VARARGS, "Decodes list of ints"},
    {"batch_decode", p_batch_decode, METH_VARARGS, "Decodes list of lists of ints"},
    {NULL, NULL, 0, NULL}};

static struct PyModuleDef moduledef = {PyModuleDef_HEAD_INIT, "huToken", "B200-native hutoken encode engine", -1,
                                       methods};

PyMODINIT_FUNC PyInit__hutoken(void) { return PyModule_Create(&moduledef); }
