"""ctypes binding of the C-ABI (include/hutoken_b200.h) for callers that hold packed buffers: bench.py, the
parity tests, pipelines that keep text and ids on the device.  Fails loudly when the library is missing."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HUTOKEN_B200_LIB") or os.path.join(_HERE, "libhutoken_b200.so")  # override: A/B runs

STATUS_NAMES = {0: "OK", 1: "FILE_NOT_FOUND", 2: "VALUE", 3: "MEMORY", 4: "CUDA", 5: "UNSUPPORTED", 6: "ARG",
                7: "STATE", 8: "CAPACITY"}
_EXC = {1: FileNotFoundError, 2: ValueError, 3: MemoryError, 6: ValueError}


class HutError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"[{STATUS_NAMES.get(status, status)}] {msg}")
        self.status = status
        self.message = msg


class _Desc(C.Structure):
    _fields_ = [("vocab_file_path", C.c_char_p), ("special_file_path", C.c_char_p), ("prefix", C.c_char_p),
                ("is_byte_encoder", C.c_int), ("special_token_id", C.c_int), ("pattern", C.c_char_p),
                ("merges_file_path", C.c_char_p)]


class _Options(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("cache_log2_slots", C.c_uint32), ("max_pretoken", C.c_uint64),
                ("big_slots", C.c_uint32), ("chunk_mb", C.c_uint32)]


class _Result(C.Structure):
    _fields_ = [("ids", C.POINTER(C.c_int32)), ("tok_offsets", C.POINTER(C.c_uint64)), ("n_ids", C.c_uint64),
                ("n_docs", C.c_uint64)]


class _TextResult(C.Structure):
    _fields_ = [("text", C.POINTER(C.c_uint8)), ("text_offsets", C.POINTER(C.c_uint64)), ("n_bytes", C.c_uint64),
                ("n_docs", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("pretokens", C.c_uint64), ("cache_hits", C.c_uint64), ("warp_merges", C.c_uint64),
                ("block_merges", C.c_uint64), ("big_merges", C.c_uint64), ("cut_documents", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("kernel_ms", C.c_double), ("main_kernel_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


NOALLOC_MAX_IDS = 1 << 24  # HUT_NOALLOC_MAX_IDS

EXPORTS = ["hut_create", "hut_destroy", "hut_last_error", "hut_encode_batch", "hut_encode_device",
           "hut_cache_reset", "hut_host_alloc", "hut_host_free", "hut_get_stats", "hut_vocab_lines",
           "hut_device_count", "hut_plan_shards", "hut_decode_batch", "hut_decode_device", "hut_decode_stats",
           "hut_decode_path", "hut_decode_device_noalloc"]

_lib = None


def load_library():
    """dlopen libhutoken_b200.so (no CUDA call is made until hut_create)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C hutoken_b200/csrc` "
                          "(or __graft_entry__.build()); the engine has no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    lib.hut_create.argtypes = [C.POINTER(_Desc), C.POINTER(_Options), C.POINTER(C.c_int), C.c_int,
                               C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]
    lib.hut_destroy.argtypes = [C.c_void_p]
    lib.hut_last_error.restype = C.c_char_p
    lib.hut_last_error.argtypes = [C.c_void_p]
    lib.hut_encode_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(_Result)]
    lib.hut_encode_device.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                      C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p]
    lib.hut_decode_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(_TextResult)]
    lib.hut_decode_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                      C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_void_p]
    lib.hut_decode_stats.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    lib.hut_decode_path.argtypes = [C.c_void_p]
    lib.hut_decode_device_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p,
                                              C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p]
    lib.hut_decode_path.restype = C.c_int
    lib.hut_cache_reset.argtypes = [C.c_void_p]
    lib.hut_host_alloc.restype = C.c_void_p
    lib.hut_host_alloc.argtypes = [C.c_size_t]
    lib.hut_host_free.argtypes = [C.c_void_p]
    lib.hut_get_stats.argtypes = [C.c_void_p, C.c_int, C.POINTER(Stats)]
    lib.hut_vocab_lines.restype = C.c_uint32
    lib.hut_vocab_lines.argtypes = [C.c_void_p]
    lib.hut_device_count.argtypes = [C.c_void_p]
    _lib = lib
    return lib


def _raise(status, msg):
    raise _EXC.get(status, HutError)(msg) if status in _EXC else HutError(status, msg)


class Engine:
    """One model on one or more GPUs of this box."""

    def __init__(self, vocab_path, special_path, prefix=None, is_byte_encoder=False, devices=None,
                 max_pretoken=0, cache_log2_slots=0, big_slots=0, chunk_mb=0, merges_path=None):
        lib = load_library()
        desc = _Desc(os.fsencode(vocab_path), os.fsencode(special_path),
                     None if prefix is None else prefix.encode("utf-8"), int(bool(is_byte_encoder)), -1, None,
                     None if merges_path is None else os.fsencode(merges_path))
        opt = _Options(C.sizeof(_Options), cache_log2_slots, max_pretoken, big_slots, chunk_mb)
        err = C.create_string_buffer(512)
        h = C.c_void_p()
        if devices:
            arr = (C.c_int * len(devices))(*devices)
            rc = lib.hut_create(C.byref(desc), C.byref(opt), arr, len(devices), C.byref(h), err, len(err))
        else:
            rc = lib.hut_create(C.byref(desc), C.byref(opt), None, 0, C.byref(h), err, len(err))
        if rc != 0:
            _raise(rc, err.value.decode("utf-8", "replace"))
        self._h, self._lib = h, lib

    def close(self):
        if getattr(self, "_h", None):
            self._lib.hut_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc != 0:
            _raise(rc, self._lib.hut_last_error(self._h).decode("utf-8", "replace"))

    def encode_packed(self, data, offs, copy=True):
        """Host buffers in, (ids int32[n], tok_offsets uint64[ndocs+1]) out.  data: uint8 array or address."""
        offs = np.ascontiguousarray(offs, dtype=np.uint64)
        nd = len(offs) - 1
        ptr = data if isinstance(data, int) else np.ascontiguousarray(data, dtype=np.uint8).ctypes.data
        res = _Result()
        self._check(self._lib.hut_encode_batch(self._h, ptr, offs.ctypes.data, nd, C.byref(res)))
        n = int(res.n_ids)
        ids = np.ctypeslib.as_array(res.ids, shape=(n,)) if n else np.zeros(0, np.int32)
        tok = np.ctypeslib.as_array(res.tok_offsets, shape=(nd + 1,))
        return (ids.copy(), tok.copy()) if copy else (ids, tok)

    def batch_encode(self, texts):
        datas = [t.encode("utf-8") if isinstance(t, str) else bytes(t) for t in texts]
        offs = np.zeros(len(datas) + 1, dtype=np.uint64)
        if datas:
            offs[1:] = np.cumsum([len(d) for d in datas])
        data = np.frombuffer(b"".join(datas) + b"\0" * 16, dtype=np.uint8)
        ids, tok = self.encode_packed(data, offs)
        return [ids[int(tok[i]):int(tok[i + 1])].tolist() for i in range(len(datas))]

    def encode(self, text):
        return self.batch_encode([text])[0]

    def encode_device(self, dev_index, d_bytes, n_bytes, d_offs, n_docs, d_ids, ids_capacity, d_tok, stream=None):
        """Everything resident on the device: raw device addresses in, number of ids out."""
        n = C.c_uint64()
        self._check(self._lib.hut_encode_device(self._h, dev_index, d_bytes, n_bytes, d_offs, n_docs, d_ids,
                                                ids_capacity, d_tok, C.byref(n), stream))
        return int(n.value)

    def decode_packed(self, ids, tok):
        """ids int32[n], tok uint64[ndocs+1] (host) -> (text uint8[n_bytes], text_offsets uint64[ndocs+1])."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        tok = np.ascontiguousarray(tok, dtype=np.uint64)
        nd = len(tok) - 1
        res = _TextResult()
        self._check(self._lib.hut_decode_batch(self._h, ids.ctypes.data, tok.ctypes.data, nd, C.byref(res)))
        n = int(res.n_bytes)
        text = np.ctypeslib.as_array(res.text, shape=(n,)).copy() if n else np.zeros(0, np.uint8)
        offs = np.ctypeslib.as_array(res.text_offsets, shape=(nd + 1,)).copy()
        return text, offs

    def batch_decode_bytes(self, lists):
        tok = np.zeros(len(lists) + 1, dtype=np.uint64)
        if lists:
            tok[1:] = np.cumsum([len(x) for x in lists])
        ids = np.fromiter((i for x in lists for i in x), dtype=np.int32, count=int(tok[-1]))
        text, offs = self.decode_packed(ids, tok)
        raw = text.tobytes()
        return [raw[int(offs[i]):int(offs[i + 1])] for i in range(len(lists))]

    def decode_device(self, d_ids, n_ids, d_tok, n_docs, stream=None):
        """ids / offsets resident on the context's first device -> (d_text address, d_text_offsets address, n_bytes);
        the outputs are device buffers owned by the context."""
        t, o, n = C.c_void_p(), C.c_void_p(), C.c_uint64()
        self._check(self._lib.hut_decode_device(self._h, d_ids, n_ids, d_tok, n_docs, C.byref(t), C.byref(o),
                                                C.byref(n), stream))
        return t.value, o.value, int(n.value)

    def decode_device_noalloc(self, d_ids, n_ids, d_tok, n_docs, d_text, text_cap, d_text_off, stream=None):
        """hut_decode_device_noalloc: caller-owned device outputs, scratch in the library's static device array, no
        allocation; at most NOALLOC_MAX_IDS ids.  Returns the text length."""
        n = C.c_uint64()
        self._check(self._lib.hut_decode_device_noalloc(self._h, d_ids, n_ids, d_tok, n_docs, d_text, text_cap,
                                                        d_text_off, C.byref(n), stream))
        return int(n.value)

    def decode_stats(self):
        ms, n = C.c_double(), C.c_uint64()
        self._lib.hut_decode_stats(self._h, C.byref(ms), C.byref(n))
        path = {0: None, 1: "token-wise", 2: "general", 3: "token-wise, one launch"}[self._lib.hut_decode_path(self._h)]
        return {"kernel_ms": ms.value, "kernel_launches": n.value, "path": path}

    def cache_reset(self):
        self._check(self._lib.hut_cache_reset(self._h))

    def stats(self, dev_index=0):
        s = Stats()
        self._lib.hut_get_stats(self._h, dev_index, C.byref(s))
        return s.as_dict()

    @property
    def vocab_lines(self):
        return self._lib.hut_vocab_lines(self._h)

    @property
    def device_count(self):
        return self._lib.hut_device_count(self._h)
