"""ctypes front-end of tests/host_harness.c: the engine's HOST-side tables (hut_model.c) and the local
pretoken-start rule (hut_classify.h) walked by a scalar test driver.  Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_SO = os.path.join(_HERE, "_build", "libhost_harness.so")
_SRCS = [os.path.join(_HERE, "host_harness.c"), os.path.join(_ROOT, "hutoken_b200", "csrc", "hut_model.c")]
_DEPS = _SRCS + [os.path.join(_ROOT, "hutoken_b200", "csrc", f)
                 for f in ("hut_model.h", "hut_classify.h", "hut_bitslice.h")]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO) or any(os.path.getmtime(_SO) < os.path.getmtime(d) for d in _DEPS):
            os.makedirs(os.path.dirname(_SO), exist_ok=True)
            subprocess.run(["gcc", "-O2", "-g", "-std=c11", "-fPIC", "-shared", "-o", _SO] + _SRCS, check=True)
        L = C.CDLL(_SO)
        L.harness_load.restype = C.c_void_p
        L.harness_load.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_char_p, C.c_size_t,
                                   C.POINTER(C.c_int)]
        L.harness_free.argtypes = [C.c_void_p]
        L.harness_encode.restype = C.c_size_t
        L.harness_encode.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.harness_decode.restype = C.c_size_t
        L.harness_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.harness_pretokens.restype = C.c_size_t
        L.harness_pretokens.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.harness_info.argtypes = [C.c_void_p, C.c_void_p]
        L.harness_pretokens_bs.restype = C.c_size_t
        L.harness_pretokens_bs.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_void_p, C.c_size_t,
                                           C.POINTER(C.c_int)]
        L.harness_bitslice_check.argtypes = [C.c_char_p]
        _lib = L
    return _lib


class HostModel:
    def __init__(self, vocab, special, prefix=None, is_byte_encoder=False, merges=None):
        L = lib()
        msg = C.create_string_buffer(512)
        st = C.c_int()
        h = L.harness_load(os.fsencode(vocab), os.fsencode(special), None if prefix is None else prefix.encode(),
                           int(bool(is_byte_encoder)), None if merges is None else os.fsencode(merges), msg,
                           len(msg), C.byref(st))
        self.status, self.message = st.value, msg.value.decode("utf-8", "replace")
        self._h = C.c_void_p(h) if h else None

    def ok(self):
        return self._h is not None and self.status == 0

    def encode(self, text):
        data = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        cap = 8 * len(data) + 64
        out = np.empty(cap, dtype=np.int32)
        n = lib().harness_encode(self._h, data, len(data), out.ctypes.data, cap)
        assert n <= cap
        return out[:n].tolist()

    def decode_bytes(self, ids):
        arr = np.ascontiguousarray(ids, dtype=np.int32)
        cap = 64 * len(arr) + 64
        out = np.empty(cap, dtype=np.uint8)
        n = lib().harness_decode(self._h, arr.ctypes.data, len(arr), out.ctypes.data, cap)
        if n == C.c_size_t(-1).value:
            raise ValueError("Element must be non-negative and less than vocab size.")
        if n > cap:
            out = np.empty(n, dtype=np.uint8)
            n = lib().harness_decode(self._h, arr.ctypes.data, len(arr), out.ctypes.data, n)
        return out[:n].tobytes()

    def info(self):
        out = np.zeros(9, dtype=np.uint32)
        lib().harness_info(self._h, out.ctypes.data)
        keys = ["n_vocab", "n_units", "pair_count", "pair_cap", "distinct_values", "identity_ids",
                "max_sym_per_char", "vocab_lines", "monotone"]
        return dict(zip(keys, out.tolist()))

    def close(self):
        if self._h:
            lib().harness_free(self._h)
            self._h = None

    __del__ = close


def pretoken_starts(data: bytes):
    cap = len(data) + 1
    out = np.empty(cap, dtype=np.uint32)
    n = lib().harness_pretokens(data, len(data), out.ctypes.data, cap)
    return out[:n].tolist()


def pretoken_starts_bitsliced(data: bytes, doc_starts=(0,)):
    """(starts, exotic) from the bit-sliced classifier; doc_starts = byte positions where documents begin."""
    isds = bytearray(len(data) + 1)
    for p in doc_starts:
        if p < len(data):
            isds[p] = 1
    cap = len(data) + 1
    out = np.empty(cap, dtype=np.uint32)
    ex = C.c_int()
    n = lib().harness_pretokens_bs(data, len(data), bytes(isds), out.ctypes.data, cap, C.byref(ex))
    return out[:n].tolist(), bool(ex.value)


def bitslice_ok(chunk32: bytes) -> bool:
    assert len(chunk32) == 32
    return lib().harness_bitslice_check(chunk32) == 0
