"""TEST INFRASTRUCTURE ONLY -- ctypes front-end for the CPU oracle.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  The shipped package (hutoken_b200/) never does.

Two checkers live here:
  * ``Oracle``      -- our plain-C restatement (oracle/hut_oracle.c) of the reference's encode path.
  * ``load_ref()``  -- the UNMODIFIED reference extension compiled by oracle/Makefile into oracle/_ref
                       (its ``_hutoken`` module: initialize / encode / batch_encode, lib.c:1128-1145).
"""
from __future__ import annotations

import ctypes as C
import glob
import importlib.machinery
import importlib.util
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libhut_oracle.so")


def build(ref: bool = True) -> None:
    """Compile the restatement and (when /root/reference exists) the reference itself."""
    targets = ["oracle"] + (["ref"] if ref else [])
    subprocess.run(["make", "-s", "-C", _HERE] + targets, check=True)


class _Stats(C.Structure):
    _fields_ = [("pretokens", C.c_uint64), ("symbols", C.c_uint64), ("tokens", C.c_uint64), ("probes", C.c_uint64)]


_lib = None


def _load():
    global _lib
    if _lib is not None:
        return _lib
    src = os.path.join(_HERE, "hut_oracle.c")
    if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        build(ref=False)
    lib = C.CDLL(_LIB)
    lib.hut_oracle_load.restype = C.c_void_p
    lib.hut_oracle_load.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_size_t]
    lib.hut_oracle_free.argtypes = [C.c_void_p]
    lib.hut_oracle_set_limit_mode.argtypes = [C.c_void_p, C.c_int]
    lib.hut_oracle_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.POINTER(C.c_int32)),
                                      C.POINTER(C.c_size_t)]
    lib.hut_oracle_free_ids.argtypes = [C.POINTER(C.c_int32)]
    lib.hut_oracle_encode_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64,
                                            C.POINTER(C.POINTER(C.c_int32)), C.c_void_p]
    lib.hut_oracle_next_pretoken.restype = C.c_size_t
    lib.hut_oracle_next_pretoken.argtypes = [C.c_char_p, C.c_size_t, C.c_size_t]
    lib.hut_oracle_get_stats.argtypes = [C.POINTER(_Stats)]
    lib.hut_oracle_vocab_lookup.restype = C.c_int32
    lib.hut_oracle_vocab_lookup.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
    lib.hut_oracle_vocab_lines.restype = C.c_size_t
    lib.hut_oracle_vocab_lines.argtypes = [C.c_void_p]
    lib.hut_oracle_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.POINTER(C.c_uint8)),
                                      C.POINTER(C.c_size_t)]
    lib.hut_oracle_free_text.argtypes = [C.POINTER(C.c_uint8)]
    lib.hut_oracle_load_merges.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_size_t]
    lib.hut_oracle_merge_rules.restype = C.c_size_t
    lib.hut_oracle_merge_rules.argtypes = [C.c_void_p]
    lib.hut_oracle_merges_active.argtypes = [C.c_void_p]
    _lib = lib
    return lib


class OracleError(Exception):
    pass


class Oracle:
    """CPU restatement of hutoken.initialize + encode / batch_encode (default parser path)."""

    def __init__(self, vocab_path, special_path, prefix=None, is_byte_encoder=False, no_limit=False,
                 merges_path=None):
        lib = _load()
        err = C.create_string_buffer(512)
        h = lib.hut_oracle_load(os.fsencode(vocab_path), os.fsencode(special_path),
                                None if prefix is None else prefix.encode("utf-8"), int(bool(is_byte_encoder)),
                                err, len(err))
        if not h:
            raise OracleError(err.value.decode("utf-8", "replace"))
        self._h = C.c_void_p(h)
        self._lib = lib
        if no_limit:
            lib.hut_oracle_set_limit_mode(self._h, 1)
        if merges_path is not None:
            if not lib.hut_oracle_load_merges(self._h, os.fsencode(merges_path), err, len(err)):
                raise OracleError(err.value.decode("utf-8", "replace"))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.hut_oracle_free(self._h)
            self._h = None

    __del__ = close

    def encode(self, text) -> list:
        data = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        ids = C.POINTER(C.c_int32)()
        n = C.c_size_t()
        buf = C.create_string_buffer(data, len(data) + 1)
        self._lib.hut_oracle_encode(self._h, buf, len(data), C.byref(ids), C.byref(n))
        out = [ids[i] for i in range(n.value)]
        self._lib.hut_oracle_free_ids(ids)
        return out

    def encode_packed(self, data: np.ndarray, offs: np.ndarray):
        """data: uint8 array, offs: uint64 [ndocs+1].  Returns (ids int32 array, tok_offs uint64 array)."""
        data = np.ascontiguousarray(data, dtype=np.uint8)
        offs = np.ascontiguousarray(offs, dtype=np.uint64)
        nd = len(offs) - 1
        tok = np.zeros(nd + 1, dtype=np.uint64)
        ids = C.POINTER(C.c_int32)()
        self._lib.hut_oracle_encode_batch(self._h, data.ctypes.data, offs.ctypes.data, nd, C.byref(ids),
                                          tok.ctypes.data)
        n = int(tok[nd])
        out = np.ctypeslib.as_array(ids, shape=(n,)).copy() if n else np.zeros(0, np.int32)
        self._lib.hut_oracle_free_ids(ids)
        return out, tok

    def batch_encode(self, texts) -> list:
        datas = [t.encode("utf-8") if isinstance(t, str) else bytes(t) for t in texts]
        offs = np.zeros(len(datas) + 1, dtype=np.uint64)
        offs[1:] = np.cumsum([len(d) for d in datas])
        data = np.frombuffer(b"".join(datas) + b"\0", dtype=np.uint8)
        ids, tok = self.encode_packed(data, offs)
        return [ids[int(tok[i]):int(tok[i + 1])].tolist() for i in range(len(datas))]

    def decode_bytes(self, ids) -> bytes:
        """decode() + pretokenizer_decode: the raw bytes (the reference's binding then cuts at the first 0x00)."""
        arr = np.ascontiguousarray(ids, dtype=np.int32)
        text = C.POINTER(C.c_uint8)()
        n = C.c_size_t()
        if self._lib.hut_oracle_decode(self._h, arr.ctypes.data, len(arr), C.byref(text), C.byref(n)) != 0:
            raise ValueError("Element must be non-negative and less than vocab size.")
        out = bytes(C.cast(text, C.POINTER(C.c_uint8 * n.value)).contents) if n.value else b""
        self._lib.hut_oracle_free_text(text)
        return out

    def decode(self, ids) -> str:
        """What hutoken.decode returns: PyUnicode_FromString of the C string (lib.c:946)."""
        return self.decode_bytes(ids).split(b"\0", 1)[0].decode("utf-8")

    def stats(self) -> dict:
        s = _Stats()
        self._lib.hut_oracle_get_stats(C.byref(s))
        return {k: getattr(s, k) for k, _ in _Stats._fields_}

    def lookup(self, key: bytes) -> int:
        return self._lib.hut_oracle_vocab_lookup(self._h, key, len(key))

    def pretokens(self, text) -> list:
        data = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        out, pos = [], 0
        while True:
            n = self._lib.hut_oracle_next_pretoken(data, len(data), pos)
            if n == 0:
                break
            out.append(data[pos:pos + n])
            pos += n
        return out


def next_pretoken(data: bytes, pos: int) -> int:
    return _load().hut_oracle_next_pretoken(data, len(data), pos)


def ref_path():
    hits = sorted(glob.glob(os.path.join(_HERE, "_ref", "_hutoken*.so")))
    return hits[0] if hits else None


def load_ref():
    """Import the unmodified reference extension from oracle/_ref (None if it was never built).

    The reference keeps its state in process globals (lib.c:73-74), so there is one live model per
    process: call ``ref.initialize(vocab, special, prefix, is_byte_encoder)`` to switch.
    """
    path = ref_path()
    if path is None:
        if os.path.isdir("/root/reference/src"):
            build(ref=True)
            path = ref_path()
        if path is None:
            return None
    loader = importlib.machinery.ExtensionFileLoader("_hutoken", path)
    spec = importlib.util.spec_from_file_location("_hutoken", path, loader=loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    return mod
