"""Python surface of the B200 engine: the public functions of the reference's hutoken.py (hutoken.py:22-191)
with the same names, arguments and error wrapping, on top of this package's `_hutoken` extension.

Only the local-vocab-file branch of initialize() (hutoken.py:26-43) exists here; the Hugging Face download
branch (hutoken.py:45-120) needs network access and is out of scope.
"""
import os
import sys
import traceback

try:
    from . import _hutoken
except ImportError as _e:  # the extension is built in-tree by __graft_entry__.build() / csrc/Makefile
    _hutoken = None
    _import_error = _e

_MISSING = "hutoken: Native C extension '_hutoken' is not installed or failed to import."


def _ext():
    if _hutoken is None:
        raise RuntimeError(f"{_MISSING} ({_import_error})")
    return _hutoken


def initialize(model_or_path, *args, **kwargs):
    """initialize(vocab_path, special_chars_path, prefix=None, is_byte_encoder=False, token_id=-1, pattern=None,
    merges_file=None).  merges_file (keyword, or the 7th positional argument after the vocab path like the
    reference's hub branch passes it, hutoken.py:112) switches to id-keyed ranks (core.c:211-337)."""
    ext = _ext()
    if not os.path.isfile(model_or_path):
        raise ValueError(f"'{model_or_path}' is not a vocab file; Hugging Face model names are not supported "
                         "by the B200 engine (no network): convert the tokenizer to hutoken's vocab format first.")
    special_chars_file = args[0] if args else kwargs.get("special_file_path")
    merges_file = args[6] if len(args) > 6 else kwargs.get("merges_file", kwargs.get("merges_file_path"))
    if special_chars_file and not os.path.isfile(special_chars_file):
        raise ValueError(f"Special characters file '{special_chars_file}' does not exist.")
    if merges_file and not os.path.isfile(merges_file):
        raise ValueError(f"The provided merges file '{merges_file}' does not exist.")
    return ext.initialize(model_or_path, special_chars_file, kwargs.get("prefix"),
                          kwargs.get("is_byte_encoder", False), kwargs.get("token_id", -1),
                          kwargs.get("pattern"), merges_file)


def encode(text):
    ext = _ext()
    try:
        return ext.encode(text)
    except Exception as e:
        traceback.print_exc(file=sys.stderr)
        raise RuntimeError(f"hutoken: Error encoding text: {e}")


def batch_encode(texts, num_threads=1):
    ext = _ext()
    try:
        return ext.batch_encode(texts, num_threads)
    except Exception as e:
        traceback.print_exc(file=sys.stderr)
        raise RuntimeError(f"hutoken: Error encoding texts: {e}")


def pack_texts(texts):
    """list of str / bytes -> (data: bytes, offsets: uint64 ndarray of len(texts) + 1).  Each text is cut at its
    first NUL, as batch_encode does (strdup at lib.c:772)."""
    import numpy as np
    if not isinstance(texts, (list, tuple)):
        raise TypeError("Invalid arguments. Expected a list of strings.")
    parts = []
    offs = np.zeros(len(texts) + 1, dtype=np.uint64)
    at = 0
    for i, t in enumerate(texts):
        if isinstance(t, str):
            t = t.encode("utf-8")
        elif not isinstance(t, (bytes, bytearray, memoryview)):
            raise TypeError("Invalid arguments. Expected a list of strings.")
        t = bytes(t)
        nul = t.find(b"\0")
        if nul >= 0:
            t = t[:nul]
        parts.append(t)
        at += len(t)
        offs[i + 1] = at
    return b"".join(parts), offs


def batch_encode_packed(texts, offsets=None):
    """Zero-object variant of batch_encode (no counterpart in hutoken.py, whose
    batch_encode builds one PyLong per id, lib.c:840-868).

    batch_encode_packed(list_of_texts)        -> (ids: int32 ndarray, tok_offsets: uint64 ndarray of n_docs + 1)
    batch_encode_packed(data, offsets)        -> the same for documents already packed back to back in one
                                                 bytes-like `data`, document i = data[offsets[i]:offsets[i + 1]]
    Document i's ids are ids[tok_offsets[i]:tok_offsets[i + 1]]; they equal batch_encode(texts)[i]."""
    import numpy as np
    ext = _ext()
    try:
        if offsets is None:
            data, offsets = pack_texts(texts)
        else:
            data = texts
            offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
            if offsets.ndim != 1 or len(offsets) < 1 or (np.diff(offsets.astype(np.int64)) < 0).any():
                raise ValueError("offsets must be a non-decreasing uint64 array with n_docs + 1 entries")
        ids, tok = ext.batch_encode_packed(data, offsets)
        return np.frombuffer(ids, dtype=np.int32), np.frombuffer(tok, dtype=np.uint64)
    except Exception as e:
        traceback.print_exc(file=sys.stderr)
        raise RuntimeError(f"hutoken: Error encoding texts: {e}")


def decode(tokens):
    ext = _ext()
    try:
        return ext.decode(tokens)
    except ValueError as e:
        traceback.print_exc(file=sys.stderr)
        raise ValueError(f"hutoken: Error decoding tokens {tokens}: {e}")
    except Exception as e:
        traceback.print_exc(file=sys.stderr)
        raise RuntimeError(f"hutoken: Error decoding tokens: {e}")


def batch_decode(tokens, num_threads=1):
    ext = _ext()
    try:
        return ext.batch_decode(tokens, num_threads)
    except Exception as e:
        traceback.print_exc(file=sys.stderr)
        raise RuntimeError(f"hutoken: Error decoding tokens: {e}")


def bpe_train(*args, **kwargs):
    return _ext().bpe_train(*args, **kwargs)


def bbpe_train(*args, **kwargs):
    return _ext().bbpe_train(*args, **kwargs)


def initialize_foma():
    raise RuntimeError("hutoken: '_hutoken' does not provide 'initialize_foma' or Foma support is not installed.")


def look_up_word(*args):
    raise RuntimeError("hutoken: '_hutoken' does not provide 'look_up_word' or Foma support is not installed.")
