"""Shared helpers for the test-suite: models, engines, oracles."""
import functools
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402

MODELS = ["gpt2_small", "llama_small", "gpt2", "llama"]


@functools.lru_cache(maxsize=None)
def oracle_for(name, no_limit=False):
    v, s = synth.PATHS[name]
    return O.Oracle(v, s, no_limit=no_limit, **synth.MODEL_KW[name])


_engines = {}


def engine_for(name, **kw):
    from hutoken_b200 import capi
    key = (name, tuple(sorted(kw.items())))
    if key not in _engines:
        v, s = synth.PATHS[name]
        _engines[key] = capi.Engine(v, s, **synth.MODEL_KW[name], **kw)
    return _engines[key]


def pack(docs):
    datas = [d.encode("utf-8") if isinstance(d, str) else bytes(d) for d in docs]
    offs = np.zeros(len(datas) + 1, dtype=np.uint64)
    if datas:
        offs[1:] = np.cumsum([len(d) for d in datas])
    data = np.frombuffer(b"".join(datas) + b"\0" * 16, dtype=np.uint8)
    return data, offs


def first_diff(a, b):
    n = min(len(a), len(b))
    neq = np.nonzero(np.asarray(a[:n]) != np.asarray(b[:n]))[0]
    return int(neq[0]) if len(neq) else n


def assert_same(ids, tok, oids, otok, what=""):
    ids, oids = np.asarray(ids), np.asarray(oids)
    if len(ids) != len(oids) or not np.array_equal(ids, oids):
        i = first_diff(ids, oids)
        raise AssertionError(f"{what}: ids differ at {i} (engine n={len(ids)}, oracle n={len(oids)}): "
                             f"{ids[max(0, i - 4):i + 6].tolist()} vs {oids[max(0, i - 4):i + 6].tolist()}")
    assert np.array_equal(np.asarray(tok, dtype=np.uint64), np.asarray(otok, dtype=np.uint64)), \
        f"{what}: per-document offsets differ"
