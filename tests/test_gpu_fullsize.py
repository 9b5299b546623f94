"""configs[1] at its full size (1 GiB of synthetic documents, GPT-2-shaped 50k vocab) through size-independent
properties, because the oracle cannot encode a gigabyte in test time:

  * sharding invariance: the batch encoded in one call == the concatenation of four independently encoded document
    ranges (the per-rank shards of a multi-GPU run), compared by SHA-256 of ids and rebased offsets;
  * round trip: decode(encode(x)) == x byte for byte (every byte is in a byte-level vocabulary, and no document of the
    generator starts with a space);
  * a prefix of the result (8 MiB of documents) against the oracle itself.

HUTOKEN_B200_FULLSIZE_MB overrides the size (the default is BASELINE.json's 1 GiB)."""
import hashlib
import os

import numpy as np
import pytest

import synth
from common import engine_for, oracle_for

pytestmark = pytest.mark.gpu

MB = int(os.environ.get("HUTOKEN_B200_FULLSIZE_MB", "1024"))


def _corpus():
    parts = [synth.corpus(2000 + i, min(256, MB - 256 * i) << 20) for i in range((MB + 255) // 256)]
    data = np.concatenate([d for d, _ in parts])
    offs, base = [np.zeros(1, np.uint64)], 0
    for d, o in parts:
        offs.append(o[1:] + np.uint64(base))
        base += len(d)
    return np.concatenate([data, np.zeros(64, np.uint8)]), np.concatenate(offs)


def test_full_size_properties():
    name = "gpt2"
    eng, orc = engine_for(name), oracle_for(name)
    data, offs = _corpus()
    nd = len(offs) - 1
    assert int(offs[-1]) >= (MB << 20) - 4096
    ids, tok = eng.encode_packed(data, offs)
    assert len(tok) == nd + 1 and int(tok[-1]) == len(ids) and (ids >= 0).all()
    whole = hashlib.sha256(ids.tobytes()).hexdigest()
    # ---- four shards, encoded independently ----
    h = hashlib.sha256()
    cuts = [nd * k // 4 for k in range(5)]
    pieces = []
    for a, b in zip(cuts[:-1], cuts[1:]):
        lo, hi = int(offs[a]), int(offs[b])
        sub = np.concatenate([data[lo:hi], np.zeros(64, np.uint8)])
        sids, stok = eng.encode_packed(sub, offs[a:b + 1] - offs[a])
        h.update(sids.tobytes())
        pieces.append(stok[:-1] + tok[a])
    assert h.hexdigest() == whole
    assert np.array_equal(np.concatenate(pieces + [tok[-1:]]), tok)
    # ---- a prefix against the oracle ----
    k = int(np.searchsorted(offs, np.uint64(8 << 20)))
    oids, otok = orc.encode_packed(data[:int(offs[k]) + 1], offs[:k + 1])
    assert np.array_equal(ids[:len(oids)], oids) and np.array_equal(tok[:k + 1], otok)
    # ---- round trip ----
    text, toff = eng.decode_packed(ids, tok)
    assert np.array_equal(toff, offs)
    assert np.array_equal(text, data[:len(text)])
