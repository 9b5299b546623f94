"""configs[0] of BASELINE.json: a vocabulary trained by the reference's own bpe_train (5256 requested; it stops early,
bpe.c:221-224), post-processed as SURVEY.md F6 documents (tools/make_c1.py), then encode of a 1 MB string.  The golden
count + SHA-256 of the id stream come from the unmodified reference."""
import hashlib
import json
import os

import numpy as np
import pytest

from common import ROOT
from oracle import oracle as O

import make_c1


def _golden():
    with open(os.path.join(ROOT, "tests", "golden", "c1_encode.json")) as f:
        return json.load(f)


def _check(ids, g):
    ids = np.asarray(ids, dtype=np.int32)
    assert len(ids) == g["n_ids"] and ids[:64].tolist() == g["first_ids"]
    assert int((ids < 0).sum()) == g["unknown_ids"]
    assert hashlib.sha256(ids.tobytes()).hexdigest() == g["sha256_int32_le"]


def test_oracle_matches_reference_on_trained_vocab():
    g = _golden()
    text = make_c1.c1_text(g["text_seed"], make_c1.TEXT_BYTES)
    assert len(text.encode("utf-8")) == g["text_bytes"]
    orc = O.Oracle(os.path.join(ROOT, g["vocab"]), os.path.join(ROOT, g["special"]))
    _check(orc.encode(text), g)


def test_host_tables_on_trained_vocab():
    import harness
    g = _golden()
    text = make_c1.c1_text(g["text_seed"], 64 << 10)
    orc = O.Oracle(os.path.join(ROOT, g["vocab"]), os.path.join(ROOT, g["special"]))
    hm = harness.HostModel(os.path.join(ROOT, g["vocab"]), os.path.join(ROOT, g["special"]))
    assert hm.ok(), hm.message
    assert hm.encode(text) == orc.encode(text)


@pytest.mark.gpu
def test_gpu_matches_reference_on_trained_vocab():
    from hutoken_b200 import capi
    g = _golden()
    text = make_c1.c1_text(g["text_seed"], make_c1.TEXT_BYTES)
    eng = capi.Engine(os.path.join(ROOT, g["vocab"]), os.path.join(ROOT, g["special"]))
    _check(eng.encode(text), g)
    # the same text as 1000 documents
    docs = text.split("\n")
    orc = O.Oracle(os.path.join(ROOT, g["vocab"]), os.path.join(ROOT, g["special"]))
    assert eng.batch_encode(docs) == orc.batch_encode(docs)
    eng.close()
