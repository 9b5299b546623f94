"""CPU suite, part 1: pins the oracle (oracle/hut_oracle.c) to the reference.

  * against the committed golden fixtures (tests/golden/, generated from the UNMODIFIED reference .so by
    tools/make_golden.py) -- runs everywhere, including the GPU box where /root/reference does not exist;
  * against the reference .so itself on fresh random inputs and adversarial random vocabularies, when
    oracle/_ref holds a build (dev container, and the GPU box because the built .so travels).
"""
import json
import os
import random
import tempfile

import numpy as np
import pytest

import fuzzmodels as F
import synth
from common import MODELS, ROOT, oracle_for
from oracle import oracle as O

GOLDEN = os.path.join(ROOT, "tests", "golden")


def _golden(name):
    with open(os.path.join(GOLDEN, f"encode_{name}.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("name", MODELS)
def test_oracle_matches_golden_encode(name):
    g = _golden(name)
    orc = oracle_for(name)
    texts = [bytes.fromhex(c["text_hex"]) for c in g["cases"]]
    for t, c in zip(texts, g["cases"]):
        assert orc.encode(t) == c["ids"], (name, t[:40])
    # the batch form agrees with the per-document form
    assert orc.batch_encode(texts) == [c["ids"] for c in g["cases"]]
    assert orc.encode(bytes.fromhex(g["extra"]["nul_cut"]["text_hex"])) == g["extra"]["nul_cut"]["ids"]


def test_oracle_matches_golden_pretokens():
    """parser_next_token (parser.c:24-88) on the reference's 28 parser cases + probes + invalid UTF-8."""
    with open(os.path.join(GOLDEN, "pretokens.json")) as f:
        g = json.load(f)
    for c in g["cases"]:
        data = bytes.fromhex(c["text_hex"])
        lens, pos = [], 0
        while True:
            n = O.next_pretoken(data, pos)
            if n == 0:
                break
            lens.append(n)
            pos += n
        assert lens == c["lengths"], data


def test_parser_cases_match_reference_regex():
    """tests/test_parser.c:10-12 pins the parser to a POSIX regex with a 1-byte fallback; same check here
    with Python's re on the same 28 strings (the only executable pretokenisation spec the reference has)."""
    import re
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from make_golden import PARSER_CASES
    hu = "áéíóúőűüöÁÉÍÓÚŐÜŰÖ"
    rx = re.compile(f"[ ]?[A-Za-z{hu}]+|[ ]?[0-9]+|[ ]?[^\\sA-Za-z{hu}0-9]+|[ ]+", re.ASCII)
    orc = oracle_for("gpt2_small")
    for t in PARSER_CASES:
        want, pos = [], 0
        while pos < len(t):
            m = rx.match(t, pos)
            if m and m.end() > pos:
                want.append(t[pos:m.end()])
                pos = m.end()
            else:
                want.append(t[pos])
                pos += 1
        assert [p.decode("utf-8") for p in orc.pretokens(t)] == want, t


needs_ref = pytest.mark.skipif(O.ref_path() is None and not os.path.isdir("/root/reference/src"),
                               reason="reference build (oracle/_ref) not available")


@needs_ref
@pytest.mark.parametrize("name", MODELS)
def test_oracle_matches_reference_on_corpus(name):
    ref = O.load_ref()
    v, s = synth.PATHS[name]
    kw = synth.MODEL_KW[name]
    ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])
    orc = oracle_for(name)
    for seed, special_every in ((101, 0), (102, 200)):
        data, offs = synth.corpus(seed, 1 << 20, special_every=special_every)
        docs = [d.decode("utf-8") for d in synth.docs_of(data, offs)]
        want = ref.batch_encode(docs, 4)
        ids, tok = orc.encode_packed(data, offs)
        flat = np.fromiter((x for w in want for x in w), dtype=np.int32)
        assert np.array_equal(flat, ids)
        assert tok.tolist() == [0] + np.cumsum([len(w) for w in want]).tolist()


@needs_ref
@pytest.mark.parametrize("byte_mode", [True, False])
def test_oracle_matches_reference_on_fuzz_models(byte_mode):
    ref = O.load_ref()
    tmp = tempfile.mkdtemp()
    for seed in range(60):
        rng = random.Random(5000 + seed)
        v, s, kw = F.random_model(rng, tmp, byte_mode, f"m{seed}")
        ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])
        orc = O.Oracle(v, s, **kw)
        texts = [F.random_text(rng, rng.randint(0, 60)) for _ in range(60)]
        want = ref.batch_encode(texts, 2)
        got = orc.batch_encode(texts)
        for t, g, w in zip(texts, got, want):
            assert g == w, (seed, kw, t, g, w)
        orc.close()


@needs_ref
def test_oracle_matches_reference_long_pretokens():
    """Inside the reference's defined domain: <= 262144 bytes and well under 80k symbols,
    plus the silent document cut just above the limit (core.c:402-407,503)."""
    ref = O.load_ref()
    name = "gpt2_small"
    v, s = synth.PATHS[name]
    kw = synth.MODEL_KW[name]
    ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])
    orc = oracle_for(name)
    rng = random.Random(3)
    for n in (1000, 16384, 60000):
        t = "go " + "".join(rng.choice("abcdeégőz") for _ in range(n)) + " end"
        assert orc.encode(t) == ref.encode(t), n
    t = "keep this " + "a" * (262144 + 8) + " dropped tail"
    want = ref.encode(t)
    assert orc.encode(t) == want and len(want) == len(ref.encode("keep this"))


def test_oracle_load_errors():
    """Loader messages of p_initialize (lib.c:248, 277-355, 466-543)."""
    tmp = tempfile.mkdtemp()
    sp = os.path.join(tmp, "s.txt")
    open(sp, "w").close()

    def err(vocab_text, special_text=None):
        vp = os.path.join(tmp, "v.txt")
        with open(vp, "w") as f:
            f.write(vocab_text)
        sp2 = sp
        if special_text is not None:
            sp2 = os.path.join(tmp, "s2.txt")
            with open(sp2, "w") as f:
                f.write(special_text)
        with pytest.raises(O.OracleError) as e:
            O.Oracle(vp, sp2)
        return str(e.value)

    assert "Vocab file is empty" in err("")
    assert "Invalid format in vocab file" in err("0x61 = 0\n")
    assert "Invalid format in special character file" in err("0x61 == 0\n", "32 == " + "x" * 40 + "\n")
    with pytest.raises(O.OracleError):
        O.Oracle(os.path.join(tmp, "missing.txt"), sp)
