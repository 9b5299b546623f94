"""CPU suite, part 2: the engine's HOST side -- the loaders and derived device tables of hut_model.c (unit
ids, pair table (L,R)->merged, initial-symbol tables, prefix handling) and the local
pretoken-start rule of hut_classify.h -- walked by the scalar test driver tests/host_harness.c and compared
with the golden fixtures and the oracle.  No CUDA involved; the shipped library has no such CPU path."""
import json
import os
import random
import tempfile

import pytest

import fuzzmodels as F
import synth
from common import MODELS, ROOT, oracle_for
from harness import HostModel, pretoken_starts
from oracle import oracle as O

GOLDEN = os.path.join(ROOT, "tests", "golden")


@pytest.mark.parametrize("name", MODELS)
def test_tables_reproduce_golden(name):
    with open(os.path.join(GOLDEN, f"encode_{name}.json")) as f:
        g = json.load(f)
    v, s = synth.PATHS[name]
    hm = HostModel(v, s, **synth.MODEL_KW[name])
    assert hm.ok(), hm.message
    for c in g["cases"]:
        t = bytes.fromhex(c["text_hex"])
        if len(t) > 3000:
            continue  # the scalar driver is O(n^2) per pretoken
        assert hm.encode(t) == c["ids"], (name, t[:40])
    info = hm.info()
    assert info["n_units"] >= info["n_vocab"] > 0 and info["pair_count"] > 0
    assert info["pair_cap"] & (info["pair_cap"] - 1) == 0 and info["pair_count"] * 2 <= info["pair_cap"]


def test_local_start_rule_matches_parser():
    with open(os.path.join(GOLDEN, "pretokens.json")) as f:
        g = json.load(f)
    for c in g["cases"]:
        data = bytes.fromhex(c["text_hex"])
        want, pos = [], 0
        for n in c["lengths"]:
            want.append(pos)
            pos += n
        assert pretoken_starts(data) == want, data


def test_local_start_rule_random_bytes():
    """Exhaustive-ish: random byte soup over a small alphabet that exercises every class transition and every
    malformed UTF-8 shape, against the oracle's sequential parser."""
    rng = random.Random(11)
    alphabet = [0x20, 0x20, 0x41, 0x61, 0x31, 0x2E, 0x0A, 0x09, 0xC3, 0xA9, 0xC5, 0x91, 0xB1, 0xE2, 0x82, 0xAC, 0xF0,
                0x9F, 0x98, 0x82, 0x80, 0xFF, 0xC0, 0xE0, 0xF8, 0xED, 0xA0]
    for _ in range(4000):
        data = bytes(rng.choice(alphabet) for _ in range(rng.randint(1, 24)))
        want, pos = [], 0
        while True:
            n = O.next_pretoken(data, pos)
            if n == 0:
                break
            want.append(pos)
            pos += n
        assert pretoken_starts(data) == want, data.hex()


@pytest.mark.parametrize("byte_mode", [True, False])
def test_tables_on_fuzz_models(byte_mode):
    tmp = tempfile.mkdtemp()
    for seed in range(80):
        rng = random.Random(9000 + seed)
        v, s, kw = F.random_model(rng, tmp, byte_mode, f"h{seed}")
        hm = HostModel(v, s, **kw)
        assert hm.ok(), hm.message
        orc = O.Oracle(v, s, **kw)
        texts = [F.random_text(rng, rng.randint(0, 40)) for _ in range(60)]
        if byte_mode:
            texts += [bytes(rng.choice([0x80, 0xC3, 0xE2, 0xF0, 0x41, 0x20, 0xA9, 0xFF, 0xC0, 0x0A, 0x61])
                            for _ in range(rng.randint(1, 40))) for _ in range(20)]
        for t in texts:
            assert hm.encode(t) == orc.encode(t), (seed, kw, t)
        hm.close()
        orc.close()


def test_loader_errors_match_oracle():
    """Same failure class and message text as the oracle's restatement of p_initialize."""
    tmp = tempfile.mkdtemp()
    good_v = os.path.join(tmp, "v.txt")
    good_s = os.path.join(tmp, "s.txt")
    with open(good_v, "w") as f:
        f.write("0x61 == 0\n0x62 == 1\n0x610x62 == 2\n")
    open(good_s, "w").close()
    assert HostModel(good_v, good_s).ok()
    cases = {
        "empty": ("", None),
        "noeq": ("0x61 = 0\n", None),
        "nul": ("0x00 == 0\n", None),
        "longspecial": ("0x61 == 0\n", "32 == " + "x" * 40 + "\n"),
        "badbyte": ("0x61 == 0\n", "999 == x\n"),
    }
    for key, (vt, st) in cases.items():
        vp = os.path.join(tmp, f"{key}_v.txt")
        with open(vp, "w") as f:
            f.write(vt)
        sp = good_s
        if st is not None:
            sp = os.path.join(tmp, f"{key}_s.txt")
            with open(sp, "w") as f:
                f.write(st)
        hm = HostModel(vp, sp)
        with pytest.raises(O.OracleError) as e:
            O.Oracle(vp, sp)
        assert not hm.ok(), key
        want_type, want_msg = str(e.value).split(": ", 1)
        assert hm.message == want_msg, (key, hm.message, want_msg)
        assert {"FileNotFoundError": 1, "ValueError": 2, "MemoryError": 3}[want_type] == hm.status, key
    hm = HostModel(os.path.join(tmp, "missing.txt"), good_s)
    assert hm.status == 1
    hm = HostModel(good_v, os.path.join(tmp, "missing_s.txt"))
    assert hm.status == 1


# ---- bit-sliced classifier (hut_bitslice.h): the fused tile kernel's fast path ----
def _oracle_starts(data, doc_starts):
    """pretoken starts of packed documents by the oracle's sequential parser, document by document"""
    bounds = sorted(set(doc_starts)) + [len(data)]
    out = []
    for a, b in zip(bounds, bounds[1:]):
        pos = 0
        doc = data[a:b]
        while True:
            n = O.next_pretoken(doc, pos)
            if n == 0:
                break
            out.append(a + pos)
            pos += n
    return out


def test_bitslice_transpose():
    from harness import bitslice_ok
    rng = random.Random(1)
    for _ in range(300):
        assert bitslice_ok(bytes(rng.randrange(256) for _ in range(32)))


def test_bitsliced_rule_matches_parser_on_golden():
    from harness import pretoken_starts_bitsliced
    with open(os.path.join(GOLDEN, "pretokens.json")) as f:
        g = json.load(f)
    for c in g["cases"]:
        data = bytes.fromhex(c["text_hex"])
        if b"\0" in data:
            continue
        want, pos = [], 0
        for n in c["lengths"]:
            want.append(pos)
            pos += n
        got, exotic = pretoken_starts_bitsliced(data)
        if not exotic:
            assert got == want, data


def test_bitsliced_rule_random_bytes_and_documents():
    """Byte soup with every class transition / malformed UTF-8 shape, packed as several documents whose
    boundaries fall anywhere (also inside multi-byte characters), compared with the sequential parser.
    Over-long forms must be FLAGGED (the kernel then switches to hut_classify.h); everything else must match."""
    from harness import pretoken_starts_bitsliced
    rng = random.Random(12)
    alphabet = [0x20, 0x20, 0x41, 0x61, 0x7A, 0x31, 0x2E, 0x0A, 0x09, 0x7F, 0x01, 0xC3, 0xA9, 0xC5, 0x91, 0xB1, 0xE2,
                0x82, 0xAC, 0xF0, 0x9F, 0x98, 0x82, 0x80, 0xFF, 0xC0, 0xC1, 0xE0, 0xF8, 0xED, 0xA0, 0xC2, 0xF4, 0x90,
                0xBC, 0x96, 0x9C, 0x8D, 0xB6]
    n_exotic = n_plain = 0
    for it in range(6000):
        n = rng.randint(1, 90)
        if it % 3 == 0:  # valid text: never exotic
            data = "".join(rng.choice("aábc őŰ1 2.€😂\n\tZé  ñ日") for _ in range(n // 2 + 1)).encode("utf-8")
        else:
            data = bytes(rng.choice(alphabet) for _ in range(n))
        ds = sorted({0} | {rng.randrange(len(data)) for _ in range(rng.randint(0, 3))})
        got, exotic = pretoken_starts_bitsliced(data, ds)
        if it % 3 == 0:
            valid_cut = all((data[p] & 0xC0) != 0x80 for p in ds)
            assert not exotic or not valid_cut
        if exotic:
            n_exotic += 1
            continue
        n_plain += 1
        assert got == _oracle_starts(data, ds), (data.hex(), ds)
    assert n_plain > 3000 and n_exotic > 50


def test_bitsliced_rule_on_corpus():
    from harness import pretoken_starts_bitsliced
    data, offs = synth.corpus(55, 1 << 18, special_every=150)
    b = data.tobytes()
    ds = [int(x) for x in offs[:-1]]
    got, exotic = pretoken_starts_bitsliced(b, ds)
    assert not exotic
    assert got == _oracle_starts(b, ds)
