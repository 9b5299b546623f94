"""Vocab file tooling: tools/convert_vocab.py against the reference's file format
(lib.c:243-378), scripts/convert.py's numbering, and the bpe_train post-processing SURVEY F6 documents."""
import json
import os
import subprocess
import sys
import tempfile

import pytest

import synth
from common import ROOT
from oracle import oracle as O

sys.path.insert(0, os.path.join(ROOT, "tools"))
import convert_vocab as C  # noqa: E402


@pytest.mark.parametrize("name", ["gpt2_small", "llama_small"])
def test_json_to_vocab_reproduces_the_file(name):
    """vocab.txt -> vocab.json (token string -> id, file order) -> convert: the same bytes as the original file."""
    v, _ = synth.PATHS[name]
    rows = C.parse(v)
    tmp = tempfile.mkdtemp()
    jp, out = os.path.join(tmp, "vocab.json"), os.path.join(tmp, "vocab.txt")
    with open(jp, "w", encoding="utf-8") as f:
        json.dump({k.decode("utf-8"): i for k, i in rows}, f, ensure_ascii=False)
    assert C.from_json(jp, out) == len(rows)
    assert open(out, "rb").read() == open(v, "rb").read()
    # --keep-ids: the JSON's own ids, sorted; shuffled key order must give the same file
    with open(jp, "w", encoding="utf-8") as f:
        json.dump({k.decode("utf-8"): i for k, i in reversed(rows)}, f, ensure_ascii=False)
    C.from_json(jp, out, keep_ids=True)
    assert open(out, "rb").read() == open(v, "rb").read()
    assert C.check(out) == []


def test_fix_trained_output_loads_and_encodes():
    """A bpe_train-shaped file (0x00 == 0 first, id 256 skipped, helper.c:175-177 / bpe.c:171) is refused as it is
    and loads after `fix`; ids are dense in file order."""
    tmp = tempfile.mkdtemp()
    src, out = os.path.join(tmp, "trained.txt"), os.path.join(tmp, "fixed.txt")
    toks = [bytes([b]) for b in range(0, 128)] + [b"th", b"the", b"he", b"in", b" t", b" th", b" the"]
    with open(src, "w") as f:
        for n, t in enumerate(toks):
            f.write(f"{C.hex_key(t)} == {n if n < 128 else n + 129}\n")  # the trainer's ids jump over 256
    assert any("0x00" in p for p in C.check(src))
    assert C.fix_trained(src, out) == len(toks) - 1
    assert C.check(out) == []
    assert [i for _, i in C.parse(out)] == list(range(len(toks) - 1))
    empty = os.path.join(tmp, "special.txt")
    open(empty, "w").close()
    with pytest.raises(O.OracleError):
        O.Oracle(src, empty, prefix=None, is_byte_encoder=False)
    orc = O.Oracle(out, empty, prefix=None, is_byte_encoder=False)
    ids = orc.encode("the thin")
    assert -1 not in ids and orc.decode_bytes(ids) == b"the thin"
    fixed = dict(C.parse(out))
    assert ids[0] == fixed[b"the"] and ids[1] == fixed[b" th"]
    orc.close()


def test_cli_and_bad_lines():
    tmp = tempfile.mkdtemp()
    bad = os.path.join(tmp, "bad.txt")
    with open(bad, "w") as f:
        f.write("0x41 == 0\nnot a vocab line\n")
    with pytest.raises(ValueError, match="bad.txt:2"):
        C.parse(bad)
    sparse = os.path.join(tmp, "sparse.txt")
    with open(sparse, "w") as f:
        f.write("0x41 == 0\n0x42 == 7\n")
    assert any("largest id 7" in p for p in C.check(sparse))
    tool = os.path.join(ROOT, "tools", "convert_vocab.py")
    r = subprocess.run([sys.executable, tool, "check", sparse], capture_output=True, text=True)
    assert r.returncode == 1 and "largest id 7" in r.stdout
    r = subprocess.run([sys.executable, tool, "check", synth.PATHS["gpt2_small"][0]], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout == ""
