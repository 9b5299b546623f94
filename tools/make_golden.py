"""Generates tests/golden/*.json from the UNMODIFIED reference (oracle/_ref/_hutoken*.so, built by
oracle/Makefile from /root/reference/src).  Run in the dev container only (the reference sources do not exist
on the GPU box); the fixtures are committed and are what pins the CPU oracle -- and through it the CUDA
engine -- to the reference's behaviour (the reference ships no golden token vectors).

    python tools/make_golden.py

Fixtures:
  encode_<model>.json   {"model", "kwargs", "cases": [{"text_hex", "ids"}]}   ids = _hutoken.encode / batch_encode
  encode_merges_<model>.json   the same with merges_file_path= (id-keyed merge path), small models only
  decode_<model>.json   {"cases": [{"ids", "text_hex"} | {"ids", "error", "message"}]}   _hutoken.decode / batch_decode
  pretokens.json        {"cases": [{"text_hex", "lengths"}]}                   parser_next_token (parser.c:24-88)

Input strings: the 28 parser cases of the reference's tests/test_parser.c:77-108, the sentences / paragraphs of
tests/test_tokenizer.py:9-40, the probes of SURVEY.md Appendix B, and seeded synthetic documents.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")]

import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

PARSER_CASES = [
    "", "Hello", " Hello", "árvíztűrő", " árvíztűrő", "ÁrvíztűrőTükör", "12345", " 123", "!!!@#$", " !!!",
    "!@#$%^&*()_+", "€", "😂", " ", "   ", "word123", "word.!_", "123word", "123.!", ".!word", ".!123",
    "\t\n\r\f\v", "A  B   C", "word ", " First Second", "Hello world 123. End!", "árvíztűrő tükörfúrógép.",
    "This is a test 123. With some special chars: !@# and spaces. árvíztűrő tükörfúrógép!",
]

SENTENCES = [
    "How can the net amount of entropy of the universe be massively decreased?",
    "How can the net", " amount of entropy of", " the universe be massively decreased?",
    "What I cannot create, I do not understand.", "What I cannot", " create, I do", " not understand.",
    "Gorcsev Iván, a Rangoon teherhajó matróza még huszonegy éves sem volt, midőn elnyerte a fizikai Nobel-díjat."
    "Ilyen nagy jelentőségű tudományos jutalmat e poétikusan ifjú korban megszerezni példátlan nagyszerű "
    "teljesítmény, még akkor is, ha egyesek előtt talán szépséghibának tűnik majd, hogy Gorcsev Iván a fizikai "
    "Nobel-díjat a makao nevű kártyajátékon nyerte el, Noah Bertinus professzortól.",
    "Hogyan hatott rátok, athéni férfiak, vádlóim beszéde, nem tudom; én bizony magam is kis híján beléjük "
    "feledkeztem, olyan meggyőzően beszéltek. Ámbár igazat úgyszólván semmit nem mondtak.",
]

PROBES = [
    "hello world", "aaa", "aaaa", "a  b", "\n", "é", "én", " hello", "<0x41>", "ő", "señor", "ñ!", "a \n b",
    "a   ", "1a.b2 c", "<|endoftext|>", "x\n\ny\t z", "€ 😂 日本", "\x01\x02\x7f", "tab\there",
    "line1\nline2\r\nline3", "  leading", "trailing  ", "don't stop", "mixed123abc456", "...!!!???", " . , ; ",
    "őő űű", "a" * 40, "ab" * 100, "é" * 50, " " * 70, "\n" * 33, "word " * 300, "ÁRVÍZTŰRŐ TÜKÖRFÚRÓGÉP",
    "x" * 4097, ("w" * 17 + " ") * 200,
]


def _hex(s) -> str:
    return (s.encode("utf-8") if isinstance(s, str) else bytes(s)).hex()


def synthetic_docs(seed, total, special_every):
    data, offs = synth.corpus(seed, total, special_every=special_every)
    return [d.decode("utf-8") for d in synth.docs_of(data, offs)]


def fuzz_texts(seed, n):
    import fuzzmodels as F
    rng = random.Random(seed)
    return [F.random_text(rng, rng.randint(0, 40)) for _ in range(n)]


def encode_golden(ref, name, merges=None):
    v, s = synth.PATHS[name]
    kw = synth.MODEL_KW[name]
    if merges:  # id-keyed path (lib.c:573-663, core.c:211-337)
        ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"], merges_file_path=merges)
    else:
        ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])
    texts = PARSER_CASES + SENTENCES + PROBES + fuzz_texts(77, 120)
    texts += synthetic_docs(11, 48 << 10, 0)[:12] + synthetic_docs(12, 48 << 10, 200)[:12]
    cases = []
    singles = [ref.encode(t) for t in texts]
    for nt in (1, 3, 8):  # batch_encode must agree with encode for every thread count (lib.c:779-794)
        assert ref.batch_encode(texts, nt) == singles, (name, nt)
    for t, ids in zip(texts, singles):
        cases.append({"text_hex": _hex(t), "ids": ids})
    # behaviours of the boundary itself
    extra = {
        "nul_cut": {"text_hex": _hex("ab\0cd"), "ids": ref.batch_encode(["ab\0cd"], 1)[0]},
        "zero_threads": ref.batch_encode(["hello", "world"], 0),
    }
    out = {"model": name, "kwargs": kw, "vocab": os.path.relpath(v, ROOT), "special": os.path.relpath(s, ROOT),
           "cases": cases, "extra": extra}
    if merges:
        out["merges"] = os.path.relpath(merges, ROOT)
    return out


def decode_golden(ref, name):
    """_hutoken.decode / batch_decode (lib.c:876-1085) on ids the reference itself produced, on random valid ids
    (byte-level pieces that do not form UTF-8 raise UnicodeDecodeError in PyUnicode_FromString) and on bad ids."""
    v, s = synth.PATHS[name]
    kw = synth.MODEL_KW[name]
    ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])
    texts = PARSER_CASES + SENTENCES + PROBES[:-2] + fuzz_texts(79, 60) + synthetic_docs(13, 32 << 10, 200)[:8]
    n_lines = sum(1 for _ in open(v, "rb"))
    rng = random.Random(17)
    id_lists = [[i for i in ref.encode(t) if i >= 0] for t in texts]
    id_lists += [[rng.randrange(n_lines) for _ in range(rng.randint(1, 30))] for _ in range(60)]
    id_lists += [[], [0], [n_lines - 1], [n_lines], [-1], [5, n_lines + 7, 3]]
    cases = []
    for ids in id_lists:
        try:
            cases.append({"ids": ids, "text_hex": _hex(ref.decode(ids))})
        except Exception as e:  # ValueError (bad id) or UnicodeDecodeError
            cases.append({"ids": ids, "error": type(e).__name__, "message": str(e)[:80]})
    ok = [c["ids"] for c in cases if "text_hex" in c and c["ids"]]
    for nt in (1, 4):
        assert [_hex(t) for t in ref.batch_decode(ok, nt)] == [c["text_hex"] for c in cases if "text_hex" in c and c["ids"]]
    return {"model": name, "kwargs": kw, "vocab_lines": n_lines, "cases": cases}


class _Slice(C.Structure):
    _fields_ = [("start", C.c_void_p), ("length", C.c_size_t)]


class _State(C.Structure):
    _fields_ = [("current_pos", C.c_void_p)]


def pretoken_golden():
    """parser_init / parser_next_token called directly in the reference .so (parser.h:7-17)."""
    lib = C.CDLL(O.ref_path())
    lib.parser_init.restype = _State
    lib.parser_init.argtypes = [C.c_char_p]
    lib.parser_next_token.restype = C.c_bool
    lib.parser_next_token.argtypes = [C.POINTER(_State), C.POINTER(_Slice)]
    texts = PARSER_CASES + PROBES + fuzz_texts(78, 200)
    rng = random.Random(5)
    # invalid UTF-8 (parser.c:144-183 rejects it byte by byte)
    texts += [bytes(rng.choice([0x80, 0xC3, 0xE2, 0xF0, 0x41, 0x20, 0xA9, 0xFF, 0xC0, 0x0A, 0x61, 0x91, 0xC5, 0xB1])
                    for _ in range(rng.randint(1, 40))) for _ in range(150)]
    cases = []
    for t in texts:
        data = t.encode("utf-8") if isinstance(t, str) else bytes(t)
        buf = C.create_string_buffer(data, len(data) + 1)
        st = lib.parser_init(C.cast(buf, C.c_char_p))
        sl = _Slice()
        lens = []
        base = C.addressof(buf)
        pos = 0
        while lib.parser_next_token(C.byref(st), C.byref(sl)):
            assert sl.start - base == pos
            lens.append(sl.length)
            pos += sl.length
        cases.append({"text_hex": data.hex(), "lengths": lens})
    return {"cases": cases}


def main():
    O.build(ref=True)
    ref = O.load_ref()
    assert ref is not None, "reference .so not built"
    os.makedirs(GOLDEN, exist_ok=True)
    for name in ("gpt2_small", "llama_small", "gpt2", "llama"):
        g = encode_golden(ref, name)
        with open(os.path.join(GOLDEN, f"encode_{name}.json"), "w") as f:
            json.dump(g, f, separators=(",", ":"))
        print(name, len(g["cases"]), "cases,", sum(len(c["ids"]) for c in g["cases"]), "ids")
    for name in ("gpt2_small", "llama_small"):  # merges-file mode; data/*_2k.merges.txt come from tools/make_merges.py
        g = encode_golden(ref, name, merges=synth.MERGES[name])
        with open(os.path.join(GOLDEN, f"encode_merges_{name}.json"), "w") as f:
            json.dump(g, f, separators=(",", ":"))
        print(name, "merges:", len(g["cases"]), "cases,", sum(len(c["ids"]) for c in g["cases"]), "ids")
    for name in ("gpt2_small", "llama_small", "gpt2", "llama"):
        g = decode_golden(ref, name)
        with open(os.path.join(GOLDEN, f"decode_{name}.json"), "w") as f:
            json.dump(g, f, separators=(",", ":"))
        print(name, "decode:", len(g["cases"]), "cases,", sum("error" in c for c in g["cases"]), "errors")
    p = pretoken_golden()
    with open(os.path.join(GOLDEN, "pretokens.json"), "w") as f:
        json.dump(p, f, separators=(",", ":"))
    print("pretokens", len(p["cases"]), "cases")


if __name__ == "__main__":
    main()
