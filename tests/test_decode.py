"""Decode path: decode() core.c:513-581 + pretokenizer_decode pretokenizer.c:197-296.

CPU part: the oracle's restatement against golden fixtures made from the unmodified reference (tests/golden/
decode_*.json, tools/make_golden.py) and against the reference .so on random models; host trie / id tables.
GPU part (-m gpu): hut_decode_batch through the C-ABI against the oracle and the fixtures, bit-exact bytes.
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
BAD_ID = "Element must be non-negative and less than vocab size."

needs_ref = pytest.mark.skipif(O.ref_path() is None and not os.path.isdir("/root/reference/src"),
                               reason="reference build (oracle/_ref) not available")


def _golden(name):
    with open(os.path.join(GOLDEN, f"decode_{name}.json")) as f:
        return json.load(f)


def _check_case(decode_bytes, c, what):
    """decode_bytes(ids) -> raw bytes or raises ValueError; the binding's cut at 0x00 + strict UTF-8 applied here."""
    if c.get("error") == "ValueError":
        with pytest.raises(ValueError, match="less than vocab size"):
            decode_bytes(c["ids"])
        return
    raw = decode_bytes(c["ids"]).split(b"\0", 1)[0]
    if c.get("error") == "UnicodeDecodeError":
        with pytest.raises(UnicodeDecodeError):
            raw.decode("utf-8")
    else:
        assert raw == bytes.fromhex(c["text_hex"]), (what, c["ids"][:12])


@pytest.mark.parametrize("name", MODELS)
def test_oracle_decode_matches_golden(name):
    orc = oracle_for(name)
    for c in _golden(name)["cases"]:
        _check_case(orc.decode_bytes, c, name)


@pytest.mark.parametrize("name", MODELS)
def test_oracle_round_trip(name):
    """decode(encode(t)) == t whenever every symbol is in the vocabulary (and, with a prefix, t has no leading
    space: the prefix is stripped only once, pretokenizer.c:209-215)."""
    orc = oracle_for(name)
    data, offs = synth.corpus(51, 256 << 10)
    n = 0
    for d in synth.docs_of(data, offs):
        ids = orc.encode(d)
        if -1 in ids or bytes(d[:1]) == b" ":
            continue
        assert orc.decode_bytes(ids) == bytes(d)
        n += 1
    assert n > 10


@needs_ref
@pytest.mark.parametrize("byte_mode", [True, False])
def test_oracle_decode_matches_reference_on_fuzz_models(byte_mode):
    """Random vocabularies: tokens cut inside multi-byte characters, replacement strings that are prefixes of one
    another or of ordinary tokens, code points above 255 ('?'), the prefix at the start only."""
    ref = O.load_ref()
    tmp = tempfile.mkdtemp()
    checked = 0
    for seed in range(40):
        rng = random.Random(7000 + seed)
        v, s, kw = F.random_model(rng, tmp, byte_mode, f"d{seed}")
        ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])
        orc = O.Oracle(v, s, **kw)
        n_lines = sum(1 for _ in open(v, "rb"))
        usable = _unambiguous_ids(v, n_lines)
        for _ in range(60):
            if rng.random() < 0.5:
                ids = [i for i in orc.encode(F.random_text(rng, rng.randint(0, 30))) if i >= 0]
            else:
                ids = [rng.randrange(n_lines) for _ in range(rng.randint(0, 25))]
            if any(i not in usable for i in ids):
                # an id no line carries is an uninitialised pointer in the reference (lib.c:396-433); an id shared by
                # several keys decodes to whichever its hash map iterates last (lib.c:423-427).  Neither is defined.
                continue
            try:
                want = ref.decode(ids).encode("utf-8")
            except UnicodeDecodeError:
                want = None
            got = orc.decode_bytes(ids).split(b"\0", 1)[0]
            if want is None:
                with pytest.raises(UnicodeDecodeError):
                    got.decode("utf-8")
            else:
                assert got == want, (seed, kw, ids)
            checked += 1
        orc.close()
    assert checked > 1000


@pytest.mark.parametrize("name", MODELS)
def test_host_decode_tables_match_golden(name):
    """The tables the device reads (id -> bytes pool, trie edges / outputs / depths, prefix), walked on the CPU."""
    import harness
    v, s = synth.PATHS[name]
    hm = harness.HostModel(v, s, **synth.MODEL_KW[name])
    assert hm.ok(), hm.message
    for c in _golden(name)["cases"]:
        _check_case(hm.decode_bytes, c, name)


@pytest.mark.parametrize("byte_mode", [True, False])
def test_host_decode_tables_match_oracle_on_fuzz_models(byte_mode):
    import harness
    tmp = tempfile.mkdtemp()
    for seed in range(40):
        rng = random.Random(7100 + seed)
        v, s, kw = F.random_model(rng, tmp, byte_mode, f"hd{seed}")
        hm = harness.HostModel(v, s, **kw)
        if not hm.ok():
            continue
        orc = O.Oracle(v, s, **kw)
        n_lines = sum(1 for _ in open(v, "rb"))
        for _ in range(60):
            ids = [rng.randrange(n_lines) for _ in range(rng.randint(0, 25))]
            try:
                want = orc.decode_bytes(ids)
            except ValueError:
                with pytest.raises(ValueError):
                    hm.decode_bytes(ids)
                continue
            assert hm.decode_bytes(ids) == want, (seed, kw, ids)
        hm.close()
        orc.close()


def _unambiguous_ids(vocab_path, n_lines):
    import collections
    import make_merges
    cnt = collections.Counter(make_merges.read_vocab(vocab_path).values())
    return {i for i, c in cnt.items() if c == 1 and i < n_lines}


def _has_id(orc, i):
    try:
        orc.decode_bytes([i])
        return True
    except ValueError:
        return False


# ---------------------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", MODELS)
def test_gpu_decode_golden(name):
    from common import engine_for
    eng = engine_for(name)

    def dec(ids):
        return eng.batch_decode_bytes([ids])[0]

    for c in _golden(name)["cases"]:
        _check_case(dec, c, name)
    # the whole fixture as one batch
    ok = [c for c in _golden(name)["cases"] if "text_hex" in c]
    got = eng.batch_decode_bytes([c["ids"] for c in ok])
    for g, c in zip(got, ok):
        assert g.split(b"\0", 1)[0] == bytes.fromhex(c["text_hex"])
    assert eng.decode_stats()["kernel_launches"] >= 1


@pytest.mark.gpu
@pytest.mark.parametrize("name", MODELS)
def test_gpu_decode_corpus_round_trip(name):
    """encode on the GPU, decode on the GPU: equal to the oracle's decode of the same ids, and to the input text
    for documents without unknown symbols.  Documents span many 1024-byte tiles."""
    from common import engine_for
    eng, orc = engine_for(name), oracle_for(name)
    data, offs = synth.corpus(61, 2 << 20, special_every=300)
    ids, tok = eng.encode_packed(data, offs)
    keep = ids >= 0
    # drop the -1 ids (undecodable by definition) and rebuild the offsets
    cum = np.concatenate([[0], np.cumsum(keep)]).astype(np.uint64)
    ids2, tok2 = ids[keep], cum[tok.astype(np.int64)]
    text, toff = eng.decode_packed(ids2, tok2)
    docs = synth.docs_of(data, offs)
    exact = 0
    for d in range(len(docs)):
        got = text[int(toff[d]):int(toff[d + 1])].tobytes()
        doc_ids = ids2[int(tok2[d]):int(tok2[d + 1])]
        assert got == orc.decode_bytes(doc_ids), (name, d)
        src = bytes(docs[d])
        if len(doc_ids) == int(tok[d + 1] - tok[d]) and src[:1] != b" ":
            assert got == src, (name, d)
            exact += 1
    assert exact > 20
    # one very long document (thousands of tiles handed over one by one) + empty documents in between
    big = np.concatenate([ids2[:200000], ids2[:200000]])
    tokb = np.array([0, 0, len(big), len(big)], dtype=np.uint64)
    text, toff = eng.decode_packed(big, tokb)
    assert toff[1] == 0 and toff[3] == toff[2]
    assert text.tobytes() == orc.decode_bytes(big)


@pytest.mark.gpu
@pytest.mark.parametrize("name", MODELS)
def test_gpu_decode_streamed_chunks(name):
    """hut_decode_batch streams a large host batch through the device in chunks of whole documents: forced here with
    tiny chunks (a few hundred ids), the text and offsets equal those of the same call in one piece.  The batch holds
    empty documents, documents larger than a chunk, and a non-zero first offset."""
    from common import engine_for
    eng = engine_for(name)
    data, offs = synth.corpus(67, 1 << 20, special_every=500)
    ids, tok = eng.encode_packed(data, offs)
    keep = ids >= 0
    cum = np.concatenate([[0], np.cumsum(keep)]).astype(np.uint64)
    ids2, tok2 = ids[keep], cum[tok.astype(np.int64)]
    nd = len(tok2) - 1
    # empty documents sprinkled in (repeated offsets); the batch starts at document 3 (first offset > 0)
    tok3 = np.sort(np.concatenate([tok2[3:], tok2[10:nd:7], tok2[-1:]])).astype(np.uint64)
    whole_text, whole_off = eng.decode_packed(ids2, tok3)
    one_piece = eng.decode_stats()
    assert one_piece["path"] == "token-wise"
    for chunk in ("300", "5000", "100000"):
        os.environ["HUTOKEN_B200_DECODE_CHUNK_IDS"] = chunk
        try:
            text, toff = eng.decode_packed(ids2, tok3)
        finally:
            del os.environ["HUTOKEN_B200_DECODE_CHUNK_IDS"]
        assert np.array_equal(toff, whole_off), (name, chunk)
        assert text.tobytes() == whole_text.tobytes(), (name, chunk)
        assert eng.decode_stats()["kernel_launches"] > one_piece["kernel_launches"]  # several chunks ran


@pytest.mark.gpu
def test_gpu_decode_streamed_over_devices():
    """A context over all distinct GPUs of the box deals the chunks of a streamed hut_decode_batch round-robin to its
    devices (every device holds its own decode tables): text and offsets equal those of a one-device context.  The
    same context, same forced chunk size, with one device listed twice and three times runs on any box; the distinct
    devices need >= 2 GPUs."""
    import torch
    from common import engine_for
    n = torch.cuda.device_count()
    name = "llama_small"
    one = engine_for(name)
    data, offs = synth.corpus(71, 2 << 20, special_every=400)
    ids, tok = one.encode_packed(data, offs)
    keep = ids >= 0
    cum = np.concatenate([[0], np.cumsum(keep)]).astype(np.uint64)
    ids2, tok2 = ids[keep], cum[tok.astype(np.int64)]
    want_text, want_off = one.decode_packed(ids2, tok2)
    lists = [(0, 0), (0, 0, 0)] + ([tuple(range(n))] if n >= 2 else [])
    for devices in lists:
        eng = engine_for(name, devices=devices)
        for chunk in ("700", "20000"):
            os.environ["HUTOKEN_B200_DECODE_CHUNK_IDS"] = chunk
            try:
                text, toff = eng.decode_packed(ids2, tok2)
            finally:
                del os.environ["HUTOKEN_B200_DECODE_CHUNK_IDS"]
            assert np.array_equal(toff, want_off), (devices, chunk)
            assert text.tobytes() == want_text.tobytes(), (devices, chunk)


@pytest.mark.gpu
@pytest.mark.parametrize("byte_mode", [True, False])
def test_gpu_decode_fuzz_models(byte_mode):
    from hutoken_b200 import capi
    tmp = tempfile.mkdtemp()
    paths = {"token-wise": 0, "general": 0, None: 0}
    for seed in range(30):
        rng = random.Random(7000 + seed)
        v, s, kw = F.random_model(rng, tmp, byte_mode, f"gd{seed}")
        eng = capi.Engine(v, s, **kw)
        orc = O.Oracle(v, s, **kw)
        n_lines = eng.vocab_lines
        lists = []
        for _ in range(80):
            if rng.random() < 0.5:
                lists.append([i for i in orc.encode(F.random_text(rng, rng.randint(0, 30))) if i >= 0])
            else:
                lists.append([rng.randrange(n_lines) for _ in range(rng.randint(0, 25))])
        usable = _unambiguous_ids(v, n_lines)  # the engine and the oracle agree on the others too (last line wins)
        lists = [x for x in lists if all(_has_id(orc, i) for i in x)]
        assert sum(all(i in usable for i in x) for x in lists) > 10
        got = eng.batch_decode_bytes(lists)
        for x, g in zip(lists, got):
            assert g == orc.decode_bytes(x), (seed, kw, x)
        # one list per call: a batch without a boundary-straddling token takes the token-wise kernels
        for x in lists[:40]:
            assert eng.batch_decode_bytes([x])[0] == orc.decode_bytes(x), (seed, kw, x)
            paths[eng.decode_stats()["path"].split(",")[0]] += 1
        with pytest.raises(ValueError, match="less than vocab size"):
            eng.batch_decode_bytes([[0], [n_lines]])
        eng.close()
        orc.close()
    assert paths["token-wise"] > 50 and paths["general"] > 10, paths


def _general_engine(v, s, **kw):
    """An engine kept to the general decode kernels (the switch is read when the decode state is created)."""
    from hutoken_b200 import capi
    eng = capi.Engine(v, s, **kw)
    os.environ["HUTOKEN_B200_DECODE"] = "general"
    try:
        eng.batch_decode_bytes([[]])
    finally:
        del os.environ["HUTOKEN_B200_DECODE"]
    return eng


@pytest.mark.gpu
@pytest.mark.parametrize("name", MODELS)
def test_gpu_decode_token_wise_equals_general(name):
    """The two sets of decode kernels on the same ids: documents of every size from empty to thousands of sub-blocks,
    offsets that put document starts on, before and after sub-block boundaries, prefix stripping on every document."""
    from common import engine_for
    eng, orc = engine_for(name), oracle_for(name)
    v, s = synth.PATHS[name]
    gen = _general_engine(v, s, **synth.MODEL_KW[name])
    data, offs = synth.corpus(67, 1 << 20, special_every=150)
    ids, _ = eng.encode_packed(data, offs)
    ids = ids[ids >= 0]
    rng = np.random.default_rng(5)
    for shape in ("one", "tiny", "ragged", "aligned"):
        n = len(ids)
        if shape == "one":
            cuts = np.array([], dtype=np.int64)
        elif shape == "tiny":
            n = 40000
            cuts = np.sort(rng.integers(0, n + 1, size=12000))  # many empty documents, most sub-blocks start several
        elif shape == "ragged":
            cuts = np.sort(rng.integers(0, n + 1, size=300))
        else:
            cuts = np.arange(256, n, 256 * 7)
        tok = np.concatenate([[0], cuts, [n]]).astype(np.uint64)
        a_text, a_off = eng.decode_packed(ids[:n], tok)
        assert eng.decode_stats()["path"] == "token-wise", (name, shape)
        b_text, b_off = gen.decode_packed(ids[:n], tok)
        assert gen.decode_stats()["path"] == "general"
        assert np.array_equal(a_off, b_off), (name, shape)
        assert a_text[:int(a_off[-1])].tobytes() == b_text[:int(b_off[-1])].tobytes(), (name, shape)
        for d in rng.integers(0, len(tok) - 1, size=20):
            got = a_text[int(a_off[d]):int(a_off[d + 1])].tobytes()
            assert got == orc.decode_bytes(ids[int(tok[d]):int(tok[d + 1])]), (name, shape, d)
    gen.close()


@pytest.mark.gpu
def test_gpu_decode_tokens_that_end_inside_a_pattern():
    """Llama-style table (`<0xHH>` patterns): a token ending in '<' decodes alone unless the following tokens spell a
    pattern with it -- inside one document only.  Path and bytes against the oracle."""
    from hutoken_b200 import capi
    tmp = tempfile.mkdtemp()
    toks = [b"a", b"<", b"0", b"x", b"A", b">", b"<0", b"\xE2\x96\x81", b"b<", b"x0A>", b"\xE2\x96\x81<0x0", b"A>b",
            b"<0x0A>", b"<0x", b"\xE2\x96"]
    v = os.path.join(tmp, "pat.vocab.txt")
    synth.write_vocab(v, [(t, i) for i, t in enumerate(toks)])
    s = synth.PATHS["llama_small"][1]
    kw = dict(prefix="\u2581", is_byte_encoder=False)
    eng, orc = capi.Engine(v, s, **kw), O.Oracle(v, s, **kw)
    T = {t: i for i, t in enumerate(toks)}
    cases = [
        ([[T[b"b<"], T[b"a"]]], "token-wise"),                                   # the walk goes on and finds nothing
        ([[T[b"<"], T[b"0"], T[b"x"], T[b"0"], T[b"A"], T[b">"]]], "general"),  # spelled across tokens
        ([[T[b"a"], T[b"<0"]], [T[b"x0A>"], T[b"a"]]], "token-wise"),            # ... across two documents: no pattern
        ([[T[b"a"], T[b"<0"], T[b"x0A>"]]], "general"),
        ([[T[b"\xE2\x96\x81<0x0"], T[b"A>b"]]], "general"),                      # prefix + pattern over the boundary
        ([[T[b"\xE2\x96\x81<0x0"], T[b"a"]], [T[b"<0x0A>"], T[b"<0x"]]], "token-wise"),
        ([[T[b"<0x"], T[b"<0x"], T[b"0"], T[b"a"]] * 200], "token-wise"),
        ([[T[b"a"], T[b"\xE2\x96"]]], "general"),                                # ends inside a character
    ]
    for lists, path in cases:
        got = eng.batch_decode_bytes(lists)
        assert eng.decode_stats()["path"].split(",")[0] == path, (lists, path)
        for x, g in zip(lists, got):
            assert g == orc.decode_bytes(x), (x, g)
    eng.close()
    orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", MODELS)
def test_gpu_decode_one_launch_path(name):
    """hut_decode_batch of a few ids (hutoken.decode of a sentence): one kernel on pinned host buffers.  Sizes on both
    sides of its limits (2048 ids, 256 lists, 32 KiB of text) against the oracle and against the large-batch kernels."""
    from common import engine_for
    eng, orc = engine_for(name), oracle_for(name)
    data, offs = synth.corpus(71, 256 << 10, special_every=90)
    ids, _ = eng.encode_packed(data, offs)
    ids = ids[ids >= 0]
    rng = np.random.default_rng(11)
    seen = set()
    for n, nd in ((1, 1), (7, 1), (20, 3), (255, 40), (2047, 256), (2048, 1), (2048, 256), (2049, 1), (900, 257), (2048, 2)):
        start = int(rng.integers(0, len(ids) - n))
        x = ids[start:start + n]
        cuts = np.sort(rng.integers(0, n + 1, size=nd - 1))
        tok = np.concatenate([[0], cuts, [n]]).astype(np.uint64)
        text, toff = eng.decode_packed(x, tok)
        path = eng.decode_stats()["path"]
        seen.add(path)
        assert (path == "token-wise, one launch") == (n <= 2048 and nd <= 256), (name, n, nd, path)
        for d in range(nd):
            assert text[int(toff[d]):int(toff[d + 1])].tobytes() == orc.decode_bytes(x[int(tok[d]):int(tok[d + 1])]), (name, n, nd, d)
    # more text than the kernel stages: the longest token 2048 times
    lens = [len(orc.decode_bytes([i])) if _has_id(orc, i) else 0 for i in range(min(eng.vocab_lines, 3000))]
    big = int(np.argmax(lens))
    if lens[big] * 2048 > 32768:
        x = np.full(2048, big, dtype=np.int32)
        text, toff = eng.decode_packed(x, np.array([0, 2048], dtype=np.uint64))
        assert eng.decode_stats()["path"] in ("token-wise", "general")
        assert text.tobytes() == orc.decode_bytes(x)
    assert "token-wise, one launch" in seen
    with pytest.raises(ValueError, match="less than vocab size"):
        eng.batch_decode_bytes([[0, eng.vocab_lines]])
    with pytest.raises(ValueError, match="less than vocab size"):
        eng.batch_decode_bytes([[-1]])


@pytest.mark.gpu
def test_gpu_decode_python_surface():
    """hutoken.decode / batch_decode with the reference's messages (lib.c:876-1085, hutoken.py:141-164)."""
    from hutoken_b200 import _hutoken as ext
    from hutoken_b200 import hutoken
    name = "gpt2_small"
    v, s = synth.PATHS[name]
    ext.initialize(v, s, None, True)
    texts = ["hello world", " árvíztűrő tükörfúrógép 123", "x\n\ny", ""]
    enc = ext.batch_encode(texts, 2)
    assert [ext.decode(e) for e in enc] == texts
    assert ext.batch_decode(enc[:3], 4) == texts[:3]
    assert hutoken.decode(enc[0]) == texts[0] and hutoken.batch_decode(enc[:2]) == texts[:2]
    with pytest.raises(ValueError, match="less than vocab size"):
        ext.decode([10 ** 6])
    with pytest.raises(ValueError, match="hutoken: Error decoding tokens"):
        hutoken.decode([-1])
    with pytest.raises(TypeError, match="Argument must be a list of integers"):
        ext.decode("abc")
    with pytest.raises(ValueError, match="No tokens provided."):
        ext.batch_decode([])
    with pytest.raises(TypeError, match="Each item must be a list of integers."):
        ext.batch_decode([[1], 2])


# ---------------------------------------------------------------------------------------------------------------
# every byte as a token: ragged, over-long and invalid UTF-8 through both decode paths
# ---------------------------------------------------------------------------------------------------------------
def _all_bytes_model(tmp):
    """Vocab of the single bytes 0x01..0xFF (id = byte - 1) + GPT-2's special-chars table."""
    v = os.path.join(tmp, "bytes.vocab.txt")
    with open(v, "w") as f:
        for b in range(1, 256):
            f.write(f"0x{b:02X} == {b - 1}\n")
    return v, synth.PATHS["gpt2_small"][1]


def _byte_cases(rng):
    """Byte strings (no 0x00): well-formed text (whole tiles take the one-pass path), the same with anomalies dropped
    in (over-long forms E0 80 xx / F0 80 80 xx / C0 xx, invalid leads F8.., stray continuation bytes, truncated
    characters, replacement strings cut short), and pure noise; short ones and ones spanning several 1 KiB tiles."""
    good = "árvíztűrő tükörfúrógép hello, world! 123 ĠĊ ▁x € 😀 ".encode("utf-8")
    odd = [b"\xE0\x80\x81", b"\xE0\x81\xBF", b"\xF0\x80\x80\x81", b"\xC0\x81", b"\xC1\xBF", b"\xF8", b"\xFF", b"\x80",
           b"\xBF\xBF", b"\xC3", b"\xE2\x96", b"\xF0\x9F\x98", b"\xC4", b"\xC4\xA0\xC4", b"\xED\xA0\x80", b"\xF4\x90\x80\x80"]
    out = [b"", b"a", b"\xC3", b"\xE0\x80\x81", good]
    for n in (5, 40, 300, 1020, 1024, 1030, 2500, 9000):
        t = (good * (n // len(good) + 1))[:n]
        out.append(t)
        for _ in range(3):
            cut = sorted(rng.randrange(len(t) + 1) for _ in range(rng.randint(1, 6)))
            parts, last = [], 0
            for c in cut:
                parts += [t[last:c], rng.choice(odd)]
                last = c
            out.append(b"".join(parts + [t[last:]]))
        out.append(bytes(rng.randrange(1, 256) for _ in range(n)))
    return out


@needs_ref
@pytest.mark.parametrize("byte_mode", [True, False])
def test_oracle_decode_matches_reference_on_all_bytes_vocab(byte_mode):
    ref = O.load_ref()
    v, s = _all_bytes_model(tempfile.mkdtemp())
    ref.initialize(v, s, None, byte_mode)
    orc = O.Oracle(v, s, prefix=None, is_byte_encoder=byte_mode)
    for raw in _byte_cases(random.Random(77)):
        ids = [b - 1 for b in raw]
        got = orc.decode_bytes(ids).split(b"\0", 1)[0]
        try:
            want = ref.decode(ids).encode("utf-8")
        except UnicodeDecodeError:
            with pytest.raises(UnicodeDecodeError):
                got.decode("utf-8")
            continue
        assert got == want, (byte_mode, raw[:40])
    orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("byte_mode", [True, False])
def test_gpu_decode_all_bytes_vocab(byte_mode):
    from hutoken_b200 import capi
    v, s = _all_bytes_model(tempfile.mkdtemp())
    eng = capi.Engine(v, s, prefix=None, is_byte_encoder=byte_mode)
    orc = O.Oracle(v, s, prefix=None, is_byte_encoder=byte_mode)
    cases = _byte_cases(random.Random(77))
    lists = [[b - 1 for b in raw] for raw in cases]
    got = eng.batch_decode_bytes(lists)
    for raw, x, g in zip(cases, lists, got):
        assert g == orc.decode_bytes(x), (byte_mode, raw[:40])
    eng.close()
    orc.close()
