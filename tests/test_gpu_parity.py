"""GPU parity: the CUDA engine (through the C-ABI) against the CPU oracle on the same inputs.
Bit-exact: ids are integers, so every comparison is equality (no tolerance)."""
import random
import tempfile

import numpy as np
import pytest

import fuzzmodels as F
import synth
from common import MODELS, assert_same, engine_for, oracle_for, pack

pytestmark = pytest.mark.gpu

BASIC = [
    "hello world", "", " ", "  ", "a", " a", "a ", "a  b", "a   b    c", "aaa", "aaaa", " Hello", "Hello, World!",
    "12345 abc!!", "árvíztűrő tükörfúrógép", "ÁRVÍZTŰRŐ TÜKÖRFÚRÓGÉP", "x\n\ny\t z", "señor ñ!", "€ 😂 日本",
    "tab\there", "line1\nline2\r\nline3", "  leading", "trailing  ", "a \n b", "1a.b2 c", "don't stop", "<0x41>",
    "<|endoftext|>", "\x01\x02\x7f", "mixed123abc456", "...!!!???", " . , ; ", "ő", "őő űű", "a" * 40, "ab" * 100,
    "é" * 50, " " * 70, "\n" * 33, "word " * 300, "How can the net amount of entropy of the universe be massively decreased?",
]


@pytest.mark.parametrize("name", MODELS)
def test_basic_strings(name):
    eng, orc = engine_for(name), oracle_for(name)
    got = eng.batch_encode(BASIC)
    want = orc.batch_encode(BASIC)
    for t, g, w in zip(BASIC, got, want):
        assert g == w, (name, t[:40], g[:20], w[:20])
    # one document at a time too (documents starting at offset 0, the prefix path)
    for t in BASIC[:20]:
        assert eng.encode(t) == orc.encode(t), (name, t)


@pytest.mark.parametrize("name", MODELS)
@pytest.mark.parametrize("seed,special_every", [(21, 0), (22, 200)])
def test_synthetic_corpus(name, seed, special_every):
    eng, orc = engine_for(name), oracle_for(name)
    data, offs = synth.corpus(seed, 3 << 20, special_every=special_every)
    ids, tok = eng.encode_packed(data, offs)
    oids, otok = orc.encode_packed(data, offs)
    assert_same(ids, tok, oids, otok, f"{name} seed {seed}")
    st = eng.stats()
    assert st["kernel_launches"] >= 2 and st["pretokens"] > 0


def test_cache_is_transparent():
    """Same answer with a cold cache, a warm cache and a tiny (thrashing) cache."""
    name = "gpt2"
    data, offs = synth.corpus(31, 2 << 20)
    oids, otok = oracle_for(name).encode_packed(data, offs)
    eng = engine_for(name)
    eng.cache_reset()
    a = eng.encode_packed(data, offs)
    b = eng.encode_packed(data, offs)  # warm
    assert eng.stats()["cache_hits"] > 0
    tiny = engine_for(name, cache_log2_slots=10)
    c = tiny.encode_packed(data, offs)
    for got in (a, b, c):
        assert_same(got[0], got[1], oids, otok, "cache variants")


@pytest.mark.parametrize("byte_mode", [True, False])
def test_fuzz_models(byte_mode):
    """Adversarial random vocabularies: non-monotone ranks, virtual units, duplicate keys / ids, -1 symbols."""
    from hutoken_b200 import capi
    from oracle import oracle as O
    tmp = tempfile.mkdtemp()
    for seed in range(40):
        rng = random.Random(1000 + seed)
        v, s, kw = F.random_model(rng, tmp, byte_mode, f"g{seed}")
        eng = capi.Engine(v, s, **kw)
        orc = O.Oracle(v, s, **kw)
        texts = [F.random_text(rng, rng.randint(0, 60)) for _ in range(80)]
        if byte_mode:  # raw invalid UTF-8 is defined in byte mode only
            texts += [bytes(rng.choice([0x80, 0xC3, 0xE2, 0xF0, 0x41, 0x20, 0xA9, 0xFF, 0xC0, 0x0A, 0x61])
                            for _ in range(rng.randint(1, 40))) for _ in range(20)]
        got, want = eng.batch_encode(texts), orc.batch_encode(texts)
        for t, g, w in zip(texts, got, want):
            assert g == w, (seed, kw, t, g, w)
        eng.close()


@pytest.mark.parametrize("name", ["gpt2_small", "llama_small"])
def test_document_edges(name):
    """Empty / one-byte / many tiny documents, documents ending inside a tile halo, tile-straddling words."""
    eng, orc = engine_for(name), oracle_for(name)
    rng = random.Random(5)
    docs = ["", "", "a", "", " ", "é", "", "x" * 4095, "y" * 4096, "z" * 4097, "", ("w" * 17 + " ") * 500, ""]
    docs += [F.random_text(rng, rng.randint(0, 3)) for _ in range(3000)]
    docs += ["", ""]
    data, offs = pack(docs)
    ids, tok = eng.encode_packed(data, offs)
    oids, otok = orc.encode_packed(data, offs)
    assert_same(ids, tok, oids, otok, name)
    # only empty documents / no documents
    for d in ([], [""], ["", "", ""]):
        data, offs = pack(d)
        ids, tok = eng.encode_packed(data, offs)
        assert len(ids) == 0 and tok.tolist() == [0] * (len(d) + 1)


@pytest.mark.parametrize("name", ["gpt2_small", "llama_small"])
def test_medium_and_long_pretokens(name):
    """Pretokens beyond 32 symbols (block-cooperative path) and beyond a tile (big path), reference-compatible."""
    eng, orc = engine_for(name), oracle_for(name)
    rng = random.Random(9)
    letters = "abcdeégőz"
    docs = []
    for n in (33, 64, 100, 129, 500, 4000, 4300, 9000, 16384, 65536):
        docs.append("start " + "".join(rng.choice(letters) for _ in range(n)) + " end")
        docs.append("".join(rng.choice("0123456789") for _ in range(n)))
        docs.append(" " * n + "x")
        docs.append("!?" * (n // 2))
    data, offs = pack(docs)
    ids, tok = eng.encode_packed(data, offs)
    oids, otok = orc.encode_packed(data, offs)
    assert_same(ids, tok, oids, otok, name)
    st = eng.stats()
    assert st["block_merges"] > 0 and st["big_merges"] > 0


def test_reference_cut_at_overlong_pretoken():
    """core.c:402-407,503: a pretoken > 262144 bytes ends its document silently; later documents are unaffected."""
    name = "gpt2_small"
    eng, orc = engine_for(name), oracle_for(name)
    docs = ["before one", "keep this " + "a" * (262144 + 8) + " dropped tail", "after one",
            "b" * 262145, "x " + "c" * 262144 + " kept",
            "configs[4]: a 16 MiB whitespace-free run " + "abcdefgh" * (2 << 20) + " tail", "last one"]
    data, offs = pack(docs)
    ids, tok = eng.encode_packed(data, offs)
    oids, otok = orc.encode_packed(data, offs)
    assert_same(ids, tok, oids, otok, name)
    assert eng.stats()["cut_documents"] == 3


@pytest.mark.parametrize("name", ["gpt2", "llama"])
def test_no_limit_long_runs(name):
    """configs[4] as north_star words it: single-class runs of 256 KiB + 8 ... 16 MiB really MERGED on the GPU (opt-in
    no-limit mode, hut_options.max_pretoken) and compared with the CPU restatement in its no-limit mode -- the
    reference itself cuts such a document (core.c:402-407,503).  Merge order: core.c:111-176, queue.c:152-199."""
    eng = engine_for(name, max_pretoken=32 << 20)
    orc = oracle_for(name, no_limit=True)
    rng = random.Random(77)
    docs = ["before"]
    for n in (262152, 1 << 20, 4 << 20, 16 << 20):
        docs.append("go " + "".join(rng.choices("abcdefghijklmnopqrstuvwxyz", k=n)) + " end")
    docs += ["0123456789" * 30000, "between", "é" * 140000 + " x", "after"]
    data, offs = pack(docs)
    ids, tok = eng.encode_packed(data, offs)
    oids, otok = orc.encode_packed(data, offs)
    assert_same(ids, tok, oids, otok, f"no-limit {name}")
    st = eng.stats()
    assert st["big_merges"] >= 4 and st["cut_documents"] == 0
    # a pretoken beyond the configured working area is an error, never a silent cut
    small = engine_for(name, max_pretoken=300000)
    from hutoken_b200 import capi
    with pytest.raises(capi.HutError, match="max_pretoken"):
        small.encode_packed(*pack(["x" * 300001]))
    ids, tok = small.encode_packed(*pack(["y" * 300000 + " z"]))
    oids, otok = orc.encode_packed(*pack(["y" * 300000 + " z"]))
    assert_same(ids, tok, oids, otok, f"no-limit at the limit {name}")


@pytest.mark.parametrize("name", ["gpt2_small", "llama_small"])
def test_no_limit_cooperative_rounds(name):
    """huge_kernel's cooperative rounds (all pairs of the minimum rank applied at once): a merges-file model has
    monotone ranks (every rule's parts are made by earlier rules), so every round of a long run takes that path.
    Runs of random letters, digits, one letter (a single chain of overlapping pairs) and a two-letter pattern, merged
    in no-limit mode and compared with the oracle (merge order: core.c:211-337, queue.c:162-164)."""
    import subprocess
    import sys
    import os
    from hutoken_b200 import capi
    from oracle import oracle as O
    from common import ROOT
    v, s = synth.PATHS[name]
    mp = os.path.join(tempfile.mkdtemp(), name + ".merges.txt")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_merges.py"), v, mp], check=True,
                   stdout=subprocess.DEVNULL)
    kw = synth.MODEL_KW[name]
    eng = capi.Engine(v, s, merges_path=mp, max_pretoken=8 << 20, **kw)
    orc = O.Oracle(v, s, merges_path=mp, no_limit=True, **kw)
    rng = random.Random(5)
    docs = ["go " + "".join(rng.choices("abcdefghijklmnopqrstuvwxyz", k=n)) + " end" for n in (5000, 300000, 2 << 20)]
    docs += ["".join(rng.choices("0123456789", k=700000)), "a" * 70000 + " " + "ab" * 40000 + " " + "aab" * 30000,
             "é" * 20000, "x"]
    data, offs = pack(docs)
    ids, tok = eng.encode_packed(data, offs)
    oids, otok = orc.encode_packed(data, offs)
    assert_same(ids, tok, oids, otok, f"cooperative rounds {name}")
    assert eng.stats()["big_merges"] >= 5
    eng.close()
    orc.close()


@pytest.mark.parametrize("byte_mode", [True, False])
def test_no_limit_rounds_on_adversarial_ranks(byte_mode):
    """huge_kernel on random vocabularies whose ranks are NOT monotone (a merge may create a pair that ranks below the
    round's rank): the rounds must notice (the check of the created pairs), apply only the occurrences left of the
    first offender and hand that one to the single-lane step.  Long runs over the models' small alphabets, string-keyed
    ranks and merges-file ranks, against the oracle's no-limit mode."""
    from hutoken_b200 import capi
    from oracle import oracle as O
    import os
    tmp = tempfile.mkdtemp()
    for seed in range(10):
        rng = random.Random(4000 + seed)
        v, s, kw = F.random_model(rng, tmp, byte_mode, f"h{seed}")
        mp = F.random_merges(rng, v, os.path.join(tmp, f"h{seed}_merges.txt")) if seed % 2 else None
        eng = capi.Engine(v, s, merges_path=mp, max_pretoken=4 << 20, **kw)
        orc = O.Oracle(v, s, merges_path=mp, no_limit=True, **kw)
        docs = []
        for n, alpha in ((6000, F.ALPHA_PLAIN[:5]), (30000, F.ALPHA_PLAIN), (9000, ["a", "b"]), (20000, F.DIGITS[:4])):
            docs.append("x " + "".join(rng.choice(alpha) for _ in range(n)) + " y")
        data, offs = pack(docs)
        ids, tok = eng.encode_packed(data, offs)
        oids, otok = orc.encode_packed(data, offs)
        assert_same(ids, tok, oids, otok, f"adversarial ranks seed {seed} byte_mode {byte_mode} merges {bool(mp)}")
        assert eng.stats()["big_merges"] >= 3
        eng.close()
        orc.close()


def test_one_context_over_distinct_devices():
    """hut_create over ALL distinct GPUs of the box (the product's own multi-GPU entry, HUTOKEN_B200_DEVICES):
    chunks dealt round-robin to physically different devices.  Skipped on a one-GPU box."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    for name in ("gpt2", "llama_small"):
        eng, orc = engine_for(name, devices=tuple(range(n)), chunk_mb=2), oracle_for(name)
        assert eng.device_count == n
        data, offs = synth.corpus(61, 24 << 20, special_every=250)
        oids, otok = orc.encode_packed(data, offs)
        for _ in range(2):
            ids, tok = eng.encode_packed(data, offs)
            assert_same(ids, tok, oids, otok, f"{name} over {n} distinct devices")
        for d in range(n):
            assert eng.stats(d)["pretokens"] > 0, f"device {d} got no work"
        ids, tok = eng.encode_packed(*pack(["only one document"]))
        assert ids.tolist() == orc.encode("only one document")


def test_device_resident_api():
    """hut_encode_device: text and ids stay on the GPU (the path bench.py times)."""
    import torch
    name = "gpt2"
    eng, orc = engine_for(name), oracle_for(name)
    data, offs = synth.corpus(41, 4 << 20)
    oids, otok = orc.encode_packed(data, offs)
    dev = torch.device("cuda:0")
    d_text = torch.from_numpy(np.concatenate([data, np.zeros(64, np.uint8)])).to(dev)
    d_offs = torch.from_numpy(offs.astype(np.int64)).to(dev)
    d_ids = torch.empty(len(data) + 64, dtype=torch.int32, device=dev)
    d_tok = torch.empty(len(offs), dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    n = eng.encode_device(0, d_text.data_ptr(), len(data), d_offs.data_ptr(), len(offs) - 1, d_ids.data_ptr(),
                          d_ids.numel(), d_tok.data_ptr())
    assert n == len(oids)
    assert_same(d_ids[:n].cpu().numpy(), d_tok.cpu().numpy().astype(np.uint64), oids, otok, "device api")
    # too small an id buffer is reported, not overrun
    from hutoken_b200 import capi
    with pytest.raises(capi.HutError):
        eng.encode_device(0, d_text.data_ptr(), len(data), d_offs.data_ptr(), len(offs) - 1, d_ids.data_ptr(), 1000,
                          d_tok.data_ptr())


def test_python_surface():
    """hutoken.initialize / encode / batch_encode drop-in (hutoken.py:22-139, lib.c:668-874)."""
    import hutoken_b200 as hutoken
    v, s = synth.PATHS["gpt2_small"]
    hutoken.initialize(v, s, is_byte_encoder=True)
    orc = oracle_for("gpt2_small")
    assert hutoken.encode("hello world") == orc.encode("hello world")
    texts = ["How can the net", " amount of entropy of", " the universe be massively decreased?", ""]
    for nt in (1, 3, 8):
        assert hutoken.batch_encode(texts, nt) == orc.batch_encode(texts)
    assert hutoken.batch_encode(texts, 0) == [[], [], [], []]  # no workers => nothing encoded (lib.c:779-794)
    assert hutoken.batch_encode(["ab\0cd"], 1) == [orc.encode("ab")]  # strdup cut (lib.c:772)
    # packed surface: same ids as the list API, as int32 / uint64 arrays
    many = texts + ["ab\0cd", "őű € 😂", "x" * 70000]
    ids, tok = hutoken.batch_encode_packed(many)
    assert ids.dtype == np.int32 and tok.dtype == np.uint64 and len(tok) == len(many) + 1
    assert [ids[tok[i]:tok[i + 1]].tolist() for i in range(len(many))] == hutoken.batch_encode(many, 1)
    data, offs = hutoken.pack_texts(many)
    ids2, tok2 = hutoken.batch_encode_packed(data, offs)
    assert np.array_equal(ids, ids2) and np.array_equal(tok, tok2)
    ids0, tok0 = hutoken.batch_encode_packed([])
    assert len(ids0) == 0 and tok0.tolist() == [0]
    with pytest.raises(RuntimeError, match="Error encoding texts"):
        hutoken.batch_encode_packed(b"ab\0cd", [0, 5])  # a NUL inside a packed document is refused, not cut
    with pytest.raises(RuntimeError, match="exceed"):
        hutoken.batch_encode_packed(b"abc", [0, 9])
    with pytest.raises(RuntimeError, match="embedded null"):
        hutoken.encode("ab\0cd")
    with pytest.raises(RuntimeError, match="Expected a list of strings"):
        hutoken.batch_encode("not a list")
    v, s = synth.PATHS["llama_small"]
    hutoken.initialize(v, s, prefix="▁")
    orc = oracle_for("llama_small")
    for t in ("hello world", " hello", "x\ny"):
        assert hutoken.encode(t) == orc.encode(t)


def test_python_surface_from_two_threads():
    """The extension's buffers and the context are process-global; calls from several Python threads are serialised
    by the module mutex (pyext.c) and must all come back right (the reference's batch_encode, lib.c:782, is
    callable from threads because it uses per-call buffers)."""
    import threading
    import hutoken_b200 as hutoken
    v, s = synth.PATHS["gpt2_small"]
    hutoken.initialize(v, s, is_byte_encoder=True)
    orc = oracle_for("gpt2_small")
    rng = random.Random(3)
    batches = [[F.random_text(rng, rng.randint(0, 400)) for _ in range(rng.randint(1, 300))] for _ in range(8)]
    batches[3] = ["x" * 300000, "tail"]  # forces the pinned packing buffer to grow under the other thread's feet
    want = [orc.batch_encode(b) for b in batches]
    errors = []

    def work(k):
        try:
            for rep in range(6):
                for i in range(k, len(batches), 2):
                    got = hutoken.batch_encode(batches[i], 2)
                    if got != want[i]:
                        errors.append((k, rep, i))
                    if hutoken.encode(batches[i][0]) != want[i][0]:
                        errors.append((k, rep, i, "encode"))
                    if hutoken.decode(want[0][0]) is None:
                        errors.append((k, rep, i, "decode"))
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    th = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors[:5]


@pytest.mark.parametrize("devices,chunk_mb", [([0], 1), ([0, 0], 1), ([0, 0, 0], 2)])
def test_host_api_streams_chunks(devices, chunk_mb):
    """hut_encode_batch streams the batch through the devices in chunks (upload | kernels | download overlap) and
    deals the chunks round-robin to the devices of the context; the result must not depend on either.  (The same
    physical GPU listed several times stands in for several GPUs: each entry has its own streams, tables and cache.)"""
    for name in ("gpt2", "llama_small"):
        eng, orc = engine_for(name, devices=tuple(devices), chunk_mb=chunk_mb), oracle_for(name)
        data, offs = synth.corpus(51, 9 << 20, special_every=300)
        oids, otok = orc.encode_packed(data, offs)
        for _ in range(2):  # second call: warm caches, reused buffers
            ids, tok = eng.encode_packed(data, offs)
            assert_same(ids, tok, oids, otok, f"{name} devices={devices}")
        assert eng.device_count == len(devices)
        # ragged batches: fewer documents than chunks / devices, empty documents, a document larger than a chunk
        docs = ["", "x" * (3 << 20) + " tail", "", "short one", ""]
        d2, o2 = pack(docs)
        ids, tok = eng.encode_packed(d2, o2)
        oids2, otok2 = orc.encode_packed(d2, o2)
        assert_same(ids, tok, oids2, otok2, f"{name} ragged devices={devices}")
        ids, tok = eng.encode_packed(*pack([]))
        assert len(ids) == 0 and tok.tolist() == [0]
