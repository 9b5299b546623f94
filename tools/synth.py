"""Synthetic corpora and vocabularies for tests and bench.py.

Data tooling only -- no tokenizer logic lives here.  Corpus bytes come from tools/hut_synth.c (seeded,
deterministic); vocab files are written in hutoken's on-disk formats (README.md:260-319 of the reference):

    vocab          (0xHH)+ == <id>\\n
    special chars  <decimal byte> == <replacement>\\n
"""
from __future__ import annotations

import collections
import ctypes as C
import heapq
import os
import re
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libhut_synth.so")
DATA_DIR = os.path.join(os.path.dirname(_HERE), "data")

LEXICON_SEED = 1234
LEXICON_TYPES = 200_000
ZIPF_S = 1.1


def _lib():
    src = os.path.join(_HERE, "hut_synth.c")
    if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", src, "-o", _LIB, "-lm"], check=True)
    lib = C.CDLL(_LIB)
    lib.hut_synth_lexicon.restype = C.c_uint64
    lib.hut_synth_lexicon.argtypes = [C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.hut_synth_corpus.restype = C.c_uint64
    lib.hut_synth_corpus.argtypes = [C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_double, C.c_uint64,
                                     C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
    return lib


_lex_cache = {}


def lexicon(seed=LEXICON_SEED, n_types=LEXICON_TYPES):
    """(bytes uint8 array, offsets uint32 array[n+1]); word i is the rank-i word."""
    key = (seed, n_types)
    if key not in _lex_cache:
        buf = np.zeros(n_types * 48, dtype=np.uint8)
        offs = np.zeros(n_types + 1, dtype=np.uint32)
        used = _lib().hut_synth_lexicon(seed, n_types, buf.ctypes.data, buf.size, offs.ctypes.data)
        assert used > 0
        _lex_cache[key] = (buf[:used].copy(), offs)
    return _lex_cache[key]


def corpus(seed, target_bytes, special_every=0, n_types=LEXICON_TYPES, zipf_s=ZIPF_S, out=None):
    """Seeded documents.  Returns (data uint8[nbytes], doc_offsets uint64[ndocs+1])."""
    lex, lex_offs = lexicon(n_types=n_types)
    cap = int(target_bytes) + 70000
    if out is None:
        out = np.empty(cap, dtype=np.uint8)
    assert out.size >= cap
    max_docs = int(target_bytes) // 64 + 16
    offs = np.zeros(max_docs + 1, dtype=np.uint64)
    nbytes = C.c_uint64()
    nd = _lib().hut_synth_corpus(seed, lex.ctypes.data, lex_offs.ctypes.data, n_types, zipf_s, int(target_bytes),
                                 int(special_every), out.ctypes.data, offs.ctypes.data, max_docs, C.byref(nbytes))
    return out[:nbytes.value], offs[:nd + 1].copy()


def docs_of(data, offs):
    b = data.tobytes()
    return [b[int(offs[i]):int(offs[i + 1])] for i in range(len(offs) - 1)]


# ------------------------------------------------------------------------------------------------
# GPT-2 byte <-> unicode table (public algorithm: printable bytes map to themselves, the other 68 to
# U+0100 + n in byte order).  hutoken carries those 68 in the special-chars file (hutoken.py:15-20).
# ------------------------------------------------------------------------------------------------
def gpt2_byte_encoder():
    keep = list(range(33, 127)) + list(range(161, 173)) + list(range(174, 256))
    table, n = {}, 0
    for b in keep:
        table[b] = chr(b)
    for b in range(256):
        if b not in table:
            table[b] = chr(256 + n)
            n += 1
    order = keep + [b for b in range(256) if b not in keep]
    return table, order


GPT2_REMAPPED = list(range(0, 33)) + list(range(127, 161)) + [173]


def write_special_gpt2(path):
    table, _ = gpt2_byte_encoder()
    with open(path, "w", encoding="utf-8") as f:
        for b in GPT2_REMAPPED:
            f.write(f"{b} == {table[b]}\n")


def write_special_llama(path):
    with open(path, "w", encoding="utf-8") as f:
        f.write("32 == ▁\n")
        for b in list(range(1, 32)) + [127]:
            f.write(f"{b} == <0x{b:02X}>\n")


def write_vocab(path, entries):
    """entries: iterable of (bytes, id) in file order."""
    with open(path, "w", encoding="ascii") as f:
        for key, idx in entries:
            f.write("".join(f"0x{b:02X}" for b in key) + f" == {idx}\n")


# ------------------------------------------------------------------------------------------------
# Offline BPE trainer over unique pretokens (word-count based, incremental pair counts).
# ------------------------------------------------------------------------------------------------
_WORD_RE = re.compile(r" ?[A-Za-záéíóúőűüöÁÉÍÓÚŐÜŰÖ]+| ?[0-9]+| ?[^\sA-Za-záéíóúőűüöÁÉÍÓÚŐÜŰÖ0-9]+| +|\s")


def count_pretokens(text: str):
    return collections.Counter(_WORD_RE.findall(text))


def train_bpe(word_counts, n_merges, min_count=2):
    """word_counts: {tuple(symbols): count}.  Returns the list of merged symbol strings, in merge order."""
    words = [list(w) for w in word_counts]
    counts = list(word_counts.values())
    pair_count = collections.Counter()
    pair_words = collections.defaultdict(set)
    for wi, w in enumerate(words):
        c = counts[wi]
        for a, b in zip(w, w[1:]):
            pair_count[(a, b)] += c
            pair_words[(a, b)].add(wi)
    heap = [(-c, p) for p, c in pair_count.items()]
    heapq.heapify(heap)
    merges, seen = [], set()
    while len(merges) < n_merges and heap:
        negc, pair = heapq.heappop(heap)
        cur = pair_count.get(pair, 0)
        if cur != -negc:
            if cur > 0:
                heapq.heappush(heap, (-cur, pair))
            continue
        if cur < min_count:
            break
        a, b = pair
        merged = a + b
        touched = set()
        for wi in list(pair_words[pair]):
            w = words[wi]
            c = counts[wi]
            for x, y in zip(w, w[1:]):
                pair_count[(x, y)] -= c
                touched.add((x, y))
            nw, i = [], 0
            while i < len(w):
                if i + 1 < len(w) and w[i] == a and w[i + 1] == b:
                    nw.append(merged)
                    i += 2
                else:
                    nw.append(w[i])
                    i += 1
            words[wi] = nw
            for x, y in zip(nw, nw[1:]):
                pair_count[(x, y)] += c
                pair_words[(x, y)].add(wi)
                touched.add((x, y))
        pair_words.pop(pair, None)
        pair_count.pop(pair, None)
        for p in touched:
            c = pair_count.get(p, 0)
            if c > 0:
                heapq.heappush(heap, (-c, p))
        if merged not in seen:
            seen.add(merged)
            merges.append(merged)
    return merges


def _sample_text(sample_mb, seed):
    data, _ = corpus(seed, sample_mb << 20)
    return data.tobytes().decode("utf-8")


def build_gpt2_vocab(vocab_path, special_path, n_merges=50_000, sample_mb=32, seed=2):
    """GPT-2-shaped byte-level vocab: 256 byte symbols (GPT-2 order) + n_merges merges, ids dense."""
    table, order = gpt2_byte_encoder()
    wc = count_pretokens(_sample_text(sample_mb, seed))
    enc_counts = collections.Counter()
    for w, c in wc.items():
        enc_counts[tuple(table[b] for b in w.encode("utf-8"))] += c
    merges = train_bpe(enc_counts, n_merges)
    entries = [(table[b].encode("utf-8"), i) for i, b in enumerate(order)]
    entries += [(m.encode("utf-8"), 256 + i) for i, m in enumerate(merges)]
    write_vocab(vocab_path, entries)
    write_special_gpt2(special_path)
    return len(entries)


def build_llama_vocab(vocab_path, special_path, vocab_size=32_000, sample_mb=32, seed=3):
    """Llama / PULI-LlumiX-shaped vocab: <unk> <s> </s>, 256 <0xHH> byte tokens, merges, then single chars
    (sentencepiece order); words carry the U+2581 prefix; plain (non byte-encoder) mode."""
    wc = count_pretokens(_sample_text(sample_mb, seed))
    enc_counts = collections.Counter()
    chars = collections.Counter()
    for w, c in wc.items():
        if w.isspace() and w != " ":
            continue  # control whitespace goes through <0xHH> replacements
        e = w.replace(" ", "▁")
        if not e.startswith("▁"):
            # a word at the start of a document gets the prefix; train both shapes, mostly prefixed
            enc_counts[tuple("▁" + e)] += c
        else:
            enc_counts[tuple(e)] += c
        for ch in e:
            chars[ch] += c
    chars["▁"] += 1
    specials = ["<unk>", "<s>", "</s>"] + [f"<0x{b:02X}>" for b in range(256)]
    n_merges = vocab_size - len(specials) - len(chars)
    merges = train_bpe(enc_counts, n_merges)
    toks = specials + merges + [ch for ch, _ in chars.most_common()]
    write_vocab(vocab_path, [(t.encode("utf-8"), i) for i, t in enumerate(toks)])
    write_special_llama(special_path)
    return len(toks)


PATHS = {
    "gpt2": (os.path.join(DATA_DIR, "gpt2_50k.vocab.txt"), os.path.join(DATA_DIR, "gpt2_special.txt")),
    "llama": (os.path.join(DATA_DIR, "llama_32k.vocab.txt"), os.path.join(DATA_DIR, "llama_special.txt")),
    "gpt2_small": (os.path.join(DATA_DIR, "gpt2_2k.vocab.txt"), os.path.join(DATA_DIR, "gpt2_special.txt")),
    "llama_small": (os.path.join(DATA_DIR, "llama_2k.vocab.txt"), os.path.join(DATA_DIR, "llama_special.txt")),
}

# merges files (lib.c:573-663) derived from the small vocabularies by tools/make_merges.py --noise
MERGES = {
    "gpt2_small": os.path.join(DATA_DIR, "gpt2_2k.merges.txt"),
    "llama_small": os.path.join(DATA_DIR, "llama_2k.merges.txt"),
}

MODEL_KW = {
    "gpt2": dict(prefix=None, is_byte_encoder=True),
    "gpt2_small": dict(prefix=None, is_byte_encoder=True),
    "llama": dict(prefix="▁", is_byte_encoder=False),
    "llama_small": dict(prefix="▁", is_byte_encoder=False),
}

if __name__ == "__main__":
    import sys
    import time
    os.makedirs(DATA_DIR, exist_ok=True)
    which = sys.argv[1:] or ["gpt2_small", "llama_small", "gpt2", "llama"]
    for name in which:
        t0 = time.time()
        v, s = PATHS[name]
        if name == "gpt2":
            n = build_gpt2_vocab(v, s)
        elif name == "gpt2_small":
            n = build_gpt2_vocab(v, s, n_merges=2000, sample_mb=4)
        elif name == "llama":
            n = build_llama_vocab(v, s)
        else:
            n = build_llama_vocab(v, s, vocab_size=2300, sample_mb=4)
        print(f"{name}: {n} entries -> {v} ({time.time() - t0:.1f}s)")
