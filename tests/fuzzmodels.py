"""Random adversarial tiny models + texts for parity fuzzing (shared by CPU and GPU tests).

The vocabularies made here are deliberately NOT BPE-trained: ids are assigned at random, so the tables
contain non-monotone ranks (a longer token with a smaller id than its parts), tokens whose parts are not in
the vocabulary (virtual units), duplicate keys (last line wins, hashmap.c:208-214), unknown
symbols (-1, core.c:207) and ambiguous splits -- the cases where a GPU pair table could diverge from the
reference's string-keyed ranks (core.c:700-722).
"""
from __future__ import annotations

import os
import random

import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tools"))
import synth  # noqa: E402

ALPHA_PLAIN = ["a", "b", "c", "d", "e", "é", "ő", "A", "Z", "ű"]
OTHERS = ["!", "?", ".", "-", "€", "ñ", "😂", "<", ">", "\"", "日"]
DIGITS = ["0", "1", "2", "7", "x"]
SPACES = [" ", " ", " ", "  ", "\n", "\t", " \n", "   ", "\r", "\x0b", "\x0c"]


def random_text(rng: random.Random, n_units: int) -> str:
    out = []
    for _ in range(n_units):
        k = rng.random()
        if k < 0.55:
            out.append("".join(rng.choice(ALPHA_PLAIN) for _ in range(rng.randint(1, 9))))
        elif k < 0.65:
            out.append("".join(rng.choice(DIGITS[:4]) for _ in range(rng.randint(1, 5))))
        elif k < 0.8:
            out.append("".join(rng.choice(OTHERS) for _ in range(rng.randint(1, 3))))
        else:
            out.append(rng.choice(SPACES))
        if rng.random() < 0.7:
            out.append(" ")
    return "".join(out)


def _encode_piece_bytes(piece: bytes, byte_mode: bool) -> bytes:
    """Encoded form of raw text bytes under the GPT-2 table (byte mode) or identity with space -> U+2581."""
    if byte_mode:
        table, _ = synth.gpt2_byte_encoder()
        return "".join(table[b] for b in piece).encode("utf-8")
    return piece.replace(b" ", "▁".encode("utf-8"))


def random_model(rng: random.Random, tmpdir: str, byte_mode: bool, name: str):
    """Writes vocab + special files; returns (vocab_path, special_path, kwargs)."""
    special = os.path.join(tmpdir, f"{name}_special.txt")
    vocab = os.path.join(tmpdir, f"{name}_vocab.txt")
    if byte_mode:
        synth.write_special_gpt2(special)
        prefix = None
    else:
        synth.write_special_llama(special)
        prefix = "▁" if rng.random() < 0.8 else None
    # candidate strings: substrings of random words (in encoded form), 1..6 symbols
    keys = set()
    table, order = synth.gpt2_byte_encoder()
    alphabet = ALPHA_PLAIN + OTHERS[:6] + DIGITS[:4] + [" "]
    if byte_mode:
        # most single byte symbols present, a few missing (-> -1)
        for b in order:
            if rng.random() < 0.93:
                keys.add(table[b].encode("utf-8"))
    else:
        for ch in alphabet + ["▁"]:
            if rng.random() < 0.8:
                keys.add(_encode_piece_bytes(ch.encode("utf-8"), False))
        for b in (1, 9, 10, 13, 127):
            if rng.random() < 0.7:
                keys.add(f"<0x{b:02X}>".encode())
    n_multi = rng.randint(20, 400)
    for _ in range(n_multi):
        lead = " " if rng.random() < 0.4 else ""
        cls = rng.random()
        src = ALPHA_PLAIN if cls < 0.75 else (DIGITS[:4] if cls < 0.85 else OTHERS[:6])
        w = lead + "".join(rng.choice(src) for _ in range(rng.randint(1, 6)))
        raw = w.encode("utf-8")
        if not byte_mode and prefix and rng.random() < 0.3 and not w.startswith(" "):
            enc = prefix.encode("utf-8") + _encode_piece_bytes(raw, False)
        else:
            enc = _encode_piece_bytes(raw, byte_mode)
        # in byte mode also add keys cut in the middle of a multi-byte char (byte-level pieces)
        if byte_mode and rng.random() < 0.3 and len(raw) > 1:
            cut = rng.randint(1, len(raw))
            enc = _encode_piece_bytes(raw[:cut], True)
        keys.add(enc)
    keys = [k for k in keys if k and b"\0" not in k]
    rng.shuffle(keys)
    ids = list(range(len(keys)))
    if rng.random() < 0.7:
        rng.shuffle(ids)  # non-monotone ranks
    entries = list(zip(keys, ids))
    n_lines = len(entries)
    # duplicate keys: re-state a few keys with the id of another line swapped (ids stay < #lines)
    for _ in range(rng.randint(0, 3)):
        i, j = rng.randrange(n_lines), rng.randrange(n_lines)
        entries.append((entries[i][0], entries[j][1]))
        entries.append((entries[j][0], entries[i][1]))
    # occasionally two different strings share one id (equal ranks -> leftmost pair wins, queue.c:162-164)
    if rng.random() < 0.3:
        for _ in range(rng.randint(1, 4)):
            i, j = rng.randrange(n_lines), rng.randrange(n_lines)
            entries.append((entries[i][0], entries[j][1]))
    synth.write_vocab(vocab, entries)
    return vocab, special, dict(prefix=prefix, is_byte_encoder=byte_mode)


def random_merges(rng: random.Random, vocab_path: str, out_path: str) -> str:
    """A merges file (lib.c:573-663) over an existing vocab: rules in random order (ranks unrelated to ids), unknown
    tokens, comments, repeated pairs (the later rank wins), rules sharing a merged token."""
    import make_merges
    vocab = make_merges.read_vocab(vocab_path)
    keys = [k for k in vocab if not any(b in k for b in b" \r\n") and not k.startswith(b"#")]
    keyset = set(keys)
    rules = []
    for tok in keys:
        cuts = make_merges._chars(tok)
        for c in cuts[:-1]:
            if tok[:c] in keyset and tok[c:] in keyset and rng.random() < 0.8:
                rules.append((tok[:c], tok[c:]))
    rng.shuffle(rules)
    lines = []
    for l, r in rules:
        lines.append(l + b" " + r)
        k = rng.random()
        if k < 0.03:
            lines.append(b"# comment")
        elif k < 0.06 and keys:
            lines.append(rng.choice(keys) + b" nosuchtoken")
        elif k < 0.10:
            pl, pr = rng.choice(rules)
            lines.append(pl + b"  " + pr)
    with open(out_path, "wb") as f:
        f.write(b"\n".join(lines) + (b"\n" if lines else b""))
    return out_path
