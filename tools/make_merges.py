"""Derives a merges file (the third on-disk format of hutoken, lib.c:573-663: lines ``left right``) from a vocab file.

Data tooling only.  For every vocab token of two or more UTF-8 characters, in id order, the rule is the split
(left, right) into two vocab tokens whose larger part id is minimal (parts with smaller ids than the token preferred) -- the usual way
merge lists are recovered from a rank-ordered BPE vocabulary.  Tokens containing a space, CR or LF cannot be written
in this format (the loader splits on spaces) and are skipped.

``noise=True`` adds the lines the loader must cope with (lib.c:587-640): comment lines, a line without a space, rules
naming unknown tokens, a rule whose concatenation is not a token, a duplicate (left, right) pair given again later
(the later rank wins), and runs of several spaces.

    python tools/make_merges.py <vocab.txt> <merges.txt> [--noise]
"""
from __future__ import annotations

import os
import re
import sys

_HEX = re.compile(r"0x([0-9a-fA-F]{2})")


def read_vocab(path):
    """{bytes: id}; the last line wins for a repeated key (hashmap_set, hashmap.c:207-214)."""
    vocab = {}
    with open(path, "rb") as f:
        for line in f:
            if line in (b"\n", b""):
                break
            key, _, val = line.rstrip(b"\n").partition(b" == ")
            vocab[bytes(int(h, 16) for h in _HEX.findall(key.decode("ascii")))] = int(val)
    return vocab


def _chars(tok: bytes):
    """UTF-8 character boundaries by lead byte (utf8_char_length, pretokenizer.c:14-28)."""
    cuts, i = [], 0
    while i < len(tok):
        c = tok[i]
        i += 1 if c < 0x80 else 2 if (c & 0xE0) == 0xC0 else 3 if (c & 0xF0) == 0xE0 else 4 if (c & 0xF8) == 0xF0 else 1
        cuts.append(min(i, len(tok)))
    return cuts


def derive(vocab):
    rules = []
    for tok, tid in sorted(vocab.items(), key=lambda kv: kv[1]):
        cuts = _chars(tok)
        if len(cuts) < 2 or any(b in tok for b in b" \r\n") or tok.startswith(b"#"):
            continue
        best = None
        for c in cuts[:-1]:
            l, r = tok[:c], tok[c:]
            if l in vocab and r in vocab:
                # parts with smaller ids first (a rank-ordered BPE vocabulary); otherwise (sentencepiece-style id
                # order) any split into two tokens, still preferring early parts
                cost = max(vocab[l], vocab[r]) + (0 if vocab[l] < tid and vocab[r] < tid else 1 << 30)
                if best is None or cost < best[0]:
                    best = (cost, l, r)
        if best:
            rules.append((best[1], best[2]))
    return rules


def write_merges(vocab_path, out_path, noise=False):
    vocab = read_vocab(vocab_path)
    rules = derive(vocab)
    lines = [b"#version: 0.2"] if noise else []
    for i, (l, r) in enumerate(rules):
        lines.append(l + b" " + r)
        if noise:
            if i == 3:
                lines += [b"# a comment in the middle", b"nospacehere", b"\xe2\x98\x83\xe2\x98\x83 zz",
                          b"q\xcc\x88 q\xcc\x88"]
            if i == 10:
                lines.append(l + b"   " + r + b" trailing words are ignored")
            if i % 97 == 50 and i >= 60:
                pl, pr = rules[i - 40]
                lines.append(pl + b" " + pr)  # the same pair again: replaces the earlier rule, with this later rank
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    with open(out_path, "wb") as f:
        f.write(b"\n".join(lines) + b"\n")
    return len(rules)


if __name__ == "__main__":
    n = write_merges(sys.argv[1], sys.argv[2], noise="--noise" in sys.argv[3:])
    print(f"{sys.argv[2]}: {n} rules")
