"""Vocab file tooling: the on-disk format `initialize` reads (lib.c:243-378) is one line per
token, ``0xHH0xHH... == id`` with upper-case hex of the token's bytes.

    python tools/convert_vocab.py json   <vocab.json> <vocab.txt>    # the job of the reference's scripts/convert.py
    python tools/convert_vocab.py fix    <trained.txt> <vocab.txt>   # make a bpe_train output loadable
    python tools/convert_vocab.py check  <vocab.txt>                 # the loader's own acceptance rules

``json``: a Hugging-Face style ``vocab.json`` (token string -> id).  Like the reference script, ids are the running
line number in the file's own key order (scripts/convert.py:3-14 ignores the JSON values); ``--keep-ids`` writes
the JSON's ids instead, sorted by id.
``fix``: drops the ``0x00 == 0`` line `initialize` rejects (lib.c:345-355) and renumbers the ids densely in file
order (bpe_train skips id 256, bpe.c:171).
Data tooling only: nothing here is on the encode path.
"""
from __future__ import annotations

import json
import re
import sys

_LINE = re.compile(rb"^((?:0x[0-9A-Fa-f]{2})+) == (-?\d+)$")


def hex_key(raw: bytes) -> str:
    return "".join(f"0x{b:02X}" for b in raw)


def from_json(json_path, out_path, keep_ids=False):
    with open(json_path, encoding="utf-8") as f:
        d = json.load(f)
    items = sorted(d.items(), key=lambda kv: kv[1]) if keep_ids else list(d.items())
    with open(out_path, "w", encoding="ascii", newline="\n") as f:
        for n, (tok, tid) in enumerate(items):
            f.write(f"{hex_key(tok.encode('utf-8'))} == {tid if keep_ids else n}\n")
    return len(items)


def parse(path):
    """[(token bytes, id)] in file order; stops at the first empty line like the loader's fgets loop."""
    out = []
    with open(path, "rb") as f:
        for ln, line in enumerate(f, 1):
            line = line.rstrip(b"\r\n")
            if not line:
                break
            m = _LINE.match(line)
            if not m:
                raise ValueError(f"{path}:{ln}: not a '0xHH.. == id' line")
            out.append((bytes.fromhex(m.group(1).decode().replace("0x", "")), int(m.group(2))))
    return out


def fix_trained(in_path, out_path):
    rows = [(k, v) for k, v in parse(in_path) if k != b"\x00"]
    with open(out_path, "w", encoding="ascii", newline="\n") as f:
        for n, (k, _) in enumerate(rows):
            f.write(f"{hex_key(k)} == {n}\n")
    return len(rows)


def check(path):
    """The conditions under which `initialize` accepts the file; returns a list of problems (empty = loadable)."""
    problems = []
    rows = parse(path)
    if not rows:
        problems.append("no vocab lines")
    for n, (k, v) in enumerate(rows, 1):
        if b"\x00" in k:
            problems.append(f"line {n}: token contains 0x00 (lib.c:345-355)")
        if v < 0:
            problems.append(f"line {n}: negative id")
    ids = [v for _, v in rows]
    if ids and max(ids) >= len(rows):
        problems.append(f"largest id {max(ids)} >= number of lines {len(rows)}: decode() bounds ids by the line count "
                        "(core.c:520-531)")
    return problems


def main(argv):
    if len(argv) >= 3 and argv[0] == "json":
        n = from_json(argv[1], argv[2], keep_ids="--keep-ids" in argv[3:])
    elif len(argv) == 3 and argv[0] == "fix":
        n = fix_trained(argv[1], argv[2])
    elif len(argv) == 2 and argv[0] == "check":
        problems = check(argv[1])
        for p in problems:
            print(p)
        return 1 if problems else 0
    else:
        print(__doc__)
        return 2
    print(f"{argv[2]}: {n} lines")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
