"""configs[0] of BASELINE.json: "hutoken.bpe_train 5000-merge vocab on synthetic ASCII/Hungarian text, then
hutoken.encode of a 1 MB string on CPU".  Run in the dev container (needs the reference build oracle/_ref):

    python tools/make_c1.py

1. trains with the UNMODIFIED reference: _hutoken.bpe_train(text, 5256, "c1.txt") on 200 KB of the synthetic
   generator's text (it fits bpe_train's stack arrays, bpe.c:252-253).  The reference writes to $HOME/config/
   (helper.c:131-162); HOME is pointed into tools/_build/ for the call.
2. post-processes the file as SURVEY.md F6 documents, because initialize() cannot load bpe_train's own output: the
   `0x00 == 0` line is dropped (lib.c:345-355 rejects it) and the ids are renumbered densely in file order (bpe_train
   skips id 256, bpe.c:171).  Result: data/c1_5k.vocab.txt (+ data/empty_special.txt).
3. encodes a 1 MB string (synthetic generator, seed 7) with the reference and records count + SHA-256 of the int32 id
   stream in tests/golden/c1_encode.json; the text itself is regenerated from the seed by the tests.
"""
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools")]
import numpy as np  # noqa: E402
import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402

HOME = os.path.join(ROOT, "tools", "_build", "c1_home")
VOCAB = os.path.join(synth.DATA_DIR, "c1_5k.vocab.txt")
SPECIAL = os.path.join(synth.DATA_DIR, "empty_special.txt")
TRAIN_BYTES, TRAIN_SEED, TEXT_BYTES, TEXT_SEED = 200 << 10, 6, 1 << 20, 7


def c1_text(seed, nbytes):
    """One string: the generator's documents joined by newlines (no NUL), cut to nbytes at a character boundary."""
    data, offs = synth.corpus(seed, nbytes + (64 << 10))
    raw = b"\n".join(bytes(d) for d in synth.docs_of(data, offs))[:nbytes]
    return raw.decode("utf-8", "ignore")


def train():
    os.makedirs(HOME, exist_ok=True)
    text = c1_text(TRAIN_SEED, TRAIN_BYTES)
    code = ("import sys; sys.path.insert(0, %r); from oracle import oracle as O; ref = O.load_ref(); "
            "ref.bpe_train(open(%r, encoding='utf-8').read(), 5256, 'c1.txt')")
    tpath = os.path.join(HOME, "train.txt")
    with open(tpath, "w", encoding="utf-8") as f:
        f.write(text)
    subprocess.run([sys.executable, "-c", code % (ROOT, tpath)], check=True, env=dict(os.environ, HOME=HOME))
    return os.path.join(HOME, "config", "c1.txt")


def postprocess(src, dst):
    lines = [l for l in open(src, "rb").read().split(b"\n") if l]
    kept = [l for l in lines if not l.startswith(b"0x00 ==")]
    with open(dst, "wb") as f:
        for i, l in enumerate(kept):
            f.write(l.split(b" == ")[0] + b" == %d\n" % i)
    return len(lines), len(kept)


def main():
    O.build(ref=True)
    raw = train()
    n_raw, n_kept = postprocess(raw, VOCAB)
    open(SPECIAL, "w").close()
    ref = O.load_ref()
    ref.initialize(VOCAB, SPECIAL, None, False)
    text = c1_text(TEXT_SEED, TEXT_BYTES)
    ids = np.asarray(ref.encode(text), dtype=np.int32)
    g = {"vocab": os.path.relpath(VOCAB, ROOT), "special": os.path.relpath(SPECIAL, ROOT), "text_seed": TEXT_SEED,
         "text_bytes": len(text.encode("utf-8")), "n_ids": int(len(ids)), "unknown_ids": int((ids < 0).sum()),
         "sha256_int32_le": hashlib.sha256(ids.tobytes()).hexdigest(), "first_ids": ids[:64].tolist(),
         "bpe_train_lines": n_raw, "vocab_lines": n_kept}
    with open(os.path.join(ROOT, "tests", "golden", "c1_encode.json"), "w") as f:
        json.dump(g, f, indent=1)
    print(g)


if __name__ == "__main__":
    main()
