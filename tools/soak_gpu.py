"""Longer randomised parity run on the GPU than the test-suite affords:  python tools/soak_gpu.py [seconds]
 * fuzz vocabularies (tests/fuzzmodels.py: non-monotone ranks, duplicate keys / ids, virtual units) on batches large enough
   for the general path (tile_kernel with ticketed tiles, batched merges), not only the one-launch small path,
 * the standard models on fresh corpora with a thrashing pretoken cache (2^10 .. 2^14 slots: inserts, evictions and
   look-ups of many warps on the same slots),
every result compared id for id with the oracle.  Prints one summary line; exits non-zero on the first difference."""
import os
import random
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import numpy as np  # noqa: E402

import fuzzmodels as F  # noqa: E402
import synth  # noqa: E402
from common import pack  # noqa: E402
from hutoken_b200 import capi  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 180.0
    t_end = time.time() + budget
    tmp = tempfile.mkdtemp()
    n_models = n_corpora = 0
    n_ids = 0
    seed = 0
    while time.time() < t_end:
        seed += 1
        rng = random.Random(50_000 + seed)
        if seed % 3:
            byte_mode = bool(seed & 1)
            v, s, kw = F.random_model(rng, tmp, byte_mode, f"soak{seed}")
            try:
                eng = capi.Engine(v, s, **kw)
            except Exception as e:  # a model shape the engine refuses: not a parity question
                if "UNSUPPORTED" in str(e) or "unsupported" in str(e) or "split" in str(e) or "span" in str(e):
                    continue
                raise
            orc = O.Oracle(v, s, **kw)
            texts = [F.random_text(rng, rng.randint(0, 80)) for _ in range(6000)]  # ~200 KB: the general path
            data, offs = pack(texts)
            ids, tok = eng.encode_packed(data, offs)
            oids, otok = orc.encode_packed(data, offs)
            n_models += 1
        else:
            name = ("gpt2", "llama", "gpt2_small", "llama_small")[(seed // 3) % 4]
            v, s = synth.PATHS[name]
            kw = synth.MODEL_KW[name]
            eng = capi.Engine(v, s, cache_log2_slots=rng.choice([10, 12, 14, 22]), **kw)
            orc = O.Oracle(v, s, **kw)
            data, offs = synth.corpus(9000 + seed, rng.choice([1, 3, 6]) << 20, special_every=rng.choice([0, 150, 2000]))
            ids, tok = eng.encode_packed(data, offs)
            ids2, tok2 = eng.encode_packed(data, offs)  # the same again on the now warm (or thrashed) cache
            assert np.array_equal(ids, ids2) and np.array_equal(tok, tok2), ("warm run differs", seed, name)
            oids, otok = orc.encode_packed(data, offs)
            n_corpora += 1
        if not (np.array_equal(ids, oids) and np.array_equal(tok, otok)):
            bad = int(np.argmax(ids[:min(len(ids), len(oids))] != oids[:min(len(ids), len(oids))]))
            print(f"DIFFERENCE at seed {seed}: {len(ids)} vs {len(oids)} ids, first at {bad}")
            sys.exit(1)
        n_ids += len(ids)
        eng.close()
        orc.close()
    print(f"soak ok: {n_models} fuzz vocabularies on the general path, {n_corpora} corpora on small / default caches, "
          f"{n_ids} ids equal to the oracle, {budget:.0f} s")


if __name__ == "__main__":
    main()
