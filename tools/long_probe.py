"""Times hut_encode_batch on documents holding one long single-class run (configs[4] shape): how the block-cooperative
path scales with the pretoken length.  usage (GPU box): python tools/long_probe.py [model]"""
import os
import random
import sys
import time

sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__))]
import numpy as np  # noqa: E402
import synth  # noqa: E402
from hutoken_b200 import capi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "gpt2"
v, s = synth.PATHS[name]
eng = capi.Engine(v, s, **synth.MODEL_KW[name])
rng = random.Random(1)
eng.batch_encode(["warm up"])
for n in (4096, 16384, 65536, 131072, 262144):
    for kind, alphabet in (("random letters", "abcdefghijklmnopqrstuvwxyz"), ("one letter", "a")):
        doc = "x " + "".join(rng.choice(alphabet) for _ in range(n - 2))
        t0 = time.perf_counter()
        ids = eng.batch_encode([doc])[0]
        dt = time.perf_counter() - t0
        st = eng.stats()
        print(f"{name} run of {n:7d} bytes ({kind}): {dt * 1e3:9.2f} ms  {len(ids)} ids  kernel_ms {st['kernel_ms']:.2f}",
              flush=True)
