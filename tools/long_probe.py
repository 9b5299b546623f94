"""Times hut_encode_batch on documents holding one long single-class run (configs[4] shape): how the block-cooperative
path scales with the pretoken length, in the opt-in no-limit mode (max_pretoken), checked against the oracle's no-limit
mode.  usage (GPU box): python tools/long_probe.py [model] [--no-check]"""
import json
import os
import random
import sys
import time

sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__))]
import numpy as np  # noqa: E402
import synth  # noqa: E402
from hutoken_b200 import capi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "gpt2"
check = "--no-check" not in sys.argv
merges = None
if "--merges" in sys.argv:  # id-keyed ranks from a merges file (core.c:211-337): monotone, so huge_kernel runs its rounds
    import subprocess
    import tempfile
    merges = os.path.join(tempfile.mkdtemp(), name + ".merges.txt")
    subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "make_merges.py"),
                    synth.PATHS[name][0], merges], check=True, stdout=subprocess.DEVNULL)
v, s = synth.PATHS[name]
eng = capi.Engine(v, s, max_pretoken=32 << 20, merges_path=merges, **synth.MODEL_KW[name])
orc = None
if check:
    from oracle import oracle as O
    orc = O.Oracle(v, s, no_limit=True, merges_path=merges, **synth.MODEL_KW[name])
rng = random.Random(1)
eng.batch_encode(["warm up " + "x" * 5000])
rows = []
for n in (4096, 65536, 262144, 1 << 20, 4 << 20, 16 << 20):
    for kind, alphabet in (("random letters", "abcdefghijklmnopqrstuvwxyz"), ("digits", "0123456789"), ("one letter", "a")):
        if kind == "one letter" and n > (1 << 20):
            continue  # one chain of overlapping pairs: the exact single-lane order, quadratic-free but serial
        doc = "x " + "".join(rng.choices(alphabet, k=n - 2))
        data = np.frombuffer(doc.encode() + b"\0" * 16, dtype=np.uint8)
        offs = np.array([0, len(doc)], dtype=np.uint64)
        eng.encode_packed(data, offs)  # sizes the buffers
        t0 = time.perf_counter()
        ids, _ = eng.encode_packed(data, offs)
        dt = time.perf_counter() - t0
        st = eng.stats()
        ok = None
        if orc is not None:
            t1 = time.perf_counter()
            oids, _ = orc.encode_packed(data, offs)
            cpu = time.perf_counter() - t1
            ok = bool(np.array_equal(ids, oids))
        else:
            cpu = None
        rows.append({"model": name + ("+merges" if merges else ""), "bytes": n, "kind": kind, "ids": int(len(ids)), "wall_ms": dt * 1e3,
                     "kernel_ms": st["kernel_ms"], "oracle_cpu_ms": None if cpu is None else cpu * 1e3, "equal": ok})
        print(json.dumps(rows[-1]), flush=True)
