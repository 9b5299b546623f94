"""End-to-end throughput of ONE process driving N devices through one context (HUTOKEN_B200_DEVICES path of the library,
hut_encode_batch dealing chunks round-robin): host buffers in pinned memory -> ids in pinned memory, compared with
the oracle on a prefix.  The counterpart of bench.py's N-process e2e leg.
usage: python tools/multi_dev_e2e.py [n_devices] [mb_per_device] [chunk_mb]"""
import json
import os
import sys
import time

sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__))]
import numpy as np  # noqa: E402
import synth  # noqa: E402
import bench  # noqa: E402
from hutoken_b200 import capi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
chunk_mb = int(sys.argv[3]) if len(sys.argv) > 3 else 0
lib = capi.load_library()
v, s = synth.PATHS["gpt2"]
eng = capi.Engine(v, s, devices=list(range(n)), chunk_mb=chunk_mb, **synth.MODEL_KW["gpt2"])
data, offs, nbytes = bench.make_corpus(2, (mb * n) << 20, 0, pinned_alloc=lambda k: bench.pinned_u8(lib, k))
for _ in range(2):
    eng.cache_reset()
    eng.encode_packed(data.ctypes.data, offs, copy=False)
steps = 5
t0 = time.perf_counter()
for _ in range(steps):
    eng.cache_reset()
    ids, tok = eng.encode_packed(data.ctypes.data, offs, copy=False)
dt = (time.perf_counter() - t0) / steps
from oracle import oracle as O  # noqa: E402
orc = O.Oracle(v, s, **synth.MODEL_KW["gpt2"])
nd = int(np.searchsorted(offs, np.uint64(8 << 20), side="right")) - 1
oids, otok = orc.encode_packed(data[:int(offs[nd]) + 1], offs[:nd + 1])
ok = bool(np.array_equal(ids[:len(oids)], oids) and np.array_equal(tok[:nd + 1], otok))
print(json.dumps({"devices_in_one_context": n, "mb_per_device": mb, "e2e_MBps": nbytes / dt / 1e6, "ms_per_step": dt * 1e3,
                  "ids": int(len(ids)), "prefix_equal_to_oracle": ok,
                  "per_device_pretokens": [eng.stats(i)["pretokens"] for i in range(n)]}))
