"""The heavy-tail workload of bench.py alone (5 % of the bytes in single-class runs of 16..31 / 40..4000 bytes), for
ncu launch lists of the rarely taken kernels.  usage (GPU box): python tools/heavy_probe.py [model] [mb]"""
import os
import sys

sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__))]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
import synth  # noqa: E402
from hutoken_b200 import capi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "gpt2"
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 96
d3, o3, _ = bench.make_corpus(400, mb << 20, 200)
data, offs = bench.heavy_tail(d3[:(mb << 20)], o3[:int(np.searchsorted(o3, np.uint64(mb << 20), side="right"))], 500)
nbytes = int(offs[-1])
v, s = synth.PATHS[name]
eng = capi.Engine(v, s, **synth.MODEL_KW[name])
dev = torch.device("cuda:0")
d_text = torch.from_numpy(data[:nbytes + 64].copy()).to(dev)
d_offs = torch.from_numpy(offs.astype(np.int64)).to(dev)
cap = nbytes + 8 * len(offs) + 4096
d_ids = torch.empty(cap, dtype=torch.int32, device=dev)
d_tok = torch.empty(len(offs), dtype=torch.int64, device=dev)
for it in range(4):
    eng.cache_reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = eng.encode_device(0, d_text.data_ptr(), nbytes, d_offs.data_ptr(), len(offs) - 1, d_ids.data_ptr(), cap, d_tok.data_ptr())
    e1.record()
    torch.cuda.synchronize()
    print(it, "ms", round(e0.elapsed_time(e1), 3), "ids", n, eng.stats())
