"""Measures what the host link gives on this box: pinned H2D alone, D2H alone, both at once (two streams per device), in
chunks of the size hut_encode_batch streams with -- on 1, 2, 4, ... GPUs AT THE SAME TIME (one host thread per
device), so that the aggregate ceiling of the end-to-end leg at N > 1 is a measured number and not a guess.
usage: python tools/pcie_probe.py [total_mb_per_gpu] [chunk_mb] [--gpus 1,2,4,8]"""
import json
import sys
import threading

import torch

args = [a for a in sys.argv[1:] if not a.startswith("--")]
total = (int(args[0]) if len(args) > 0 else 1024) << 20
chunk = (int(args[1]) if len(args) > 1 else 32) << 20
ngpu = torch.cuda.device_count()
counts = [1]
for a in sys.argv[1:]:
    if a.startswith("--gpus"):
        counts = [int(x) for x in (a.split("=", 1)[1] if "=" in a else sys.argv[sys.argv.index(a) + 1]).split(",")]
counts = [c for c in counts if c <= ngpu]


class Dev:
    def __init__(self, i):
        self.dev = torch.device("cuda", i)
        torch.cuda.set_device(i)
        self.h_in = torch.empty(total, dtype=torch.uint8).pin_memory()
        self.h_out = torch.empty(total, dtype=torch.uint8).pin_memory()
        self.d_in = torch.empty(total, dtype=torch.uint8, device=self.dev)
        self.d_out = torch.ones(total, dtype=torch.uint8, device=self.dev)
        self.s1, self.s2 = torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev)

    def run(self, up, down, barrier, out, k):
        torch.cuda.set_device(self.dev)
        torch.cuda.synchronize(self.dev)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        barrier.wait()
        e0.record(torch.cuda.current_stream(self.dev))
        self.s1.wait_event(e0)
        self.s2.wait_event(e0)
        for o in range(0, total, chunk):
            if up:
                with torch.cuda.stream(self.s1):
                    self.d_in[o:o + chunk].copy_(self.h_in[o:o + chunk], non_blocking=True)
            if down:
                with torch.cuda.stream(self.s2):
                    self.h_out[o:o + chunk].copy_(self.d_out[o:o + chunk], non_blocking=True)
        e1.record(self.s1)
        e2.record(self.s2)
        torch.cuda.synchronize(self.dev)
        out[k] = max(e0.elapsed_time(e1), e0.elapsed_time(e2))


devs = [Dev(i) for i in range(max(counts))]


def measure(n, up, down, reps=4):
    best = 1e18
    for _ in range(reps):
        out = [0.0] * n
        bar = threading.Barrier(n)
        ths = [threading.Thread(target=devs[k].run, args=(up, down, bar, out, k)) for k in range(n)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        best = min(best, max(out))
    return n * total / best / 1e6  # aggregate GB/s per direction


res = {"total_mb_per_gpu": total >> 20, "chunk_mb": chunk >> 20, "gpus_visible": ngpu, "runs": []}
for n in counts:
    res["runs"].append({"gpus": n, "h2d_alone_GBs": measure(n, True, False), "d2h_alone_GBs": measure(n, False, True),
                        "both_per_direction_GBs": measure(n, True, True)})
print(json.dumps(res))
