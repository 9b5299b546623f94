"""Measures what the host link gives on this box: pinned H2D alone, D2H alone, both at once (two streams), in
chunks of the size hut_encode_batch streams with.  Evidence for the e2e leg of bench.py (PCIe-bound).
usage: python tools/pcie_probe.py [total_mb] [chunk_mb]"""
import json
import sys

import torch

total = (int(sys.argv[1]) if len(sys.argv) > 1 else 1024) << 20
chunk = (int(sys.argv[2]) if len(sys.argv) > 2 else 32) << 20
dev = torch.device("cuda:0")
h_in = torch.empty(total, dtype=torch.uint8).pin_memory()
h_out = torch.empty(total, dtype=torch.uint8).pin_memory()
d_in = torch.empty(total, dtype=torch.uint8, device=dev)
d_out = torch.ones(total, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(up, down, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        s1.wait_event(e0)
        s2.wait_event(e0)
        for o in range(0, total, chunk):
            if up:
                with torch.cuda.stream(s1):
                    d_in[o:o + chunk].copy_(h_in[o:o + chunk], non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    h_out[o:o + chunk].copy_(d_out[o:o + chunk], non_blocking=True)
        e1.record(s1)
        e2.record(s2)
        torch.cuda.synchronize()
        best = min(best, max(e0.elapsed_time(e1), e0.elapsed_time(e2)))
    return total / best / 1e6  # GB/s per direction


res = {"total_mb": total >> 20, "chunk_mb": chunk >> 20, "h2d_alone_GBs": run(True, False),
       "d2h_alone_GBs": run(False, True), "both_per_direction_GBs": run(True, True)}
print(json.dumps(res))
