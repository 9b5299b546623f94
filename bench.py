#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mb 1024] [--model gpt2]

A "step" is one batch_encode pass of the hot path over one batch of synthetic documents: at N=1 the workload
is BASELINE.json configs[1] -- GPT-2-shaped byte-level BPE (50k merges, is_byte_encoder=True, 68 special-char
lines), 1 GB of synthetic documents on one B200.  Under torchrun (N>1) every rank encodes its own 1 GB shard of
documents (weak scaling; documents are independent, so there is no data-path collective -- SURVEY.md 8(e));
torch.distributed/NCCL is used only for the barrier and the max-over-ranks of the timings.

Printed JSON (rank 0, one line): value = input MB/s with the text resident in HBM (hut_encode_device);
e2e = the same through hut_encode_batch with HOST buffers (pinned), H2D of the text and D2H of the ids inside
the timed region; roofline = the fused tile kernel against the measured HBM copy bandwidth, on ALGORITHMIC bytes
B_in + 4*T + 8*(N_docs+1); cpu_baseline = the unmodified reference (oracle/_ref) batch_encode with all host
threads on a bounded sample of the same documents.  The oracle / reference are only ever the baseline or the
checker here, never the measured product.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tools")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "encode_input_MB_per_s"
UNIT = "MB/s"
CHUNK = 256 << 20


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mb", type=int, default=1024, help="synthetic documents per GPU, MiB (configs[1]: 1 GB)")
    ap.add_argument("--model", default="gpt2", choices=["gpt2", "llama", "gpt2_small", "llama_small"])
    ap.add_argument("--special-every", type=int, default=0)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--chunk-mb", type=int, default=0, help="hut_options.chunk_mb of the e2e leg (0 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-decode", action="store_true", help="skip the decode leg (ids -> text on the device)")
    ap.add_argument("--check", action="store_true", help="also verify a sample of the output against the oracle")
    return ap.parse_args()


def make_corpus(seed, total_bytes, special_every, pinned_alloc=None):
    """Seeded documents, generated in parallel 256 MiB chunks (tools/hut_synth.c releases the GIL)."""
    import synth
    nchunks = max(1, (total_bytes + CHUNK - 1) // CHUNK)
    sizes = [min(CHUNK, total_bytes - i * CHUNK) for i in range(nchunks)]
    outs = [None] * nchunks
    synth.lexicon()

    def work(i):
        outs[i] = synth.corpus(seed * 1000 + i, sizes[i], special_every=special_every)

    ths = [threading.Thread(target=work, args=(i,)) for i in range(nchunks)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    nbytes = sum(len(d) for d, _ in outs)
    ndocs = sum(len(o) - 1 for _, o in outs)
    if pinned_alloc is not None:
        data = pinned_alloc(nbytes + 64)
    else:
        data = np.empty(nbytes + 64, dtype=np.uint8)
    offs = np.empty(ndocs + 1, dtype=np.uint64)
    at, dn = 0, 0
    for d, o in outs:
        data[at:at + len(d)] = d
        offs[dn:dn + len(o) - 1] = o[:-1] + np.uint64(at)
        at += len(d)
        dn += len(o) - 1
    offs[ndocs] = at
    data[at:at + 64] = 0
    return data, offs, nbytes


class ClockSampler:
    """SM clock + throttle reasons during the timed region (NVML; same fields as the nvidia-smi recipe)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._th = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def start(self):
        if self._nv:
            self._th = threading.Thread(target=self._run, daemon=True)
            self._th.start()

    def stop(self):
        self._stop.set()
        if self._th:
            self._th.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ----------------------------------------------------------------------------------------------------
# CPU reference: the UNMODIFIED reference extension (oracle/_ref), batch_encode with every host thread
# ----------------------------------------------------------------------------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def load_reference(model):
    """(callable docs -> n_tokens, kind).  kind 'reference' = oracle/_ref .so; 'port' = oracle restatement."""
    import synth
    from oracle import oracle as O
    v, s = synth.PATHS[model]
    kw = synth.MODEL_KW[model]
    nthreads = host_threads()
    ref = O.load_ref()
    if ref is not None:
        ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])

        def run(docs):
            out = ref.batch_encode(docs, nthreads)  # lib.c:722-874: pthreads over a shared task queue
            return sum(len(x) for x in out)
        return run, "reference", nthreads
    orc = O.Oracle(v, s, **kw)

    def run_port(docs):
        data = np.frombuffer(b"".join(docs) + b"\0", dtype=np.uint8)
        offs = np.zeros(len(docs) + 1, dtype=np.uint64)
        offs[1:] = np.cumsum([len(d) for d in docs])
        ids, _ = orc.encode_packed(data, offs)
        return len(ids)
    return run_port, "port", 1


def sample_docs(data, offs, max_bytes, as_str):
    nd = int(np.searchsorted(offs, np.uint64(max_bytes), side="right")) - 1
    nd = max(nd, 1)
    b = data[:int(offs[nd])].tobytes()
    docs = [b[int(offs[i]):int(offs[i + 1])] for i in range(nd)]
    if as_str:
        docs = [d.decode("utf-8") for d in docs]
    return docs, int(offs[nd])


def time_reference(run, kind, data, offs, seconds):
    """Bounded sample: a 4 MB probe sets the rate, then a sample sized for ~`seconds` of CPU work (<= 256 MB)."""
    as_str = kind == "reference"
    docs, nb = sample_docs(data, offs, 4 << 20, as_str)
    t0 = time.perf_counter()
    run(docs)
    rate = nb / (time.perf_counter() - t0)
    target = int(min(max(rate * seconds, 4 << 20), 256 << 20, int(offs[-1])))
    docs, nb = sample_docs(data, offs, target, as_str)
    t0 = time.perf_counter()
    ntok = run(docs)
    dt = time.perf_counter() - t0
    return {"bytes": nb, "docs": len(docs), "tokens": ntok, "seconds": dt}


def run_reference_arm(args):
    """`--impl reference`: the reference's own CPU batch_encode on this box's host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    run, kind, nthreads = load_reference(args.model)
    data, offs, nbytes = make_corpus(2, min(args.mb, 256) << 20, args.special_every)
    per_step = max(2.0, min(args.cpu_seconds, 60.0 / max(1, args.steps + args.warmup)))
    as_str = kind == "reference"
    probe, nb0 = sample_docs(data, offs, 4 << 20, as_str)
    t0 = time.perf_counter()
    run(probe)
    rate = nb0 / (time.perf_counter() - t0)
    docs, nb = sample_docs(data, offs, int(min(max(rate * per_step, 4 << 20), nbytes)), as_str)
    for _ in range(args.warmup):
        run(docs)
    t0 = time.perf_counter()
    ntok = 0
    for _ in range(args.steps):
        ntok = run(docs)
    dt = (time.perf_counter() - t0) / args.steps
    val = nb / dt / 1e6
    sample = f"{nb} bytes / {len(docs)} docs of the same synthetic corpus per step (bounded CPU sample)"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args, nbytes, len(offs) - 1),
        "tokens_per_s": ntok / dt,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": nthreads, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, nbytes, ndocs):
    names = {"gpt2": "configs[1]: GPT-2-shaped byte-level BPE (50k merges, is_byte_encoder=True, 68 special chars) "
                     "batch_encode of synthetic docs",
             "llama": "configs[2]: Llama/PULI-LlumiX-shaped 32K vocab (prefix U+2581, plain mode) batch_encode of "
                      "synthetic docs"}
    return {"workload": names.get(args.model, args.model), "model_files": args.model,
            "mb_per_gpu": args.mb, "special_every": args.special_every,
            "l2": "input (>= 1 GB at the default size) is larger than the 126 MB L2; no flush needed",
            "pretoken_cache": "reset at the start of every step (inside the timed region)",
            "sharding": f"by document, {args.gpus} rank(s), no collective on the data path"}


# ----------------------------------------------------------------------------------------------------
# our arm: sm_100a kernels through the C-ABI
# ----------------------------------------------------------------------------------------------------
def pinned_u8(capi_lib, n):
    import ctypes as C
    p = capi_lib.hut_host_alloc(n)
    if not p:
        raise MemoryError("hut_host_alloc failed")
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n,))


_ALL_CPUS = None


def bind_to_gpu_numa_node(index):
    """Restricts this process to the CPUs NVML reports as local to GPU `index`, so that the pinned text / id
    buffers of the e2e leg are allocated on the socket the GPU hangs off (8 ranks otherwise share one node's
    memory controllers and the inter-socket link).  Best effort: any failure leaves the affinity unchanged."""
    global _ALL_CPUS
    try:
        import pynvml
        _ALL_CPUS = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None


class _DevView:
    """A device buffer owned by the engine, exposed to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _device_equal(torch, ptr, d_text, n, dev):
    got = torch.as_tensor(_DevView(ptr, n), device=dev)
    return bool(torch.equal(got, d_text[:n]))


def run_ours(args):
    import torch
    import torch.distributed as dist
    import synth
    from hutoken_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    bind_to_gpu_numa_node(local)  # before any pinned allocation: first touch places it on the GPU's own node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    lib = capi.load_library()
    v, s = synth.PATHS[args.model]
    eng = capi.Engine(v, s, devices=[local], chunk_mb=args.chunk_mb, **synth.MODEL_KW[args.model])

    # ---- this rank's documents (pinned host memory: the e2e leg copies from it) ----
    t0 = time.time()
    data, offs, nbytes = make_corpus(2 + rank, args.mb << 20, args.special_every,
                                     pinned_alloc=lambda n: pinned_u8(lib, n))
    ndocs = len(offs) - 1
    gen_s = time.time() - t0

    d_text = torch.empty(nbytes + 64, dtype=torch.uint8, device=dev)
    d_text.copy_(torch.from_numpy(data[:nbytes + 64]), non_blocking=False)
    d_offs = torch.from_numpy(offs.astype(np.int64)).to(dev)
    ids_cap = nbytes + 4096  # byte mode, one symbol per byte: T <= B_in
    if args.model.startswith("llama"):
        ids_cap = nbytes + 8 * ndocs + 4096
    d_ids = torch.empty(ids_cap, dtype=torch.int32, device=dev)
    d_tok = torch.empty(ndocs + 1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev)
    sh = stream.cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step_device():
        eng.cache_reset()
        return eng.encode_device(0, d_text.data_ptr(), nbytes, d_offs.data_ptr(), ndocs, d_ids.data_ptr(), ids_cap,
                                 d_tok.data_ptr(), stream=sh)

    # ---- device-resident leg: `value` ----
    n_ids = 0
    for _ in range(args.warmup):
        n_ids = step_device()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main_ms, launches = [], 0
    barrier()
    clocks.start()
    e0.record(stream)
    for _ in range(args.steps):
        n_ids = step_device()
        st = eng.stats()
        main_ms.append(st["main_kernel_ms"])
        launches += int(st["kernel_launches"])
    e1.record(stream)
    barrier()
    clk = clocks.stop()
    dev_ms = e0.elapsed_time(e1) / args.steps

    # ---- end-to-end leg through the host-buffer C-ABI call: H2D text + kernels + D2H ids every step ----
    e2e_ms = None
    if not args.no_e2e:
        for _ in range(min(args.warmup, 3)):
            eng.cache_reset()
            eng.encode_packed(data.ctypes.data, offs, copy=False)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            eng.cache_reset()
            ids_h, tok_h = eng.encode_packed(data.ctypes.data, offs, copy=False)
            last = int(ids_h[-1]) if len(ids_h) else 0  # the host reads the result  # noqa: F841
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps

    # ---- decode leg: the ids just produced, resident on the device, back to text ----
    dec = None
    if not args.no_decode:
        try:
            n_text = 0
            for _ in range(2):
                _, _, n_text = eng.decode_device(d_ids.data_ptr(), n_ids, d_tok.data_ptr(), ndocs, stream=sh)
            d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            dsteps = max(1, min(args.steps, 5))
            torch.cuda.synchronize(dev)
            d0.record(stream)
            for _ in range(dsteps):
                t_ptr, _, n_text = eng.decode_device(d_ids.data_ptr(), n_ids, d_tok.data_ptr(), ndocs, stream=sh)
            d1.record(stream)
            torch.cuda.synchronize(dev)
            dms = d0.elapsed_time(d1) / dsteps
            dst = eng.decode_stats()
            # round trip: byte mode with a complete byte alphabet decodes back to the input exactly
            same = None
            if n_text == nbytes:
                same = _device_equal(torch, t_ptr, d_text, nbytes, dev)
            dec = {"value": n_text / (dms * 1e-3) / 1e6, "unit": "decoded MB/s (ids resident in HBM)", "ms_per_step": dms,
                   "kernel_ms": dst["kernel_ms"], "gpu_launches_per_step": int(dst["kernel_launches"]),
                   "text_bytes": int(n_text), "round_trip_equals_input": same}
        except ValueError as e:  # ids the vocabulary cannot print (-1): nothing to decode
            dec = {"value": None, "note": str(e)}

    # ---- max over ranks / whole-job sums (hutoken_b200/dist.py; covered on CPU by tests/test_dist_cpu.py) ----
    from hutoken_b200 import dist as D

    def allmax(x):
        return D.all_reduce_scalar(x, "max", dev)

    def allsum(x):
        return D.all_reduce_scalar(x, "sum", dev)

    dev_ms_max = allmax(dev_ms)
    total_bytes = allsum(float(nbytes))
    total_ids = allsum(float(n_ids))
    e2e_ms_max = allmax(e2e_ms) if e2e_ms is not None else None
    launches_total = int(allsum(float(launches)))

    if args.check:
        from oracle import oracle as O
        orc = O.Oracle(v, s, **synth.MODEL_KW[args.model])
        nd = int(np.searchsorted(offs, np.uint64(8 << 20), side="right")) - 1
        oids, otok = orc.encode_packed(data[:int(offs[nd]) + 1], offs[:nd + 1])
        got = d_ids[:len(oids)].cpu().numpy()
        assert np.array_equal(got, oids), "bench --check: GPU ids differ from the oracle"
        assert np.array_equal(d_tok[:nd + 1].cpu().numpy().astype(np.uint64), otok)

    if rank == 0:
        peak, peak_src = measured_peak()
        k_ms = float(np.mean(main_ms))
        algo = nbytes + 4 * n_ids + 8 * (ndocs + 1)
        achieved = algo / (k_ms * 1e-3) / 1e9
        traffic = None
        try:  # DRAM bytes of the dominant kernel from the committed ncu capture of this very workload
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["tile_kernel"]
            if args.model == "gpt2" and args.mb == 1024 and args.special_every == 0:
                traffic = int(tj["dram_bytes_read"] + tj["dram_bytes_write"])
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": total_bytes / (dev_ms_max * 1e-3) / 1e6, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms_max, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args, nbytes, ndocs),
            "tokens_per_s": total_ids / (dev_ms_max * 1e-3),
            "tokens_per_step": int(total_ids), "bytes_per_token": total_bytes / max(total_ids, 1),
            "bytes_per_gpu": int(nbytes), "docs_per_gpu": int(ndocs),
            "gpu_launches": launches_total,
            "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "hut::tile_kernel (+ miss_kernel)", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(algo), "kernel_ms": k_ms,
                         "kernel_share_of_step": k_ms / dev_ms},
            "stats_last_step": {k: st[k] for k in ("pretokens", "cache_hits", "warp_merges", "block_merges",
                                                   "big_merges")},
            "corpus_gen_s": gen_s,
        }
        if dec is not None:
            line["decode"] = dec
        if e2e_ms_max is not None:
            line["e2e"] = {"value": total_bytes / (e2e_ms_max * 1e-3) / 1e6, "unit": UNIT,
                           "h2d_bytes_per_step": int(nbytes + 8 * (ndocs + 1)),
                           "d2h_bytes_per_step": int(4 * n_ids + 8 * (ndocs + 1)),
                           "ms_per_step": e2e_ms_max,
                           "api": "hut_encode_batch (host buffers in pinned memory -> ids in pinned memory)"}
        if not args.no_cpu_baseline and world >= 1:
            try:
                if _ALL_CPUS:
                    os.sched_setaffinity(0, _ALL_CPUS)  # the CPU baseline gets every host core
                run, kind, nthreads = load_reference(args.model)
                r = time_reference(run, kind, data, offs, args.cpu_seconds)
                line["cpu_baseline"] = {
                    "value": r["bytes"] / r["seconds"] / 1e6, "unit": UNIT, "cores": nthreads, "kind": kind,
                    "tokens_per_s": r["tokens"] / r["seconds"],
                    "sample": f"first {r['bytes']} bytes / {r['docs']} docs of rank 0's corpus, "
                              f"{r['seconds']:.1f} s wall, batch_encode(num_threads={nthreads})"}
            except Exception as e:  # the baseline must never take the measurement down
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable",
                                        "sample": f"{type(e).__name__}: {e}"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
