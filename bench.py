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
    ap.add_argument("--no-workloads", action="store_true", help="skip the short runs of the other configs (workloads object)")
    ap.add_argument("--no-latency", action="store_true", help="skip the per-call latency leg (latency object)")
    ap.add_argument("--llama-mb", type=int, default=10240, help="configs[2] corpus in all, MiB (split over the ranks)")
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


def heavy_tail(data, offs, seed, frac=0.05):
    """Splices long single-class runs into a corpus so that the rarely taken paths get real work: about `frac` of the
    bytes end up in pretokens of 16..31 bytes (second cache) and of 40..4000 bytes (long_kernel): long words, digit
    runs, hex / base64-like blobs with one character class.  Returns (data, offs) with the same documents, grown."""
    rng = np.random.default_rng(seed)
    nd = len(offs) - 1
    total = int(offs[-1])
    budget = int(total * frac)
    lens = []
    while budget > 0:
        n = int(rng.integers(16, 32)) if rng.random() < 0.85 else int(min(4000, max(40, rng.lognormal(5.3, 1.0))))
        lens.append(n)
        budget -= n + 1
    lens = np.array(lens, dtype=np.int64)
    where = np.sort(rng.integers(0, nd, size=len(lens)))  # run k is appended to document where[k]
    alph = [np.frombuffer(b"abcdefghijklmnopqrstuvwxyz", np.uint8), np.frombuffer(b"0123456789", np.uint8),
            np.frombuffer(b"ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz", np.uint8)]
    kinds = rng.integers(0, 3, size=len(lens))
    out = np.empty(total + int(lens.sum()) + len(lens) + 64, dtype=np.uint8)
    new_offs = np.empty(nd + 1, dtype=np.uint64)
    at, k = 0, 0
    for d in range(nd):
        new_offs[d] = at
        lo, hi = int(offs[d]), int(offs[d + 1])
        out[at:at + hi - lo] = data[lo:hi]
        at += hi - lo
        while k < len(lens) and where[k] == d:
            out[at] = 0x20
            a = alph[kinds[k]]
            out[at + 1:at + 1 + lens[k]] = a[rng.integers(0, len(a), size=lens[k])]
            at += 1 + int(lens[k])
            k += 1
    new_offs[nd] = at
    out[at:at + 64] = 0
    return out[:at + 64], new_offs


def long_run_docs(seed, sizes):
    """configs[4] shape: documents holding one single-class run of the given sizes inside normal text."""
    import synth
    rng = np.random.default_rng(seed)
    base, boffs = synth.corpus(seed, 64 << 10)
    filler = base[:int(boffs[4])].tobytes()
    docs = []
    letters = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz", np.uint8)
    for n in sizes:
        run = letters[rng.integers(0, 26, size=n)].tobytes()
        docs.append(filler[:300] + b" " + run + b" " + filler[300:600])
    offs = np.zeros(len(docs) + 1, dtype=np.uint64)
    offs[1:] = np.cumsum([len(d) for d in docs])
    return np.frombuffer(b"".join(docs) + b"\0" * 64, dtype=np.uint8), offs


def run_ours(args):
    import hashlib
    import torch
    import torch.distributed as dist
    import synth
    import ctypes as C
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
    stream = torch.cuda.current_stream(dev)
    sh = stream.cuda_stream
    peak, peak_src = measured_peak()
    check_bytes = 8 << 20

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def make_engine(model, **kw):
        v, s = synth.PATHS[model]
        return capi.Engine(v, s, devices=[local], chunk_mb=args.chunk_mb, **synth.MODEL_KW[model], **kw)

    def make_oracle(model, **kw):
        from oracle import oracle as O
        v, s = synth.PATHS[model]
        return O.Oracle(v, s, **synth.MODEL_KW[model], **kw)

    def device_leg(eng, model, data, offs, nbytes, steps, warmup, sync_all):
        """`value` leg: text and offsets resident in HBM, ids stay there (hut_encode_device).  CUDA events on the launch
        stream; the pretoken cache is reset inside every step."""
        ndocs = len(offs) - 1
        d_text = torch.empty(nbytes + 64, dtype=torch.uint8, device=dev)
        d_text.copy_(torch.from_numpy(data[:nbytes + 64]), non_blocking=False)
        d_offs = torch.from_numpy(offs.astype(np.int64)).to(dev)
        ids_cap = nbytes + 4096  # byte mode, one symbol per byte: T <= B_in
        if model.startswith("llama"):
            ids_cap = nbytes + 8 * ndocs + 4096
        d_ids = torch.empty(ids_cap, dtype=torch.int32, device=dev)
        d_tok = torch.empty(ndocs + 1, dtype=torch.int64, device=dev)

        def step():
            eng.cache_reset()
            return eng.encode_device(0, d_text.data_ptr(), nbytes, d_offs.data_ptr(), ndocs, d_ids.data_ptr(), ids_cap,
                                     d_tok.data_ptr(), stream=sh)

        n_ids = 0
        for _ in range(warmup):
            n_ids = step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        main_ms, launches, st = [], 0, None
        if sync_all:
            barrier()
        else:
            torch.cuda.synchronize(dev)
        e0.record(stream)
        for _ in range(steps):
            n_ids = step()
            st = eng.stats()
            main_ms.append(st["main_kernel_ms"])
            launches += int(st["kernel_launches"])
        e1.record(stream)
        if sync_all:
            barrier()
        else:
            torch.cuda.synchronize(dev)
        return {"ms": e0.elapsed_time(e1) / steps, "main_ms": float(np.mean(main_ms)), "n_ids": int(n_ids),
                "launches": launches, "stats": st, "d_text": d_text, "d_ids": d_ids, "d_tok": d_tok, "ndocs": ndocs}

    def parity(orc, data, offs, leg, limit):
        """ids + offsets of the first `limit` bytes against the oracle: (equal, sha256 of the engine's id prefix, probe stats)"""
        nd = int(np.searchsorted(offs, np.uint64(limit), side="right")) - 1
        nd = max(1, min(nd, len(offs) - 1))
        oids, otok = orc.encode_packed(data[:int(offs[nd]) + 1], offs[:nd + 1])
        after = orc.stats()  # the oracle's counters cover its last call only (reset at the start of every call)
        got = leg["d_ids"][:len(oids)].cpu().numpy()
        gtok = leg["d_tok"][:nd + 1].cpu().numpy().astype(np.uint64)
        ok = bool(len(got) == len(oids) and np.array_equal(got, oids) and np.array_equal(gtok, otok))
        probes = {k: after[k] for k in ("pretokens", "symbols", "tokens", "probes")}
        probes["bytes"] = int(offs[nd])
        return ok, hashlib.sha256(got.tobytes()).hexdigest(), probes

    def summarize(leg, nbytes, probes=None):
        algo = nbytes + 4 * leg["n_ids"] + 8 * (leg["ndocs"] + 1)
        out = {"value": nbytes / (leg["ms"] * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": leg["ms"],
               "tokens_per_s": leg["n_ids"] / (leg["ms"] * 1e-3), "bytes": int(nbytes), "docs": int(leg["ndocs"]),
               "roofline": {"kernel": "hut::tile_kernel", "kernel_ms": leg["main_ms"],
                            "achieved": algo / (leg["main_ms"] * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                            "frac": algo / (leg["main_ms"] * 1e-3) / 1e9 / peak,
                            "frac_whole_step": algo / (leg["ms"] * 1e-3) / 1e9 / peak,
                            "algorithmic_bytes_per_launch": int(algo)},
               "stats_last_step": {k: leg["stats"][k] for k in ("pretokens", "cache_hits", "warp_merges", "block_merges",
                                                                "big_merges", "cut_documents")}}
        if probes and probes["bytes"]:
            # SURVEY 8(d) secondary figure: lookups the ALGORITHM needs (no dedup credit), 8 B per (key, value) slot,
            # counted by the oracle on the parity prefix and scaled to the shard by bytes
            p_total = probes["probes"] * (nbytes / probes["bytes"])
            out["roofline"]["bytes_incl_probes"] = int(algo + 8 * p_total)
            out["roofline"]["probes_per_input_byte"] = probes["probes"] / probes["bytes"]
        return out

    # =============================== headline workload: configs[1] (or --model / --mb) ===============================
    eng = make_engine(args.model)
    t0 = time.time()
    data, offs, nbytes = make_corpus(2 + rank, args.mb << 20, args.special_every,
                                     pinned_alloc=lambda n: pinned_u8(lib, n))
    ndocs = len(offs) - 1
    gen_s = time.time() - t0
    clocks = ClockSampler(local)
    clocks.start()
    leg = device_leg(eng, args.model, data, offs, nbytes, args.steps, args.warmup, sync_all=True)
    clk = clocks.stop()
    dev_ms, n_ids, launches, st = leg["ms"], leg["n_ids"], leg["launches"], leg["stats"]
    d_text, d_ids, d_tok = leg["d_text"], leg["d_ids"], leg["d_tok"]

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
        # the inverse through the host-buffer call: the ids just returned (pinned host memory) -> text in pinned host
        # memory, H2D of the ids and D2H of the text inside the timed region (hut_decode_batch streams document chunks)
        e2e_dec = None
        if not args.no_decode and len(ids_h):
            try:
                nd_h = len(tok_h) - 1
                res_t = capi._TextResult()
                call = lambda: eng._check(eng._lib.hut_decode_batch(eng._h, ids_h.ctypes.data, tok_h.ctypes.data, nd_h,  # noqa: E731
                                                                    C.byref(res_t)))
                call()
                t0 = time.perf_counter()
                reps = max(1, min(args.steps, 3))
                for _ in range(reps):
                    call()
                dec_ms = (time.perf_counter() - t0) * 1e3 / reps
                txt = np.ctypeslib.as_array(res_t.text, shape=(int(res_t.n_bytes),))
                same = bool(int(res_t.n_bytes) == nbytes and np.array_equal(txt, data[:nbytes]))
                e2e_dec = {"value": int(res_t.n_bytes) / (dec_ms * 1e-3) / 1e6, "unit": "decoded MB/s", "ms_per_call": dec_ms,
                           "h2d_bytes": int(4 * len(ids_h) + 8 * (nd_h + 1)), "d2h_bytes": int(res_t.n_bytes) + 8 * (nd_h + 1),
                           "round_trip_equals_input": same, "path": eng.decode_stats()["path"],
                           "scope": "rank 0's own shard (not summed over ranks; at N > 1 all ranks share the host link)",
                           "api": "hut_decode_batch (ids in pinned host memory -> text in pinned host memory)"}
            except Exception as ex:  # the decode leg must never take the encode numbers down with it
                e2e_dec = {"error": str(ex)[:200]}

    # ---- decode leg: the ids just produced, resident on the device, back to text ----
    def decode_once(eng_x, lg, tok_ptr, nd, steps):
        """(ms per call, text length, text address, text_off address, stats) of hut_decode_device on ids left in HBM"""
        n_text = 0
        for _ in range(2):
            eng_x.decode_device(lg["d_ids"].data_ptr(), lg["n_ids"], tok_ptr, nd, stream=sh)
        d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        d0.record(stream)
        for _ in range(steps):
            t_ptr, o_ptr, n_text = eng_x.decode_device(lg["d_ids"].data_ptr(), lg["n_ids"], tok_ptr, nd, stream=sh)
        d1.record(stream)
        torch.cuda.synchronize(dev)
        return d0.elapsed_time(d1) / steps, int(n_text), t_ptr, o_ptr, eng_x.decode_stats()

    def decode_leg(eng_x, lg, nb, steps, orc_x=None):
        """The batch as the encode leg left it, then the same ids as ONE document (the reference's own benchmark shape,
        scripts/benchmark.py:52-56).  Checked against the input text where the round trip is exact, and against the
        oracle's decode of the first documents when an oracle is given."""
        try:
            nd = lg["ndocs"]
            ms, n_text, t_ptr, o_ptr, dst = decode_once(eng_x, lg, lg["d_tok"].data_ptr(), nd, steps)
            # round trip: byte mode with a complete byte alphabet decodes back to the input exactly
            same = _device_equal(torch, t_ptr, lg["d_text"], nb, dev) if n_text == nb else None
            oracle_ok = None
            if orc_x is not None:
                k = min(nd, 64)
                tok_h = lg["d_tok"][:k + 1].cpu().numpy()
                ids_h = lg["d_ids"][:int(tok_h[-1])].cpu().numpy()
                toff = torch.as_tensor(_DevView(o_ptr, 8 * (k + 1)), device=dev).cpu().numpy().view(np.uint64)
                text_h = torch.as_tensor(_DevView(t_ptr, max(1, int(toff[-1]))), device=dev).cpu().numpy()
                oracle_ok = all(text_h[int(toff[d]):int(toff[d + 1])].tobytes() ==
                                orc_x.decode_bytes(ids_h[int(tok_h[d]):int(tok_h[d + 1])]) for d in range(k))
            one = torch.tensor([0, lg["n_ids"]], dtype=torch.int64, device=dev)
            ms1, n1, _, _, dst1 = decode_once(eng_x, lg, one.data_ptr(), 1, steps)
            return {"value": n_text / (ms * 1e-3) / 1e6, "unit": "decoded MB/s (ids resident in HBM)", "ms_per_step": ms,
                    "kernel_ms": dst["kernel_ms"], "gpu_launches_per_step": int(dst["kernel_launches"]),
                    "text_bytes": n_text, "round_trip_equals_input": same, "first_docs_equal_oracle": oracle_ok,
                    "path": dst.get("path"),
                    "single_document": {"value": n1 / (ms1 * 1e-3) / 1e6, "ms_per_step": ms1, "text_bytes": n1,
                                        "path": dst1.get("path"), "ms_vs_batch": ms1 / ms}}
        except ValueError as e:  # ids the vocabulary cannot print (-1): nothing to decode
            return {"value": None, "note": str(e)}

    leg["n_ids"], leg["ndocs"] = n_ids, ndocs
    dec = None

    # ---- id check of THIS rank's shard (every rank, so the scaling runs carry one per GPU) ----
    check_ok, check_sha, probes = None, None, None
    if args.check or world > 1 or not args.no_workloads:
        orc = make_oracle(args.model)
        check_ok, check_sha, probes = parity(orc, data, offs, leg, check_bytes)
        if not args.no_decode:
            dec = decode_leg(eng, leg, nbytes, max(1, min(args.steps, 5)), orc)
        orc.close()
        if not check_ok:
            raise SystemExit(f"bench.py: rank {rank}: GPU ids differ from the oracle on the first {check_bytes} bytes")
    elif not args.no_decode:
        dec = decode_leg(eng, leg, nbytes, max(1, min(args.steps, 5)))

    # ---- max over ranks / whole-job sums ----
    import dist_torch as D  # tools/dist_torch.py: torch.distributed plumbing (the product package imports no PyTorch)

    def allmax(x):
        return D.all_reduce_scalar(x, "max", dev)

    def allsum(x):
        return D.all_reduce_scalar(x, "sum", dev)

    dev_ms_max = allmax(dev_ms)
    total_bytes = allsum(float(nbytes))
    total_ids = allsum(float(n_ids))
    e2e_ms_max = allmax(e2e_ms) if e2e_ms is not None else None
    launches_total = int(allsum(float(launches)))
    checks_ok = int(allsum(1.0 if check_ok else 0.0)) if check_ok is not None else None
    head = summarize(leg, nbytes, probes)
    del d_text, d_ids, d_tok, leg
    torch.cuda.empty_cache()

    # =============================== the other configs, one short run each (same JSON line) ===============================
    workloads = {}
    if not args.no_workloads:
        def extra(name, model, data_x, offs_x, eng_x, orc_kw, steps, limit, note, with_decode=False):
            nb = int(offs_x[-1])
            lg = device_leg(eng_x, model, data_x, offs_x, nb, steps, 1, sync_all=False)
            orc_x = make_oracle(model, **orc_kw)
            ok, sha, pr = parity(orc_x, data_x, offs_x, lg, limit)
            dec_x = decode_leg(eng_x, lg, nb, 2, orc_x) if with_decode and not args.no_decode else None
            orc_x.close()
            w = summarize(lg, nb, pr)
            if dec_x is not None:
                w["decode"] = dec_x
            w["parity"] = {"equal_to_oracle": ok, "prefix_bytes": int(pr["bytes"]), "sha256_int32_ids": sha}
            w["config"] = note
            ms_all = allmax(lg["ms"])
            w["whole_job_value"] = allsum(float(nb)) / (ms_all * 1e-3) / 1e6
            workloads[name] = w
            del lg
            torch.cuda.empty_cache()

        # configs[2]: Llama-shaped 32k vocab, 10 GiB in all, split over the ranks (so the scaling run yields the
        # strong-scaling curve north_star asks for)
        if args.model != "llama":
            mb2 = max(64, args.llama_mb // world)
            eng2 = make_engine("llama")
            d2, o2, nb2 = make_corpus(300 + rank, mb2 << 20, 0, pinned_alloc=lambda n: pinned_u8(lib, n))
            extra("configs[2] llama_10GiB_total", "llama", d2, o2, eng2, {}, 2, check_bytes,
                  f"Llama/PULI-LlumiX-shaped 32K vocab (prefix U+2581, plain mode), {mb2} MiB on this rank = "
                  f"{mb2 * world} MiB over {world} rank(s) (strong scaling)", with_decode=True)
            eng2.close()
            lib.hut_host_free(d2.ctypes.data)
            del d2, o2
        # configs[3]: a special byte / multibyte character every ~200 bytes
        d3, o3, nb3 = make_corpus(400 + rank, 256 << 20, 200, pinned_alloc=None)
        extra("configs[3] special_every_200", args.model, d3, o3, eng, {}, 3, check_bytes,
              "configs[1] model, one replaced byte / multibyte character every ~200 bytes, 256 MiB per rank",
              with_decode=True)
        # heavy tail: 5 % of the bytes in pretokens of 16..31 and 40..4000 bytes (second cache, long_kernel)
        d5, o5 = heavy_tail(d3[:(96 << 20)], o3[:int(np.searchsorted(o3, np.uint64(96 << 20), side="right"))], 500 + rank)
        extra("heavy_tail_mix", args.model, d5, o5, eng, {}, 3, check_bytes,
              "configs[1] model, 96 MiB + 5 % of the bytes spliced in as single-class runs of 16..31 / 40..4000 bytes")
        del d3, o3, d5, o5
        # configs[4]: long single-class runs really merged (opt-in no-limit mode; the reference cuts such documents)
        eng4 = make_engine(args.model, max_pretoken=32 << 20)
        d4, o4 = long_run_docs(600 + rank, [16 << 10, 64 << 10, (256 << 10) - 8, (256 << 10) + 8, 16 << 20])
        extra("configs[4] long_runs_no_limit", args.model, d4, o4, eng4, {"no_limit": True}, 1, 1 << 30,
              "5 documents holding one letters-only run of 16 Ki / 64 Ki / 256 Ki - 8 / 256 Ki + 8 / 16 Mi bytes (SURVEY 8(d) C5), "
              "max_pretoken = 32 MiB (no cut), checked in full against the oracle's no-limit mode")
        eng4.close()
        del d4, o4

    # =============================== latency of the drop-in calls (rank 0) ===============================
    latency = None
    if rank == 0 and not args.no_latency:
        latency = measure_latency(args.model)

    if rank == 0:
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
            "roofline": dict(head["roofline"], bound="hbm", traffic=ncu_traffic(args), peak_source=peak_src,
                             kernel_share_of_step=head["roofline"]["kernel_ms"] / dev_ms),
            "stats_last_step": head["stats_last_step"],
            "check": None if checks_ok is None else {"ranks_equal_to_oracle": checks_ok, "ranks": world,
                                                     "prefix_bytes": check_bytes, "sha256_rank0": check_sha},
            "corpus_gen_s": gen_s,
        }
        if workloads:
            line["workloads"] = workloads
        if latency is not None:
            line["latency"] = latency
        if dec is not None:
            line["decode"] = dec
        if e2e_ms_max is not None:
            line["e2e"] = {"value": total_bytes / (e2e_ms_max * 1e-3) / 1e6, "unit": UNIT,
                           "h2d_bytes_per_step": int(nbytes + 8 * (ndocs + 1)),
                           "d2h_bytes_per_step": int(4 * n_ids + 8 * (ndocs + 1)),
                           "ms_per_step": e2e_ms_max,
                           "api": "hut_encode_batch (host buffers in pinned memory -> ids in pinned memory)"}
            if e2e_dec is not None:
                line["e2e"]["decode"] = e2e_dec
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


def ncu_traffic(args):
    """DRAM bytes of the dominant kernel from the committed ncu capture of this very workload (per launch), else None"""
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["tile_kernel"]
        if args.model == "gpt2" and args.mb == 1024 and args.special_every == 0:
            return int(tj["dram_bytes_read"] + tj["dram_bytes_write"])
    except Exception:
        pass
    return None


def measure_latency(model):
    """The reference's own per-call bound (tests/test_tokenizer.py:107-130: mean encode(sentence1) < 1 ms over 10 000
    calls, the same for decode(encode(..))), through the drop-in Python surface, with the unmodified reference beside
    it; plus the list API (one PyLong per id) on a 64 MB sample beside the packed call."""
    import synth
    import hutoken_b200 as hutoken
    v, s = synth.PATHS[model]
    kw = synth.MODEL_KW[model]
    sentence1 = "How can the net amount of entropy of the universe be massively decreased?"
    out = {"sentence_bytes": len(sentence1), "calls": 10000}

    def timed(fn, n):
        ts = np.empty(n)
        for i in range(n):
            t0 = time.perf_counter()
            fn()
            ts[i] = time.perf_counter() - t0
        return {"mean_us": float(ts.mean() * 1e6), "p50_us": float(np.percentile(ts, 50) * 1e6),
                "p99_us": float(np.percentile(ts, 99) * 1e6)}

    hutoken.initialize(v, s, prefix=kw["prefix"], is_byte_encoder=kw["is_byte_encoder"])
    ids = hutoken.encode(sentence1)
    for _ in range(200):
        hutoken.encode(sentence1)
    out["encode"] = timed(lambda: hutoken.encode(sentence1), 10000)
    try:
        hutoken.decode(ids)
        out["decode_of_encode"] = timed(lambda: hutoken.decode(hutoken.encode(sentence1)), 10000)
    except Exception as e:  # a vocabulary that cannot print every id
        out["decode_of_encode"] = {"note": str(e)}
    # list API vs packed API on the same 64 MB of documents
    data, offs = synth.corpus(77, 64 << 20)
    docs = [d.decode("utf-8") for d in synth.docs_of(data, offs)]
    hutoken.batch_encode(docs[:64], 1)
    t0 = time.perf_counter()
    res = hutoken.batch_encode(docs, 1)
    t_first = time.perf_counter() - t0  # includes growing the pinned / device buffers to this batch's size
    del res
    t0 = time.perf_counter()
    res = hutoken.batch_encode(docs, 1)
    t_list = time.perf_counter() - t0
    t0 = time.perf_counter()
    pid, ptok = hutoken.batch_encode_packed(data[:int(offs[-1])], offs)
    t_packed = time.perf_counter() - t0
    nb = int(offs[-1])
    out["batch_encode_list_api"] = {"value": nb / t_list / 1e6, "unit": UNIT, "seconds": t_list, "bytes": nb,
                                    "ids": int(sum(len(r) for r in res)),
                                    "first_call_seconds": t_first,
                                    "note": "hutoken.batch_encode(list of str), second call of this size: two UTF-8 passes "
                                            "+ a list entry per id (one int object per id VALUE)"}
    out["batch_encode_packed_api"] = {"value": nb / t_packed / 1e6, "unit": UNIT, "seconds": t_packed,
                                      "ids": int(len(pid))}
    # the unmodified reference on the same calls
    try:
        from oracle import oracle as O
        ref = O.load_ref()
        if ref is not None:
            ref.initialize(v, s, kw["prefix"], kw["is_byte_encoder"])
            for _ in range(200):
                ref.encode(sentence1)
            r = {"encode": timed(lambda: ref.encode(sentence1), 10000)}
            try:
                rids = ref.encode(sentence1)
                ref.decode(rids)
                r["decode_of_encode"] = timed(lambda: ref.decode(ref.encode(sentence1)), 10000)
            except Exception as e:
                r["decode_of_encode"] = {"note": str(e)}
            nthreads = host_threads()
            sub = docs[:len(docs) // 8]
            nbs = sum(len(d.encode("utf-8")) for d in sub)
            t0 = time.perf_counter()
            ref.batch_encode(sub, nthreads)
            dt = time.perf_counter() - t0
            r["batch_encode_list_api"] = {"value": nbs / dt / 1e6, "unit": UNIT, "seconds": dt, "bytes": nbs,
                                          "threads": nthreads}
            out["reference"] = r
    except Exception as e:
        out["reference"] = {"note": f"{type(e).__name__}: {e}"}
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
