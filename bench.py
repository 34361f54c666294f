#!/usr/bin/env python
"""bench.py -- scored requests/s of the scoring hot path on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this build (CUDA, C ABI)
    python bench.py --impl reference --steps K --warmup W    # the CPU port of the reference path

A "step" is one pass of the hot path over one batch of the workload: BASELINE.json config 2
(4,096 requests x 16 candidates x 1536-dim fp32 embeddings, 4 weighted voters) per GPU.
With N > 1 (launched under torchrun, one rank per GPU) every rank scores its own batch of
4,096 requests -- requests are independent, so the path shards with no collective -- and
`value` is all ranks' requests / max-over-ranks device time ("scaling": "weak").

  value     device-resident: inputs already in HBM when the timed region starts; K back-to-back
            launches timed with CUDA events on the launching stream; three distinct input
            batches (each 403 MB > the 126 MB L2) are rotated so no step re-reads cached lines.
  e2e       the same metric through lwc_score_batch(LWC_MEM_HOST): pinned host buffers in,
            host buffers out, host<->device copies inside the timed region.
  roofline  algorithmic bytes per launch (98,788 B/request) / mean launch time,
            against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline  the oracle's reference-typed port (Decimal tally + f64 embeddings) on all host
            cores, timed on a bounded sample (rank 0, N = 1 only).
  parity_sample a 256-request sample of the timed launch's own outputs against the f64 oracle.
  cfg1      BASELINE.json configs[0] literally: ONE request, single thread, on the CPU port
            (cpu_baseline_cfg1) and through lwc_score_batch(LWC_MEM_HOST) on the GPU (gpu_cfg1).
  extra     the other BASELINE.json configs in the same record: cfg3 (65,536 x 32 x 3072, sharded by
            request over the N ranks, strong scaling), cfg5's pool shape (32,768 x N rows x 1024,
            rows sharded, NCCL all-gather route and the gather fused into the tensor kernel), and at
            N = 1 cfg4 (8,192 x 128 x 3072 on tcgen05), each with its own parity check.
PyTorch is used only for events/streams bookkeeping and the torch.distributed barrier.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "llm-weighted-consensus_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "scored requests/sec (16×1536 emb)"  # BASELINE.json's metric, without its "at 1/2/4/8 B200; % of roofline" tail
UNIT = "requests/s"
WORKLOAD = "cfg2"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        d = json.load(open(path))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(kernel, R):
    """dram bytes per launch from the committed ncu --set full capture, if one exists."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = d.get(kernel)
        if ent and int(ent.get("R", -1)) == R:
            return float(ent["dram_bytes_per_launch"])
    except Exception:
        pass
    return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed regions run."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.stop_flag = False
        self.active = False  # set by the main thread around the timed regions: nothing is queried outside them
        self.max_mhz = None
        self.ok = False
        self.first_delay_s = float(os.environ.get("LWC_BENCH_SAMPLER_DELAY_US", "300")) * 1e-6
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        was_active = False
        while not self.stop_flag:
            if not self.active:
                was_active = False
                time.sleep(0.0002)
                continue
            if not was_active:
                # first sample of a timed region: let the launching thread enqueue its work first (an NVML query takes
                # driver locks; colliding with the first launches of a 1.3 ms region costs microseconds per step)
                was_active = True
                time.sleep(self.first_delay_s)
                if not self.active:
                    continue
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                util = 1
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((mhz, util))
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.003)  # a few samples per timed region; every query contends with the launch path

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        loaded = [m for m, u in self.samples if u > 0] or [m for m, _ in self.samples]
        return {"sm_mhz": float(np.median(loaded)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def host_cores():
    """Threads the CPU arm uses: every core this process may run on.  (torchrun exports OMP_NUM_THREADS=1 to
    its workers; the reference arm is one process that owns the box, so it does not follow that.)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def bind_to_gpu_numa_node(index):
    """Pin this process to the CPUs next to its GPU (NVML's ideal affinity) before any pinned buffer is allocated
    and touched, so the host side of every H2D copy is node-local memory: with 8 ranks pushing 55 GB/s each the
    end-to-end rate is bound by host memory, and remote-node pages halve it.  Returns the previous CPU set."""
    try:
        import pynvml

        before = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(index))
        if not os.sched_getaffinity(0):
            os.sched_setaffinity(0, before)
        return before
    except Exception:
        return None


def cpu_baseline(E, V, w, budget_s=10.0):
    """Reference-typed CPU port on all host cores, bounded sample of the same workload."""
    import oracle

    R = V.shape[0]
    cores = host_cores()
    probe = min(R, 256)
    t = oracle.score_port(E[:probe], V[:probe], w[:probe], oracle.SIM_COSINE, nthreads=cores)["seconds"]
    per_req = max(t / probe, 1e-9)
    n = int(min(R, max(probe, budget_s / per_req)))
    reps, total_req, total_s = 0, 0, 0.0
    while total_s < budget_s and reps < 4096:
        total_s += oracle.score_port(E[:n], V[:n], w[:n], oracle.SIM_COSINE, nthreads=cores)["seconds"]
        total_req += n
        reps += 1
    return {"value": total_req / total_s, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{reps} pass(es) over the first {n} requests of the {WORKLOAD} batch "
                      f"({total_s:.1f} s of compute; Decimal tally + f64 Gram, OpenMP over requests)"}


def workload_string(R, N, D, M):
    """The same string in both arms: the driver compares config.workload to decide whether they ran the same thing."""
    return f"{WORKLOAD}: {R} requests x {N} candidates x {D}-dim fp32, {M} voters, per GPU"


def cfg1_cpu(reps=2000):
    """BASELINE.json configs[0] literally: one request (16 x 1536, 4 weighted voters) on ONE thread -- the reference
    runs its tally on one task (client.rs:384-456) -- Decimal tally + f64 normalise/Gram/vote, per-call compute time."""
    import oracle
    from lwc_b200 import synth

    R, N, D, M = synth.CONFIGS["cfg1"]
    E, V, w, _ = synth.make_batch(R, N, D, M, seed=0x5EED0001, edge_cases=False)
    for _ in range(50):
        oracle.score_port(E, V, w, oracle.SIM_COSINE, nthreads=1)
    ns = np.array([oracle.score_port(E, V, w, oracle.SIM_COSINE, nthreads=1)["seconds"] for _ in range(reps)]) * 1e9
    return {"median_ns": float(np.median(ns)), "p99_ns": float(np.percentile(ns, 99)), "min_ns": float(ns.min()), "reps": reps,
            "threads": 1, "kind": "port", "what": "1 request, 16 x 1536, 4 voters: Decimal tally + f64 Gram, compute phase of one call"}


def run_reference(args):
    """--impl reference: the CPU port of the reference path, all host threads, the product arm's config: every step
    scores the full batch."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # before the OpenMP runtime starts: one thread per core, neighbours close (step times varied 2x without it)
    os.environ.setdefault("OMP_PROC_BIND", "close")
    os.environ.setdefault("OMP_PLACES", "cores")
    import oracle
    from lwc_b200 import synth

    R, N, D, M = synth.CONFIGS[WORKLOAD]
    E, V, w, _ = synth.make_batch(R, N, D, M)
    cores = host_cores()
    for _ in range(args.warmup):
        oracle.score_port(E, V, w, oracle.SIM_COSINE, nthreads=cores)
    secs = np.array([oracle.score_port(E, V, w, oracle.SIM_COSINE, nthreads=cores)["seconds"] for _ in range(args.steps)])
    total = float(np.sum(secs))
    value = R * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64+decimal96",
        "data": "synthetic",
        "config": {"workload": workload_string(R, N, D, M), "step": f"the full batch of {R} requests per step"},
        "step_ms": {"min": float(secs.min() * 1e3), "median": float(np.median(secs) * 1e3), "max": float(secs.max() * 1e3)},
        "value_at_median_step": R / float(np.median(secs)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {R} requests (the whole batch), OMP_PROC_BIND="
                                   f"{os.environ.get('OMP_PROC_BIND')}; the Rust reference cannot be built "
                                   "here (no cargo/rustc), so this is the C restatement in oracle/"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "cpu_baseline_cfg1": cfg1_cpu(),
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def parity_sample(oracle_mod, E, V, w, got, rtol=1e-5, atol=1e-7):
    """A sample of the timed launch's own outputs against the f64 oracle (checker only; runs after the timed
    regions).  tolerance_used = max |got - want| / (rtol |want| + atol): <= 1 means the stated bar holds."""
    want = oracle_mod.score_f64(E, V, w, oracle_mod.SIM_COSINE)
    err = np.abs(got["conf"].astype(np.float64) - want["conf"])
    top = np.maximum(np.abs(want["conf"]).max(axis=1, keepdims=True), 1e-30)
    scored = want["winner"] >= 0
    # a winner is "the same" when it is the oracle's, or the oracle cannot separate the two (gap below fp32 resolution)
    same = got["winner"] == want["winner"]
    rows = np.arange(len(same))
    gw = np.where(got["winner"] >= 0, got["winner"], 0)
    near = scored & (got["winner"] >= 0) & (np.abs(want["conf"][rows, gw] - want["conf"].max(axis=1)) <= 2e-6 * top[:, 0])
    return {"requests": int(len(same)), "max_rel_conf_err": float(np.max(err / np.maximum(np.abs(want["conf"]), 1e-30) *
                                                                           (np.abs(want["conf"]) > 1e-3 * top))),
            "max_conf_err_over_top": float(np.max(err / top)),
            "tolerance_used": float(np.max(err / (rtol * np.abs(want["conf"]) + atol))),
            "tolerance": {"rtol": rtol, "atol": atol},
            "winner_agreement": float(np.mean(same | near)), "winner_exact": float(np.mean(same))}


def h2d_peak(torch, nbytes, reps=5):
    """Plain pinned cudaMemcpyAsync host->device rate of this box (best of `reps`), the ceiling of any host-memory path."""
    src = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    best = 0.0
    for _ in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dst.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    del src, dst
    return best


def cfg1_gpu(ctx, lwc_b200, synth, reps=2000):
    """configs[0] on the GPU: the latency ONE request pays through lwc_score_batch(LWC_MEM_HOST) -- pinned buffers,
    H2D, one launch, D2H, stream synchronise -- wall clock around the C call."""
    import ctypes

    R, N, D, M = synth.CONFIGS["cfg1"]
    E, V, w, _ = synth.make_batch(R, N, D, M, seed=0x5EED0001, edge_cases=False)
    hb = [lwc_b200.HostBuffer(ctx, a.shape, np.float32) for a in (E, V, w)]
    for b, a in zip(hb, (E, V, w)):
        b.array[...] = a
    outs = [lwc_b200.HostBuffer(ctx, sh, t) for sh, t in (((R, N), np.float32), ((R, N), np.float32), ((R, M), np.float32),
                                                           ((R,), np.int32), ((R, N), np.int32))]
    desc = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, lwc_b200.PREC_FP32, lwc_b200.MEM_HOST, 1)
    vp = ctypes.c_void_p
    argv = (ctx.handle, ctypes.byref(desc), *[vp(b.ptr) for b in hb], *[vp(o.ptr) for o in outs], None)
    fn = ctx.lib.lwc_score_batch
    for _ in range(100):
        fn(*argv)
    us = np.empty(reps)
    for i in range(reps):
        t0 = time.perf_counter_ns()
        rc = fn(*argv)
        us[i] = (time.perf_counter_ns() - t0) * 1e-3
        if rc:
            raise lwc_b200.LwcError(rc, ctx.lib.lwc_last_error_message(ctx.handle).decode())
    winner = int(outs[3].array[0])
    for b in hb + outs:
        b.free()
    return {"median_us": float(np.median(us)), "p99_us": float(np.percentile(us, 99)), "min_us": float(us.min()), "reps": reps,
            "api": "lwc_score_batch(LWC_MEM_HOST), 1 request, pinned buffers, blocking", "winner": winner}


def run_b200(args):
    import torch

    import lwc_b200
    from lwc_b200 import synth

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"WORLD_SIZE={world} but --gpus {args.gpus}")
    torch.cuda.set_device(local_rank)
    all_cpus = bind_to_gpu_numa_node(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if args.workload != WORKLOAD:
        return run_other_workload(args, torch, lwc_b200, synth, local_rank, rank, world, dist)
    R, N, D, M = synth.CONFIGS[WORKLOAD]
    # the device-resident leg keeps its batches in HBM for the whole run, so the context may let
    # consecutive launches overlap (LWC_CTX_INPUTS_RESIDENT); the host-memory leg ignores the flag
    ctx = lwc_b200.Context(device=local_rank, flags=0 if args.no_overlap else lwc_b200.CTX_INPUTS_RESIDENT)
    kernel = ctx.kernel_for(R, N, D, M)

    # ---- inputs: one pinned host batch (e2e) + three distinct device-resident batches
    hE = lwc_b200.HostBuffer(ctx, (R, N, D), np.float32)
    E, V, w, _ = synth.make_batch(R, N, D, M, seed=0x5EED0000 + 2 + 1000 * rank, E_out=hE.array)
    hV = lwc_b200.HostBuffer(ctx, V.shape, np.float32)
    hw = lwc_b200.HostBuffer(ctx, w.shape, np.float32)
    hV.array[...] = V
    hw.array[...] = w
    out_host = {k: lwc_b200.HostBuffer(ctx, s, t) for k, s, t in
                (("conf", (R, N), np.float32), ("cw", (R, N), np.float32), ("agree", (R, M), np.float32),
                 ("winner", (R,), np.int32), ("rank", (R, N), np.int32))}

    # Distinct input batches: a fused launch walks up to 32 batches at once with different CTAs in different batches,
    # so every batch of such a run must be its own memory (two list entries sharing a buffer would be read
    # concurrently and meet in L2); across runs, and with one launch per batch, any three batches of 403 MB already
    # defeat the 126 MB L2.
    n_rot = 3 if (args.launch == "loop" or os.environ.get("LWC_LIST_FUSION") == "0") else max(3, min(args.steps, 32))
    dE = [torch.empty((R, N, D), dtype=torch.float32, device="cuda") for _ in range(n_rot)]
    dE[0].copy_(torch.from_numpy(hE.array))
    for i in range(1, n_rot):  # distinct data per rotating batch: a request-wise roll of batch 0
        dE[i].copy_(torch.roll(dE[0], shifts=i * 17, dims=0))
    dV = torch.from_numpy(V).cuda()
    dw = torch.from_numpy(w).cuda()
    d_outs = [{"conf": torch.empty((R, N), dtype=torch.float32, device="cuda"),
               "cw": torch.empty((R, N), dtype=torch.float32, device="cuda"),
               "agree": torch.empty((R, M), dtype=torch.float32, device="cuda"),
               "winner": torch.empty((R,), dtype=torch.int32, device="cuda"),
               "rank": torch.empty((R, N), dtype=torch.int32, device="cuda")} for _ in range(n_rot)]
    d_out = d_outs[0]
    desc_dev = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, lwc_b200.PREC_FP32, lwc_b200.MEM_DEVICE, 1)
    desc_host = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, lwc_b200.PREC_FP32, lwc_b200.MEM_HOST, 1)
    stream = torch.cuda.Stream()

    # The timed region is ONE call of lwc_score_batches with a list of K batches (the three resident batches in
    # rotation).  The library scores a run of same-shaped batches with one launch per 32 batches (persistent CTAs
    # walk the run), so nothing depends on how fast this interpreter -- or the launch path eight ranks share --
    # comes round (round 1: 0.87 scaling efficiency at 8 GPUs with K = 20).  LWC_LIST_FUSION=0 restores one launch
    # per batch, --launch loop one library call per batch.
    def batch_list(k):
        return ctx.batch_list([desc_dev] * k, [dE[j % n_rot].data_ptr() for j in range(k)], [dV.data_ptr()] * k,
                              [dw.data_ptr()] * k, *[[d_outs[j % n_rot][o].data_ptr() for j in range(k)]
                                                     for o in ("conf", "cw", "agree", "winner", "rank")],
                              stream=stream.cuda_stream)

    run_warm, run_timed, run_one = batch_list(args.warmup), batch_list(args.steps), batch_list(1)
    if args.launch == "loop":  # A/B: one lwc_score_batch call per step from this interpreter (round 1's timed loop)
        import ctypes
        _vp, _fn = ctypes.c_void_p, ctx.lib.lwc_score_batch
        _argv = [(ctx.handle, ctypes.byref(desc_dev), _vp(dE[j].data_ptr()), _vp(dV.data_ptr()), _vp(dw.data_ptr()),
                  _vp(d_out["conf"].data_ptr()), _vp(d_out["cw"].data_ptr()), _vp(d_out["agree"].data_ptr()),
                  _vp(d_out["winner"].data_ptr()), _vp(d_out["rank"].data_ptr()), _vp(stream.cuda_stream)) for j in range(n_rot)]

        def run_timed():
            for i in range(args.steps):
                if _fn(*_argv[i % n_rot]):
                    raise lwc_b200.LwcError(500, ctx.lib.lwc_last_error_message(ctx.handle).decode())

    def step_host():
        ctx.score_raw(desc_host, hE.ptr, hV.ptr, hw.ptr, out_host["conf"].ptr, out_host["cw"].ptr,
                      out_host["agree"].ptr, out_host["winner"].ptr, out_host["rank"].ptr)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(local_rank)
    sampler.start()

    # ---- device-resident: W warm-up steps, then exactly K timed steps
    run_warm()
    barrier()
    launches0 = ctx.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active = True
    with torch.cuda.stream(stream):
        ev0.record(stream)
        t_enq = time.perf_counter()
        run_timed()
        enqueue_ms = (time.perf_counter() - t_enq) * 1e3
        ev1.record(stream)
    barrier()
    sampler.active = False
    launches = ctx.launch_count() - launches0
    my_ms = ev0.elapsed_time(ev1)
    dev_ms = max_over_ranks(my_ms)
    per_rank_ms = [my_ms / args.steps]
    if dist is not None:
        tt = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
        dist.all_gather(tt, torch.tensor([my_ms / args.steps], dtype=torch.float64, device="cuda"))
        per_rank_ms = [float(x.item()) for x in tt]
    ms_per_step = dev_ms / args.steps
    value = world * R * args.steps / (dev_ms * 1e-3)
    spread = (max(per_rank_ms) - min(per_rank_ms)) / min(per_rank_ms)
    if spread > 0.03 and rank == 0:
        print(f"[bench] WARNING: per-rank step times differ by {100 * spread:.1f} % (> 3 %): {per_rank_ms}", file=sys.stderr, flush=True)

    # ---- the same K steps with one launch per batch (round 1's mode, LWC_LIST_FUSION=0), for comparison in the same line
    unfused = None
    if args.launch == "list" and os.environ.get("LWC_LIST_FUSION") != "0":
        os.environ["LWC_LIST_FUSION"] = "0"
        ctx_u = lwc_b200.Context(device=local_rank, flags=0 if args.no_overlap else lwc_b200.CTX_INPUTS_RESIDENT)
        os.environ.pop("LWC_LIST_FUSION", None)

        def list_u(k):
            return ctx_u.batch_list([desc_dev] * k, [dE[j % n_rot].data_ptr() for j in range(k)], [dV.data_ptr()] * k,
                                    [dw.data_ptr()] * k, *[[d_outs[j % n_rot][o].data_ptr() for j in range(k)]
                                                           for o in ("conf", "cw", "agree", "winner", "rank")],
                                    stream=stream.cuda_stream)

        list_u(args.warmup)()
        barrier()
        u0, u1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = ctx_u.launch_count()
        run_u = list_u(args.steps)
        with torch.cuda.stream(stream):
            u0.record(stream)
            run_u()
            u1.record(stream)
        barrier()
        u_ms = max_over_ranks(u0.elapsed_time(u1)) / args.steps
        unfused = {"ms_per_step": u_ms, "value": world * R / (u_ms * 1e-3), "unit": UNIT, "gpu_launches": int(ctx_u.launch_count() - n0),
                   "frac": R * synth.algorithmic_bytes_per_request(N, D, M) / (u_ms * 1e-3) / 1e9 / load_peaks()[0],
                   "what": "the same K steps and buffers with LWC_LIST_FUSION=0: one launch per batch, launches overlapped (round 1's timed loop)"}
        ctx_u.close()

    # ---- end to end: host buffers in, host buffers out, copies inside the timed region
    e2e_steps = max(3, min(args.steps, 20))
    for _ in range(2):
        step_host()
    barrier()
    ext = torch.cuda.ExternalStream(ctx.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active = True
    e0.record(ext)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()
    e1.record(ext)
    barrier()
    sampler.active = False
    wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), 0.0))
    if e2e_ms <= 0:
        e2e_ms = max_over_ranks(wall_ms)
    e2e_value = world * R * e2e_steps / (e2e_ms * 1e-3)
    h2d = hE.nbytes + hV.nbytes + hw.nbytes
    d2h = sum(b.nbytes for b in out_host.values())

    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- the same call with the embeddings crossing PCIe as f64 (the reference's element type, rounded on the device)
    # and as bf16 (opt-in transport, 1e-2 tolerance): what the element type on the wire costs / buys end to end
    transports = {}
    if world == 1 and not args.no_extras:
        import ctypes

        for name, et, dt in (("f64", lwc_b200.ELEM_F64, np.float64), ("bf16", lwc_b200.ELEM_BF16, np.uint16)):
            hT = lwc_b200.HostBuffer(ctx, (R, N, D), dt)
            if et == lwc_b200.ELEM_F64:
                hT.array[...] = hE.array
            else:
                hT.array[...] = ctx.stage_f64_bf16(hE.array.astype(np.float64))
            argv = (ctx.handle, ctypes.byref(desc_host), et, ctypes.c_void_p(hT.ptr), ctypes.c_void_p(hV.ptr), ctypes.c_void_p(hw.ptr),
                    *[ctypes.c_void_p(out_host[k].ptr) for k in ("conf", "cw", "agree", "winner", "rank")], None)
            for _ in range(2):
                ctx._check(ctx.lib.lwc_score_batch_typed(*argv))
            t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t1[0].record(ext)
            for _ in range(e2e_steps):
                ctx._check(ctx.lib.lwc_score_batch_typed(*argv))
            t1[1].record(ext)
            torch.cuda.synchronize()
            ms = t1[0].elapsed_time(t1[1]) / e2e_steps
            transports[name] = {"value": R / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
                                "h2d_bytes_per_step": hT.nbytes + hV.nbytes + hw.nbytes}
            transports[name]["winners"] = out_host["winner"].array.copy()
            hT.free()
        step_host()  # leave the fp32 transport's results in the host buffers for the checks below
        for name in transports:
            wn = transports[name].pop("winners")
            transports[name]["winner_agreement_with_fp32_transport"] = float(np.mean(wn == out_host["winner"].array))

    # the two paths must agree bit for bit on the same batch (device batch 0 == host batch)
    run_one()
    torch.cuda.synchronize()
    dev_np = {k: v.cpu().numpy() for k, v in d_out.items()}
    same = all(bool(np.array_equal(dev_np[k], out_host[k].array, equal_nan=True)) for k in dev_np)

    # plain pinned H2D rate of this box: what "PCIe-bound" is measured against (all ranks copy at once, as in the e2e leg)
    barrier()
    pcie = h2d_peak(torch, hE.nbytes)
    barrier()

    peak, peak_src = load_peaks()
    bytes_per_launch = R * synth.algorithmic_bytes_per_request(N, D, M)
    achieved = bytes_per_launch / (ms_per_step * 1e-3) / 1e9
    e2e_gbs = (h2d + d2h) / (e2e_ms / e2e_steps * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(R, N, D, M),
                   "similarity": "cosine", "precision": "fp32-grade (3xTF32 compensated Gram)",
                   "kernel": kernel, "parallelism": f"request-sharded x{world}, no collectives",
                   "l2": f"{n_rot} distinct input batches of {hE.nbytes / 1e6:.0f} MB each, visited in rotation (> 126 MB L2; no buffer "
                         "appears twice in one fused launch)",
                   "launches": ("serialised" if args.no_overlap else
                                "programmatic dependent launch: a launch streams its (resident) inputs while the previous one drains")
                               + "; the K timed steps are one lwc_score_batches call: a run of same-shaped batches is scored by one launch per 32 batches"},
        "clocks": sampler.summary(), "per_rank_ms_per_step": per_rank_ms, "per_rank_spread": spread,
        "host_enqueue_ms": enqueue_ms, "one_launch_per_batch": unfused,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps, "wall_ms_per_step": wall_ms / e2e_steps,
                "api": "lwc_score_batch(LWC_MEM_HOST), pinned host buffers",
                "achieved_gbs": e2e_gbs, "h2d_peak_gbs": pcie, "frac_of_h2d_peak": e2e_gbs / pcie,
                "bound": "PCIe: plain pinned cudaMemcpyAsync H2D of the same bytes on this box = h2d_peak_gbs",
                "other_transports": transports or None},
        "gpu_launches": int(launches),
        # per STEP (one batch): a fused launch scores steps_per_launch batches, so bytes and time both scale with it
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": load_traffic(kernel, R), "kernel": kernel, "peak_source": peak_src,
                     "bytes_per_launch": bytes_per_launch * args.steps / max(int(launches), 1),
                     "bytes_per_step": bytes_per_launch, "steps_per_launch": args.steps / max(int(launches), 1),
                     "traffic_is": "DRAM bytes per step (batch) from the committed ncu capture",
                     "frac_of_nominal_8000": achieved / 8000.0},
        "host_device_paths_agree": same,
    }
    if rank == 0:
        import oracle  # checker only: after every timed region

        sl = slice(0, 256)
        line["parity_sample"] = parity_sample(oracle, hE.array[sl], V[sl], w[sl], {k: v[sl] for k, v in dev_np.items()})
    if world == 1 and not args.no_cfg1:
        line["gpu_cfg1"] = cfg1_gpu(ctx, lwc_b200, synth)
    if not args.no_extras:
        line["extra"] = run_extras(args, torch, lwc_b200, synth, ctx, dist, rank, world, local_rank)
    if world == 1 and not args.no_cpu_baseline:
        if all_cpus:
            os.sched_setaffinity(0, all_cpus)  # the CPU arm owns the whole box again
        line["cpu_baseline"] = cpu_baseline(hE.array, V, w)
        line["cpu_baseline_cfg1"] = cfg1_cpu()
        if not args.no_host_leg:
            line["e2e_from_reference_types"] = host_types_leg(local_rank)
    if rank == 0:
        print(json.dumps(line), flush=True)
    for b in (hE, hV, hw, *out_host.values()):
        b.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def host_types_leg(device, with_cpu=True):
    """End to end FROM THE REFERENCE'S OWN TYPES: tools/e2e_host (C++) builds each shape as per-request
    CreateEmbeddingResponse-shaped rows of Vec<f64> plus Option<Vec<f64>> votes, submits them to lwc::Batcher from
    several host threads and collects Tally futures; wall clock.  cfg2's shape is the headline; cfg3's and cfg4's
    shapes show how the ratio to the CPU port moves with the pool size (the CPU's Gram cost grows with N^2, the bytes
    the host has to convert and ship with N).  Returns the tool's JSON per shape (or why it did not run)."""
    import subprocess

    exe = os.path.join(ROOT, "tools", "e2e_host")
    if not os.path.exists(exe):
        return {"unavailable": "tools/e2e_host is not built (python __graft_entry__.py)"}
    out = {}
    shapes = (("cfg2_shape", 4096, 16, 1536, 4, 5), ("cfg3_shape", 1024, 32, 3072, 4, 4), ("cfg4_shape", 256, 128, 3072, 16, 4))
    for name, R, N, D, M, rounds in shapes:
        try:
            run = subprocess.run([exe, "--device", str(device), "--requests", str(R), "--n", str(N), "--d", str(D), "--m", str(M),
                                  "--rounds", str(rounds), "--batch", str(max(32, min(512, R // 4)))],
                                 capture_output=True, text=True, timeout=300)
            lines = [l for l in run.stdout.splitlines() if l.startswith("{")]
            if run.returncode != 0 or not lines:
                out[name] = {"unavailable": f"tools/e2e_host exited {run.returncode}: {run.stderr[-300:]}"}
                continue
            out[name] = json.loads(lines[-1])
            if out[name].get("convert_only_requests_per_s"):
                # the bare f64 -> fp32 pass over the same heap rows by the same threads: what "host-bound" means here
                out[name]["frac_of_convert_only"] = out[name]["value"] / out[name]["convert_only_requests_per_s"]
            if with_cpu:
                import oracle
                from lwc_b200 import synth

                n = min(R, 256 if N <= 32 else 64)
                E, V, w, _ = synth.make_batch(n, N, D, M, seed=9, edge_cases=False)
                cores = host_cores()
                t, reps = 0.0, 0
                while t < 3.0 and reps < 64:
                    t += oracle.score_port(E, V, w, oracle.SIM_COSINE, nthreads=cores)["seconds"]
                    reps += 1
                out[name]["cpu_port_same_shape"] = {"value": n * reps / t, "unit": UNIT, "cores": cores,
                                                    "sample": f"{reps} pass(es) over {n} requests (fp32 inputs already contiguous)"}
                out[name]["ratio_to_cpu_port"] = out[name]["value"] / out[name]["cpu_port_same_shape"]["value"]
        except Exception as exc:  # a side leg must never cost the bench line
            out[name] = {"unavailable": repr(exc)}
    return out


def device_batch(torch, R, N, D, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    e = torch.randn((R, N, D), dtype=torch.float32, device="cuda", generator=g)
    e *= torch.exp(0.5 * torch.randn((R, N, 1), dtype=torch.float32, device="cuda", generator=g))
    return e


def timed_side_shape(torch, lwc_b200, synth, ctx, dist, R, N, D, M, prec, steps, warmup, seed=7, n_rot=1):
    """K device-resident launches of one shape (inputs generated in HBM), max over ranks; plus a 256-request
    parity sample of the outputs.  Returns (ms_per_step, kernel name, parity dict or None)."""
    from lwc_b200 import shard

    # side shapes rotate one or two big batches: a fused launch would read the same buffer from several CTAs at once
    # (L2 hits), so these are timed with one launch per batch
    prev = os.environ.get("LWC_LIST_FUSION")
    os.environ["LWC_LIST_FUSION"] = "0"
    ctx = lwc_b200.Context(device=torch.cuda.current_device())
    if prev is None:
        os.environ.pop("LWC_LIST_FUSION", None)
    else:
        os.environ["LWC_LIST_FUSION"] = prev
    kernel = ctx.kernel_for(R, N, D, M, precision=prec)
    dE = [device_batch(torch, R, N, D, seed + i) for i in range(n_rot)]
    V = synth.votes(R, N, M, seed=11)
    w = synth.weights(R, M, seed=11)
    dV, dw = torch.from_numpy(V).cuda(), torch.from_numpy(w).cuda()
    outs = [torch.empty((R, N), dtype=torch.float32, device="cuda"), torch.empty((R, N), dtype=torch.float32, device="cuda"),
            torch.empty((R, M), dtype=torch.float32, device="cuda"), torch.empty((R,), dtype=torch.int32, device="cuda"),
            torch.empty((R, N), dtype=torch.int32, device="cuda")]
    desc = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, prec, lwc_b200.MEM_DEVICE, 1)
    stream = torch.cuda.Stream()

    def lst(k):
        return ctx.batch_list([desc] * k, [dE[j % n_rot].data_ptr() for j in range(k)], [dV.data_ptr()] * k, [dw.data_ptr()] * k,
                              *[[o.data_ptr()] * k for o in outs], stream=stream.cuda_stream)

    lst(warmup)()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    run = lst(steps)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        run()
        ev1.record(stream)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = shard.max_over_ranks(dist, ev0.elapsed_time(ev1) / steps, device="cuda")
    par = None
    if dist is None or dist.get_rank() == 0:
        import oracle  # checker only

        j = (steps - 1) % n_rot  # the batch the last launch scored
        n = min(R, 256)
        got = {"conf": outs[0][:n].cpu().numpy(), "winner": outs[3][:n].cpu().numpy()}
        tf32 = prec == lwc_b200.PREC_TF32
        par = parity_sample(oracle, dE[j][:n].cpu().numpy(), V[:n], w[:n], got, rtol=2e-3 if tf32 else 1e-5,
                            atol=1e-5 if tf32 else 1e-7)
    del dE, outs
    ctx.close()
    torch.cuda.empty_cache()
    return ms, kernel, par


def run_extras(args, torch, lwc_b200, synth, ctx, dist, rank, world, local_rank):
    """The other BASELINE.json configs, in the record the driver writes: cfg3 strong-scaled by request over the
    ranks, cfg5's pool shape with rows sharded over the ranks (both exchange routes), cfg4 at N = 1."""
    from lwc_b200 import shard

    peak, _ = load_peaks()
    extra = {}
    steps, warmup = max(3, min(args.steps, 10)), 3

    # ---- cfg3: 65,536 x 32 x 3072, fp32-grade, requests sharded over the ranks (no collective)
    R_total, N, D, M = synth.CONFIGS["cfg3"]
    b, e = shard.request_block(R_total, rank, world)
    ms, kernel, par = timed_side_shape(torch, lwc_b200, synth, ctx, dist, e - b, N, D, M, lwc_b200.PREC_FP32, steps, warmup)
    gbs = (e - b) * synth.algorithmic_bytes_per_request(N, D, M) / (ms * 1e-3) / 1e9
    extra["cfg3_strong"] = {"ms": ms, "value": R_total / (ms * 1e-3), "unit": UNIT, "requests_total": R_total,
                            "requests_per_gpu": e - b, "kernel": kernel, "achieved_gbs_per_gpu": gbs, "frac": gbs / peak,
                            "efficiency_input": "value at n_gpus / (n_gpus x value at 1)", "steps": steps,
                            "l2": f"one batch of {(e - b) * N * D * 4 / 1e9:.2f} GB per GPU (> 126 MB L2)", "parity_sample": par}

    # ---- cfg4 (1 GPU only): 8,192 x 128 x 3072, 16 voters, tcgen05, both precisions
    if world == 1:
        R4, N4, D4, M4 = synth.CONFIGS["cfg4"]
        for name, prec in (("cfg4_fp32", lwc_b200.PREC_FP32), ("cfg4_tf32", lwc_b200.PREC_TF32)):
            ms, kernel, par = timed_side_shape(torch, lwc_b200, synth, ctx, None, R4, N4, D4, M4, prec, steps, warmup)
            gbs = R4 * synth.algorithmic_bytes_per_request(N4, D4, M4) / (ms * 1e-3) / 1e9
            extra[name] = {"ms": ms, "value": R4 / (ms * 1e-3), "unit": UNIT, "kernel": kernel, "achieved_gbs": gbs,
                           "frac": gbs / peak, "algorithmic_tflops": R4 * synth.algorithmic_flops_per_request(N4, D4, M4) / (ms * 1e-3) / 1e12,
                           "steps": steps, "parity_sample": par}

    # ---- cfg5's shape: one pool of 32,768 x world rows x 1024, 16 voters, rows sharded over the ranks
    try:
        extra["cfg5_pool"] = pool_extra(torch, lwc_b200, dist, rank, world, local_rank, steps=5, warmup=2)
    except Exception as exc:  # never lose the bench line to a side leg
        extra["cfg5_pool"] = {"failed": repr(exc)}
    return extra


def pool_extra(torch, lwc_b200, dist, rank, world, local_rank, steps, warmup, n_own=32768, D=1024, M=16):
    import ctypes

    from lwc_b200 import shard

    n_total = n_own * world
    ctx = lwc_b200.Context(device=local_rank)
    if world > 1:
        ident = lwc_b200.comm_unique_id() if rank == 0 else bytes(128)
        ident = shard.broadcast_bytes(dist, ident, src=0, device="cuda")
        ctx.comm_init(world, rank, ident)
    # host-generated rows (so that rank 0 can check the whole pool against the CPU oracle's O(N D) identity)
    rng = np.random.default_rng(100 + rank)
    cent = np.random.default_rng(99).standard_normal((8, D)).astype(np.float32)
    E_h = cent[rng.integers(0, 8, n_own)] * 0.5 + rng.standard_normal((n_own, D), dtype=np.float32)
    V_h = rng.random((M, n_own), dtype=np.float32) ** 6
    V_h /= V_h.sum(axis=1, keepdims=True) * np.float32(world)  # a voter's whole row (all ranks) sums to 1
    w_h = np.linspace(0.5, 2.0, M).astype(np.float32)
    E, V, w = torch.from_numpy(E_h).cuda(), torch.from_numpy(V_h).cuda(), torch.from_numpy(w_h).cuda()
    conf = torch.empty((n_own,), dtype=torch.float32, device="cuda")
    cw = torch.empty((n_own,), dtype=torch.float32, device="cuda")
    agree = torch.empty((M,), dtype=torch.float32, device="cuda")
    winner = torch.zeros((1,), dtype=torch.int64, device="cuda")
    desc = lwc_b200.PoolDesc(n_own, D, M, lwc_b200.PREC_BF16, lwc_b200.MEM_DEVICE)
    vp = ctypes.c_void_p

    def step():
        rc = ctx.lib.lwc_pool_consensus(ctx.handle, ctypes.byref(desc), vp(E.data_ptr()), vp(V.data_ptr()), vp(w.data_ptr()),
                                        vp(conf.data_ptr()), vp(cw.data_ptr()), vp(agree.data_ptr()), vp(winner.data_ptr()), None)
        if rc:
            raise lwc_b200.LwcError(rc, ctx.lib.lwc_last_error_message(ctx.handle).decode())

    def timed():
        for _ in range(warmup):
            step()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        ext = torch.cuda.ExternalStream(ctx.stream())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        for _ in range(steps):
            step()  # blocking: returns when the call's stream has drained
        e1.record(ext)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        return shard.max_over_ranks(dist, e0.elapsed_time(e1) / steps, device="cuda")

    def gather_result():
        """(whole-pool confidences on rank 0, winners of all ranks, sum of confidences)"""
        if dist is None:
            return conf.cpu().numpy(), [int(winner.item())], float(conf.sum().item())
        parts = [torch.empty_like(conf) for _ in range(world)]
        dist.all_gather(parts, conf)
        wins = [torch.zeros_like(winner) for _ in range(world)]
        dist.all_gather(wins, winner)
        allc = torch.cat(parts)
        return allc.cpu().numpy(), [int(x.item()) for x in wins], float(allc.double().sum().item())

    try:
        bf16_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
    except Exception:
        bf16_peak = 1590.0
    flops_per_gpu = 2.0 * n_total * n_total * D / world
    routes = [("nccl_allgather" if world > 1 else "single_gpu", False)] + ([("fused_p2p_gather", True)] if world > 1 else [])
    want = None
    if world > 1:  # rank 0 needs the whole pool for the check
        Es = [torch.empty_like(E) for _ in range(world)]
        Vs = [torch.empty_like(V) for _ in range(world)]
        dist.all_gather(Es, E)
        dist.all_gather(Vs, V)
        if rank == 0:
            E_all = torch.cat(Es).cpu().numpy()
            V_all = torch.cat(Vs, dim=1).cpu().numpy()
        del Es, Vs
    else:
        E_all, V_all = E_h, V_h
    if rank == 0:
        import oracle  # checker only

        want = oracle.pool_f64(E_all, V_all, w_h, method=1)  # O(N D) centroid identity (SURVEY 8.0 note)
    out = {"rows_total": n_total, "rows_per_gpu": n_own, "D": D, "voters": M, "routes": []}
    for name, fused in routes:
        if fused:
            mine = ctx.pool_p2p_export(n_own, D)
            buf = torch.frombuffer(bytearray(mine), dtype=torch.uint8).cuda()
            bufs = [torch.empty_like(buf) for _ in range(world)]
            dist.all_gather(bufs, buf)
            ctx.pool_p2p_attach(b"".join(bytes(x.cpu().numpy().tobytes()) for x in bufs))
        ms = timed()
        allc, wins, total = gather_result()
        rec = {"route": name, "ms": ms, "tflops_per_gpu": flops_per_gpu / (ms * 1e-3) / 1e12,
               "frac_of_bf16_peak": flops_per_gpu / (ms * 1e-3) / 1e12 / bf16_peak, "winner_same": len(set(wins)) == 1,
               "sum_conf": total, "steps": steps}
        if rank == 0:
            top = float(want["conf"].max())
            rec["max_conf_err_over_top"] = float(np.max(np.abs(allc - want["conf"])) / top)
            rec["winner_is_oracles"] = bool(wins[0] == want["winner"])
            rec["check_passed"] = bool(rec["max_conf_err_over_top"] <= 1e-2 and rec["winner_same"] and abs(total - 1.0) < 1e-3)
        out["routes"].append(rec)
    if world > 1:
        ctx.comm_destroy()
    ctx.close()
    return out


def run_other_workload(args, torch, lwc_b200, synth, local_rank, rank=0, world=1, dist=None):
    """Kernel-only measurement of another BASELINE.json shape (cfg3/cfg4), inputs generated in HBM.  Under
    torchrun the R requests of the config are sharded by request over the ranks (strong scaling, no
    collective; BASELINE config 3 is quoted that way).  Not the driver's bench line: it feeds profiles/."""
    from lwc_b200 import shard

    if args.workload == "pool":
        return run_pool_workload(args, torch, lwc_b200, local_rank)
    if args.workload == "table":
        return run_table_workload(args, torch, lwc_b200, local_rank)
    if args.workload == "vote":
        return run_vote_workload(args, torch, lwc_b200, local_rank)
    R_total, N, D, M = synth.CONFIGS[args.workload]
    if args.requests:
        R_total = args.requests
    b, e = shard.request_block(R_total, rank, world)
    R = e - b
    prec = {"fp32": lwc_b200.PREC_FP32, "tf32": lwc_b200.PREC_TF32}[args.precision]
    ctx = lwc_b200.Context(device=local_rank)
    n_rot = 2
    ms, kernel, par = timed_side_shape(torch, lwc_b200, synth, ctx, dist, R, N, D, M, prec, args.steps, args.warmup, n_rot=n_rot)
    peak, peak_src = load_peaks()
    nbytes = R * synth.algorithmic_bytes_per_request(N, D, M)  # per GPU: the roofline is a per-GPU statement
    flops = R * synth.algorithmic_flops_per_request(N, D, M)
    achieved = nbytes / (ms * 1e-3) / 1e9
    extra = {}
    if args.e2e and world == 1:
        extra = side_e2e_and_cpu(args, torch, lwc_b200, synth, ctx, N, D, M, prec)
    if rank == 0:
        print(json.dumps({**extra,
            "metric": f"scored requests/sec ({N}×{D} emb)", "value": R_total / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "scaling": "strong", "dtype": "f32",
            "data": "synthetic (generated in HBM)",
            "config": {"workload": f"{args.workload}: {R_total} requests x {N} candidates x {D}-dim fp32, {M} voters"
                                   + (f", sharded by request x{world} ({R} per GPU), no collective" if world > 1 else ""),
                       "kernel": kernel, "precision": args.precision,
                       "l2": f"{n_rot} rotating batches of {R * N * D * 4 / 1e9:.2f} GB per GPU"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": peak_src, "bytes_per_launch": nbytes, "scope": "per GPU",
                         "algorithmic_tflops": flops / (ms * 1e-3) / 1e12},
            "parity_sample": par,
            "gpu_launches": args.steps}), flush=True)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def side_e2e_and_cpu(args, torch, lwc_b200, synth, ctx, N, D, M, prec):
    """--e2e on a side workload: the same shape through lwc_score_batch(LWC_MEM_HOST) with pinned host buffers
    (copies inside the timed region), and the CPU port on a bounded sample, so that the end-to-end ratio can be
    read at pool sizes where the Gram dominates the CPU cost."""
    Rh = 2048 if N * D <= 32 * 3072 else 1024
    E, V, w, _ = synth.make_batch(Rh, N, D, M, seed=5, edge_cases=False)
    hE, hV, hw = (lwc_b200.HostBuffer(ctx, a.shape, np.float32) for a in (E, V, w))
    hE.array[...], hV.array[...], hw.array[...] = E, V, w
    outs = [lwc_b200.HostBuffer(ctx, (Rh, N), np.float32), lwc_b200.HostBuffer(ctx, (Rh, N), np.float32),
            lwc_b200.HostBuffer(ctx, (Rh, M), np.float32), lwc_b200.HostBuffer(ctx, (Rh,), np.int32),
            lwc_b200.HostBuffer(ctx, (Rh, N), np.int32)]
    desc = lwc_b200.BatchDesc(Rh, N, D, M, lwc_b200.SIM_COSINE, prec, lwc_b200.MEM_HOST, 1)

    def step():
        ctx.score_raw(desc, hE.ptr, hV.ptr, hw.ptr, *[o.ptr for o in outs])

    for _ in range(2):
        step()
    ext = torch.cuda.ExternalStream(ctx.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(5):
        step()
    e1.record(ext)
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1) / 5
    out = {"e2e": {"value": Rh / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms, "requests_per_step": Rh,
                   "h2d_bytes_per_step": hE.nbytes + hV.nbytes + hw.nbytes, "d2h_bytes_per_step": sum(o.nbytes for o in outs),
                   "api": "lwc_score_batch(LWC_MEM_HOST), pinned host buffers"}}
    if not args.no_cpu_baseline:
        import oracle

        cores = host_cores()
        n = min(Rh, 256)
        t, reps = 0.0, 0
        while t < 5.0 and reps < 64:
            t += oracle.score_port(E[:n], V[:n], w[:n], oracle.SIM_COSINE, nthreads=cores)["seconds"]
            reps += 1
        out["cpu_baseline"] = {"value": n * reps / t, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{reps} pass(es) over {n} requests of this shape"}
    for b in (hE, hV, hw, *outs):
        b.free()
    return out


def run_pool_workload(args, torch, lwc_b200, local_rank):
    """Kernel-side measurement of the single-pool path (BASELINE config 5 shape, D = 1024, 16 voters) on one
    GPU: N rows x N columns all-pairs pass.  Side measurement for profiles/ and DESIGN.md."""
    import ctypes

    N, D, M = (args.requests or 65536), 1024, 16
    ctx = lwc_b200.Context(device=local_rank)
    g = torch.Generator(device="cuda").manual_seed(5)
    E = torch.randn((N, D), dtype=torch.float32, device="cuda", generator=g)
    V = torch.rand((M, N), dtype=torch.float32, device="cuda", generator=g) ** 6
    V /= V.sum(dim=1, keepdim=True)
    w = torch.rand((M,), dtype=torch.float32, device="cuda", generator=g) + 0.5
    conf = torch.empty((N,), dtype=torch.float32, device="cuda")
    cw = torch.empty((N,), dtype=torch.float32, device="cuda")
    agree = torch.empty((M,), dtype=torch.float32, device="cuda")
    winner = torch.zeros((1,), dtype=torch.int64, device="cuda")
    desc = lwc_b200.PoolDesc(N, D, M, lwc_b200.PREC_BF16, lwc_b200.MEM_DEVICE)
    vp = ctypes.c_void_p

    def step():
        rc = ctx.lib.lwc_pool_consensus(ctx.handle, ctypes.byref(desc), vp(E.data_ptr()), vp(V.data_ptr()), vp(w.data_ptr()),
                                        vp(conf.data_ptr()), vp(cw.data_ptr()), vp(agree.data_ptr()), vp(winner.data_ptr()), None)
        if rc:
            raise lwc_b200.LwcError(rc, ctx.lib.lwc_last_error_message(ctx.handle).decode())

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    ext = torch.cuda.ExternalStream(ctx.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(args.steps):
        step()  # blocking: returns after the call's own stream drained
    e1.record(ext)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    flops = 2.0 * N * N * D
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
        src = "measured (MEASURED_PEAKS.json bf16_tflops, burst)"
    except Exception:
        peak, src = 1590.0, "fallback (B200_PROFILING.md)"
    ach = flops / (ms * 1e-3) / 1e12
    print(json.dumps({
        "metric": "pool_all_pairs_per_sec", "value": 1e3 / ms, "unit": "pools/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "dtype": "bf16 operands, f32 accumulate", "data": "synthetic (generated in HBM)",
        "config": {"workload": f"pool: {N} candidates x {D}-dim, {M} voters, all pairs on one GPU", "kernel": "pool_gemm (tcgen05 kind::f16)"},
        "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "peak_source": src,
                     "flops_per_launch": flops},
        "winner": int(winner.item())}), flush=True)
    ctx.close()
    return 0


def run_vote_workload(args, torch, lwc_b200, local_rank):
    """Side measurement of vote production: rows of 20 top-logprob alternatives -> vote rows of 16
    choices, written where the scorer reads them.  HBM-bound: 8 bytes in per alternative, 4 out per choice."""
    import ctypes

    rows, top, N = (args.requests or 4_194_304), 20, 16
    ctx = lwc_b200.Context(device=local_rank)
    g = torch.Generator(device="cuda").manual_seed(3)
    lp = -8.0 * torch.rand((rows, top), dtype=torch.float32, device="cuda", generator=g)
    leaf = torch.randint(-4, N, (rows, top), dtype=torch.int32, device="cuda", generator=g)
    V = torch.empty((rows, N), dtype=torch.float32, device="cuda")
    vp = ctypes.c_void_p
    st = torch.cuda.ExternalStream(ctx.stream())

    def step():
        rc = ctx.lib.lwc_vote_from_logprobs(ctx.handle, rows, top, N, vp(lp.data_ptr()), vp(leaf.data_ptr()), vp(V.data_ptr()),
                                            lwc_b200.MEM_DEVICE, None)
        if rc:
            raise lwc_b200.LwcError(rc, ctx.lib.lwc_last_error_message(ctx.handle).decode())

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(args.steps):
        step()
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    nbytes = rows * (top * 8 + N * 4)
    peak, src = load_peaks()
    ach = nbytes / (ms * 1e-3) / 1e9
    print(json.dumps({
        "metric": "vote_rows_per_sec", "value": rows * 1e3 / ms, "unit": "rows/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "dtype": "f32", "data": "synthetic (generated in HBM)",
        "config": {"workload": f"vote: {rows} voter rows x {top} alternatives -> {N} choices", "kernel": "vote_from_logprobs"},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "peak_source": src,
                     "bytes_per_launch": nbytes},
        "row_sum": float(V[0].sum().item())}), flush=True)
    ctx.close()
    return 0


def run_table_workload(args, torch, lwc_b200, local_rank):
    """Side measurement of the training-table weight lookup: 8 prompt embeddings per call
    against a resident table; the score pass streams the whole table from HBM once per call."""
    import time
    import numpy as np

    rows, D, M, top, nq = (args.requests or 1_000_000), 1536, 16, 16, 8
    ctx = lwc_b200.Context(device=local_rank)
    rng = np.random.default_rng(8)
    T = rng.standard_normal((rows, D), dtype=np.float32)
    Q = rng.random((rows, M), dtype=np.float32)
    q = rng.standard_normal((nq, D), dtype=np.float32)
    base = np.ones(M, np.float32)
    t0 = time.perf_counter()
    ctx.table_load(T, Q)
    load_s = time.perf_counter() - t0
    for _ in range(args.warmup):
        ctx.table_weights(q, top, base, base * 0.5, base * 2)
    ext = torch.cuda.ExternalStream(ctx.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(args.steps):
        w, idx, sim = ctx.table_weights(q, top, base, base * 0.5, base * 2)
    e1.record(ext)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    bytes_alg = rows * (D * 4 + 4) + nq * rows * 4 * 2  # table + inverse norms once; scores written then scanned
    peak, src = load_peaks()
    ach = bytes_alg / (ms * 1e-3) / 1e9
    print(json.dumps({
        "metric": "table_weight_lookups_per_sec", "value": nq * 1e3 / ms, "unit": "prompts/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"table: {rows} rows x {D}-dim, {M} LLMs, top {top}, {nq} prompts per call (host in/out)",
                   "kernel": "table_scores + table_topk + table_weights", "table_load_s": load_s},
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "peak_source": src,
                     "bytes_per_call": bytes_alg, "note": "whole call incl. top-k rounds and host copies"},
        "first_neighbour": int(idx[0, 0])}), flush=True)
    ctx.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e", action="store_true", help="side workloads cfg3/cfg4: add the host-buffer path and the CPU port")
    ap.add_argument("--launch", default="list", choices=["list", "loop"], help="timed launches: one lwc_score_batches call (default) or a Python loop of lwc_score_batch calls")
    ap.add_argument("--no-extras", action="store_true", help="skip the cfg3 / cfg4 / cfg5 side records of the main line")
    ap.add_argument("--no-cfg1", action="store_true", help="skip the single-request latency leg (2,100 tiny launches; used for the ncu launch list)")
    ap.add_argument("--no-host-leg", action="store_true", help="skip tools/e2e_host (end to end from the reference's types)")
    ap.add_argument("--no-overlap", action="store_true", help="do not set LWC_CTX_INPUTS_RESIDENT (launches fully serialised)")
    ap.add_argument("--workload", default=WORKLOAD, choices=["cfg2", "cfg3", "cfg4", "cfg2x4", "pool", "table", "vote"],
                    help="cfg2 is the bench line; cfg3/cfg4 are kernel-only side measurements")
    ap.add_argument("--requests", type=int, default=0, help="override R for --workload cfg3/cfg4")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
