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
        while not self.stop_flag:
            if not self.active:
                time.sleep(0.0005)
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


def run_reference(args):
    """--impl reference: the CPU port of the reference path, all host threads, same config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import oracle
    from lwc_b200 import synth

    R, N, D, M = synth.CONFIGS[WORKLOAD]
    sample = min(R, 1024)
    E, V, w, _ = synth.make_batch(sample, N, D, M)
    cores = host_cores()
    for _ in range(args.warmup):
        oracle.score_port(E, V, w, oracle.SIM_COSINE, nthreads=cores)
    secs = [oracle.score_port(E, V, w, oracle.SIM_COSINE, nthreads=cores)["seconds"] for _ in range(args.steps)]
    total = float(np.sum(secs))
    value = sample * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64+decimal96",
        "data": "synthetic",
        "config": {"workload": f"{WORKLOAD}: {R} requests x {N} candidates x {D}-dim, {M} voters",
                   "step": f"bounded sample: first {sample} requests of the batch per step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{args.steps} steps x {sample} requests; the Rust reference cannot be built "
                                   "here (no cargo/rustc), so this is the C restatement in oracle/"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


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

    n_rot = 3
    dE = [torch.empty((R, N, D), dtype=torch.float32, device="cuda") for _ in range(n_rot)]
    dE[0].copy_(torch.from_numpy(hE.array))
    for i in range(1, n_rot):  # distinct data per rotating batch: a request-wise roll of batch 0
        dE[i].copy_(torch.roll(dE[0], shifts=i * 17, dims=0))
    dV = torch.from_numpy(V).cuda()
    dw = torch.from_numpy(w).cuda()
    d_out = {"conf": torch.empty((R, N), dtype=torch.float32, device="cuda"),
             "cw": torch.empty((R, N), dtype=torch.float32, device="cuda"),
             "agree": torch.empty((R, M), dtype=torch.float32, device="cuda"),
             "winner": torch.empty((R,), dtype=torch.int32, device="cuda"),
             "rank": torch.empty((R, N), dtype=torch.int32, device="cuda")}
    desc_dev = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, lwc_b200.PREC_FP32, lwc_b200.MEM_DEVICE, 1)
    desc_host = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, lwc_b200.PREC_FP32, lwc_b200.MEM_HOST, 1)
    stream = torch.cuda.Stream()

    # the timed loop calls the C entry point with pre-built arguments: the launches have to stay ahead of a
    # 62 us kernel for consecutive launches to overlap, so nothing is re-wrapped per step
    import ctypes
    _vp = ctypes.c_void_p
    _fn = ctx.lib.lwc_score_batch
    _dev_args = [(ctx.handle, ctypes.byref(desc_dev), _vp(dE[j].data_ptr()), _vp(dV.data_ptr()), _vp(dw.data_ptr()),
                  _vp(d_out["conf"].data_ptr()), _vp(d_out["cw"].data_ptr()), _vp(d_out["agree"].data_ptr()),
                  _vp(d_out["winner"].data_ptr()), _vp(d_out["rank"].data_ptr()), _vp(stream.cuda_stream))
                 for j in range(n_rot)]

    def step_device(i):
        rc = _fn(*_dev_args[i % n_rot])
        if rc:
            raise lwc_b200.LwcError(rc, ctx.lib.lwc_last_error_message(ctx.handle).decode())

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
    for i in range(args.warmup):
        step_device(i)
    barrier()
    launches0 = ctx.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active = True
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for i in range(args.steps):
            step_device(i)
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

    # the two paths must agree bit for bit on the same batch (device batch 0 == host batch)
    step_device(0)
    torch.cuda.synchronize()
    same = bool(np.array_equal(d_out["winner"].cpu().numpy(), out_host["winner"].array))

    peak, peak_src = load_peaks()
    bytes_per_launch = R * synth.algorithmic_bytes_per_request(N, D, M)
    achieved = bytes_per_launch / (ms_per_step * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{WORKLOAD}: {R} requests x {N} candidates x {D}-dim fp32, {M} voters, per GPU",
                   "similarity": "cosine", "precision": "fp32-grade (3xTF32 compensated Gram)",
                   "kernel": kernel, "parallelism": f"request-sharded x{world}, no collectives",
                   "l2": f"{n_rot} rotating input batches of {hE.nbytes / 1e6:.0f} MB each (> 126 MB L2)",
                   "launches": "serialised" if args.no_overlap else
                               "programmatic dependent launch: a launch streams its (resident) inputs while the previous one drains"},
        "clocks": sampler.summary(), "per_rank_ms_per_step": per_rank_ms,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps, "wall_ms_per_step": wall_ms / e2e_steps,
                "api": "lwc_score_batch(LWC_MEM_HOST), pinned host buffers"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": load_traffic(kernel, R), "kernel": kernel, "peak_source": peak_src,
                     "bytes_per_launch": bytes_per_launch, "frac_of_nominal_8000": achieved / 8000.0},
        "host_device_paths_agree": same,
    }
    if world == 1 and not args.no_cpu_baseline:
        if all_cpus:
            os.sched_setaffinity(0, all_cpus)  # the CPU arm owns the whole box again
        line["cpu_baseline"] = cpu_baseline(hE.array, V, w)
    if rank == 0:
        print(json.dumps(line), flush=True)
    for b in (hE, hV, hw, *out_host.values()):
        b.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


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
    kernel = ctx.kernel_for(R, N, D, M, precision=prec)
    g = torch.Generator(device="cuda").manual_seed(7)
    n_rot = 2
    dE = []
    for _ in range(n_rot):
        e = torch.randn((R, N, D), dtype=torch.float32, device="cuda", generator=g)
        e *= torch.exp(0.5 * torch.randn((R, N, 1), dtype=torch.float32, device="cuda", generator=g))
        dE.append(e)
    V = synth.votes(R, N, M, seed=11)
    w = synth.weights(R, M, seed=11)
    dV, dw = torch.from_numpy(V).cuda(), torch.from_numpy(w).cuda()
    outs = [torch.empty((R, N), dtype=torch.float32, device="cuda"), torch.empty((R, N), dtype=torch.float32, device="cuda"),
            torch.empty((R, M), dtype=torch.float32, device="cuda"), torch.empty((R,), dtype=torch.int32, device="cuda"),
            torch.empty((R, N), dtype=torch.int32, device="cuda")]
    desc = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, prec, lwc_b200.MEM_DEVICE, 1)
    stream = torch.cuda.Stream()

    def step(i):
        ctx.score_raw(desc, dE[i % n_rot].data_ptr(), dV.data_ptr(), dw.data_ptr(), *[o.data_ptr() for o in outs],
                      stream=stream.cuda_stream)

    for i in range(args.warmup):
        step(i)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for i in range(args.steps):
            step(i)
        ev1.record(stream)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = shard.max_over_ranks(dist, ev0.elapsed_time(ev1) / args.steps, device="cuda")
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
                       "l2": f"{n_rot} rotating batches of {dE[0].numel() * 4 / 1e9:.2f} GB per GPU"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": peak_src, "bytes_per_launch": nbytes, "scope": "per GPU",
                         "algorithmic_tflops": flops / (ms * 1e-3) / 1e12},
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
