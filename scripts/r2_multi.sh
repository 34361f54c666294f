#!/bin/bash
# the bench exactly as the driver launches it, at N = all visible GPUs and at N = 1 on the same box
N=$(python -c "import torch; print(torch.cuda.device_count())")
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err
tail -c 600 gpurun_out/r2_bench_${N}gpu.err
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras --no-host-leg > gpurun_out/r2_bench_1of${N}.json 2>/dev/null
python - <<PY
import json
for f in ("gpurun_out/r2_bench_${N}gpu.json","gpurun_out/r2_bench_1of${N}.json"):
    d=json.loads([l for l in open(f) if l.startswith("{")][-1])
    print(f, "value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "per_rank", [round(x*1e3,1) for x in d["per_rank_ms_per_step"]], "spread", round(d["per_rank_spread"],4), "launches", d["gpu_launches"], "e2e", d["e2e"]["value"])
    if "extra" in d:
        e=d["extra"]; print(" cfg3", e["cfg3_strong"]["ms"], e["cfg3_strong"]["frac"], [ (r["route"], r["ms"], r.get("check_passed")) for r in e["cfg5_pool"]["routes"]])
PY
