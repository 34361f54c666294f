#!/bin/bash
# N-GPU pass (N = number of visible GPUs): the multi-GPU tests, then the bench exactly as the driver launches it
N=$(python -c "import torch; print(torch.cuda.device_count())")
mkdir -p gpurun_out
echo "GPUs: $N"
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu --timeout 800 2>&1 | tail -15 | cut -c1-400
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err
tail -c 1500 gpurun_out/r2_bench_${N}gpu.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2_bench_${N}gpu.json") if l.startswith("{")][-1])
print("value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "per_rank", d["per_rank_ms_per_step"], "spread", d["per_rank_spread"])
print("e2e", d["e2e"]["value"], d["e2e"]["h2d_peak_gbs"], d["e2e"]["frac_of_h2d_peak"])
print(json.dumps(d["extra"], indent=0)[:3000])
PY
