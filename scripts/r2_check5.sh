#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 2>&1 | tail -30 > gpurun_out/gpu_all_6.log; tail -6 gpurun_out/gpu_all_6.log | cut -c1-300
for i in 1 2; do
for l in list loop; do echo "launch=$l K=20"; python bench.py --steps 20 --warmup 5 --launch $l --no-extras --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['frac'], d['value'])"; done
done
python bench.py --steps 200 --warmup 10 --no-extras --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('K=200', d['ms_per_step'], d['roofline']['frac'])"
python bench.py --workload vote --steps 20 --warmup 3 | tee gpurun_out/r2_vote4.json | cut -c1-330
LWC_VOTE_IMPL=tile python bench.py --workload vote --steps 20 --warmup 3 | cut -c1-330
python bench.py --workload cfg4 --steps 10 --warmup 3 | tee gpurun_out/r2_cfg4_f16_b.json | cut -c1-260
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_c.json 2> gpurun_out/r2_bench_c.err; tail -c 400 gpurun_out/r2_bench_c.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2_bench_c.json") if l.startswith("{")][-1])
print("value", d["value"], d["ms_per_step"], d["roofline"]["frac"]); print("e2e", json.dumps(d["e2e"])[:900]); print("host", json.dumps(d.get("e2e_from_reference_types"))[:2500]); print("cpu", d.get("cpu_baseline",{}).get("value"))
PY
