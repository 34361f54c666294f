#!/bin/bash
# Round-2 single-GPU record: tests, the bench line, the ncu launch list of the bench command, full captures
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 2>&1 | tail -30 > gpurun_out/gpu_all_8.log; tail -5 gpurun_out/gpu_all_8.log | cut -c1-300
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_d.json 2> gpurun_out/r2_bench_d.err; tail -c 300 gpurun_out/r2_bench_d.err
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r2_bench_ref.json 2>/dev/null
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-host-leg > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-host-leg > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -f -k regex:score_stream -c 2 -o gpurun_out/r2_stream python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --no-host-leg > gpurun_out/ncu_stream.log 2>&1
ncu --set full --clock-control none --import-source on -f -k regex:vote_ring -c 1 -o gpurun_out/r2_vote_ring python bench.py --workload vote --steps 1 --warmup 3 > gpurun_out/ncu_vote_ring.log 2>&1
for f in r2_stream r2_vote_ring; do ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/${f}_raw.csv 2>/dev/null; done
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2_bench_d.json") if l.startswith("{")][-1])
print("value", d["value"], d["ms_per_step"], d["roofline"]["frac"], "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"])
for k,v in d["extra"].items():
    if k!="cfg5_pool": print(k, v["ms"], v["frac"], v["parity_sample"]["tolerance_used"])
print(json.dumps(d["extra"]["cfg5_pool"])[:400])
r=json.loads(open("gpurun_out/r2_bench_ref.json").read().strip().splitlines()[-1]); print("ref", r["value"], r["step_ms"], r["config"]["workload"]==d["config"]["workload"])
PY
head -3 gpurun_out/r2_launches.csv | cut -c1-300; wc -l gpurun_out/r2_launches.csv
