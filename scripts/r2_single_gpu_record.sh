#!/bin/bash
# Round-2 single-GPU record: tests, the bench line and the reference arm, the ncu launch list of the bench command,
# --set full captures of the headline (fused list), cfg3 and vote kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 2>&1 | tail -30 > gpurun_out/gpu_all_final.log; tail -5 gpurun_out/gpu_all_final.log | cut -c1-300
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 300 gpurun_out/r2_bench_n1.err
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r2_bench_ref.json 2>/dev/null
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-host-leg --no-cfg1 > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-host-leg --no-cfg1 > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -f -k regex:score_small_kernel -c 1 -o gpurun_out/r2_small32_pipe python bench.py --workload cfg3 --requests 16384 --steps 1 --warmup 3 > gpurun_out/ncu_small32_pipe.log 2>&1
ncu -i gpurun_out/r2_small32_pipe.ncu-rep --page raw --csv > gpurun_out/r2_small32_pipe_raw.csv 2>/dev/null
python - <<PY
import json, csv
d=json.loads([l for l in open("gpurun_out/r2_bench_n1.json") if l.startswith("{")][-1])
print("value", d["value"], d["ms_per_step"], d["roofline"]["frac"], "launches", d["gpu_launches"], "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"])
for k,v in d["extra"].items():
    if k!="cfg5_pool": print(k, v["ms"], v["frac"], v["parity_sample"]["tolerance_used"], v["kernel"])
print(json.dumps(d["extra"]["cfg5_pool"])[:300])
print({k: (v.get("value"), v.get("ratio_to_cpu_port")) for k,v in d["e2e_from_reference_types"].items()})
print("cfg1", d["gpu_cfg1"]["median_us"], d["cpu_baseline_cfg1"]["median_ns"])
r=json.loads(open("gpurun_out/r2_bench_ref.json").read().strip().splitlines()[-1]); print("ref", r["value"], r["step_ms"])
rows=list(csv.reader(open("gpurun_out/r2_small32_pipe_raw.csv"))); h=rows[0]; x=rows[2]
for k in ["Kernel Name","gpu__time_duration.sum","dram__bytes_read.sum","sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active","smsp__issue_active.avg.pct_of_peak_sustained_active","sm__warps_active.avg.pct_of_peak_sustained_active","launch__registers_per_thread","smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]:
    print(k, x[h.index(k)], rows[1][h.index(k)])
PY
grep -c "score_stream" gpurun_out/r2_launches.csv
