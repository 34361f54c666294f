#!/bin/bash
# Round-2 GPU pass: tcgen05 fp16-piece route, vote kernel, host-types tool, cfg3 kernel timing
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 2>&1 | tail -40 > gpurun_out/gpu_all_3.log; tail -25 gpurun_out/gpu_all_3.log | cut -c1-300
rm -f gpurun_out/r2_e2e_host2.jsonl
for t in fp32 f64 bf16; do timeout 200 ./tools/e2e_host --transport $t | tee -a gpurun_out/r2_e2e_host2.jsonl | cut -c1-260; done
timeout 200 ./tools/e2e_host --threads 8 | tee -a gpurun_out/r2_e2e_host2.jsonl | cut -c1-200
python bench.py --workload vote --steps 20 --warmup 3 | tee gpurun_out/r2_vote2.json | cut -c1-330
python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | tee gpurun_out/r2_cfg3_slice2.json | cut -c1-300
python bench.py --workload cfg4 --steps 10 --warmup 3 | tee gpurun_out/r2_cfg4_f16.json | cut -c1-1200
LWC_TC_X3=tf32 python bench.py --workload cfg4 --steps 10 --warmup 3 | tee gpurun_out/r2_cfg4_tf32x3.json | cut -c1-300
python bench.py --workload cfg4 --precision tf32 --steps 10 --warmup 3 | tee gpurun_out/r2_cfg4_tf32.json | cut -c1-300
python bench.py --steps 200 --warmup 10 --no-extras --no-cpu-baseline | cut -c1-300
python bench.py --steps 20 --warmup 10 --no-extras --no-cpu-baseline | cut -c1-300
