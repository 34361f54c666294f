#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 2>&1 | tail -30 > gpurun_out/gpu_all_5.log; tail -6 gpurun_out/gpu_all_5.log | cut -c1-300
python bench.py --workload cfg4 --steps 10 --warmup 3 | tee gpurun_out/r2_cfg4_f16_pf.json | cut -c1-260
LWC_TC_STAGES=4 python bench.py --workload cfg4 --steps 10 --warmup 3 | cut -c1-260
python bench.py --workload cfg4 --precision tf32 --steps 10 --warmup 3 | cut -c1-260
for kw in 2 4; do for u in 2 4; do echo "KW=$kw U=$u"; LWC_SMALL_KW=$kw LWC_SMALL_U=$u python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | cut -c1-240; done; done
LWC_SMALL_IMPL=stream python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | cut -c1-240
python bench.py --workload cfg3 --requests 16384 --precision tf32 --steps 20 --warmup 3 | cut -c1-240
