#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 2>&1 | tail -40 > gpurun_out/gpu_all_7.log; tail -12 gpurun_out/gpu_all_7.log | cut -c1-300
for i in 1 2; do
echo "cfg3 default (ring, fp16 pieces + fix-up)"; python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | tee gpurun_out/r2_cfg3_ring16.json | cut -c1-230
echo "cfg3 direct (fp16 pieces)"; LWC_SMALL_IMPL=direct python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | cut -c1-230
done
echo "cfg3 tf32 single pass"; python bench.py --workload cfg3 --requests 16384 --precision tf32 --steps 20 --warmup 3 | cut -c1-230
echo "cfg3 full"; python bench.py --workload cfg3 --steps 10 --warmup 3 | tee gpurun_out/r2_cfg3_full.json | cut -c1-1200
