#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 2>&1 | tail -30 > gpurun_out/gpu_all_4.log; tail -8 gpurun_out/gpu_all_4.log | cut -c1-300
python bench.py --workload vote --steps 20 --warmup 3 | tee gpurun_out/r2_vote3.json | cut -c1-330
for i in 1 2; do
python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | cut -c1-240
LWC_B200_LIB=$PWD/build/liblwc_noflush.so python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | cut -c1-240
done
bash scripts/r2_ncu.sh
