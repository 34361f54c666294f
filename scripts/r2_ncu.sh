#!/bin/bash
# ncu --set full captures of the kernels VERDICT r1 asked counters for (one GPU; each command has already exited 0 without ncu)
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
$NCU -k regex:score_tc16 -c 1 -o gpurun_out/r2_tc16 python bench.py --workload cfg4 --requests 2048 --steps 1 --warmup 3 > gpurun_out/ncu_tc16.log 2>&1
$NCU -k regex:score_small -c 1 -o gpurun_out/r2_small32 python bench.py --workload cfg3 --requests 16384 --steps 1 --warmup 3 > gpurun_out/ncu_small32.log 2>&1
$NCU -k regex:vote_tile -c 1 -o gpurun_out/r2_vote python bench.py --workload vote --steps 1 --warmup 3 > gpurun_out/ncu_vote.log 2>&1
for f in r2_tc16 r2_small32 r2_vote; do ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/${f}_raw.csv 2>/dev/null; done
ls -la gpurun_out/*.ncu-rep
