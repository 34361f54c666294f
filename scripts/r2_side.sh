#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --workload table --steps 10 --warmup 3 | tee gpurun_out/r2_table.json | cut -c1-700
python bench.py --workload pool --steps 5 --warmup 2 | tee gpurun_out/r2_pool_64k.json | cut -c1-500
python bench.py --workload cfg3 --precision tf32 --steps 5 --warmup 3 | tee gpurun_out/r2_cfg3_full_tf32.json | cut -c1-300
