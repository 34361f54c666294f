#!/bin/bash
# Round-2 GPU pass on one B200: full -m gpu suite, the host-types end-to-end tool, vote workload, main bench line.
mkdir -p gpurun_out
nproc
timeout 1200 python -m pytest tests -q -m gpu --timeout 400 -x 2>&1 | tail -25 > gpurun_out/gpu_all_2.log; tail -12 gpurun_out/gpu_all_2.log
rm -f gpurun_out/r2_e2e_host.jsonl
for b in 256 512 1024 4096; do timeout 200 ./tools/e2e_host --batch $b | tee -a gpurun_out/r2_e2e_host.jsonl | cut -c1-330; done
timeout 200 ./tools/e2e_host --threads 8 | tee -a gpurun_out/r2_e2e_host.jsonl | cut -c1-200
for t in f64 bf16; do timeout 200 ./tools/e2e_host --transport $t | tee -a gpurun_out/r2_e2e_host.jsonl | cut -c1-330; done
python bench.py --workload vote --steps 20 --warmup 3 | tee gpurun_out/r2_vote.json | cut -c1-500
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_b.json 2> gpurun_out/r2_bench_b.err; tail -c 600 gpurun_out/r2_bench_b.err; cut -c1-4000 gpurun_out/r2_bench_b.json
