set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 200 python bench.py --steps 300 --warmup 10 --no-cpu-baseline > gpurun_out/bench_pdl.log 2>&1; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_pdl.log | head -1; grep -o '"frac": [0-9.]*' gpurun_out/bench_pdl.log | head -1; grep -o '"host_device_paths_agree": [a-z]*' gpurun_out/bench_pdl.log
timeout 200 python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-overlap > gpurun_out/bench_nopdl.log 2>&1; grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_nopdl.log | head -1
