set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest17.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest17.log; tail -25 gpurun_out/pytest17.log
B="timeout 200 python bench.py"
$B --workload cfg4 --requests 2048 --steps 10 --warmup 3 --precision fp32 > gpurun_out/cfg4_x3.log 2>&1; tail -1 gpurun_out/cfg4_x3.log
LWC_TC_KO=1 $B --workload cfg4 --requests 2048 --steps 10 --warmup 3 --precision fp32 > gpurun_out/cfg4_x3_ko1.log 2>&1; tail -1 gpurun_out/cfg4_x3_ko1.log
$B --workload cfg4 --requests 2048 --steps 10 --warmup 3 --precision tf32 > gpurun_out/cfg4_tf32b.log 2>&1; tail -1 gpurun_out/cfg4_tf32b.log
