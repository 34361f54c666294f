set -x
B="timeout 300 python bench.py"
$B --workload pool --requests 32768 --steps 5 --warmup 2 > gpurun_out/pool_32k.log 2>&1; tail -1 gpurun_out/pool_32k.log
$B --workload pool --requests 65536 --steps 3 --warmup 1 > gpurun_out/pool_64k.log 2>&1; tail -1 gpurun_out/pool_64k.log
