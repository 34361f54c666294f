set -x
B="timeout 200 python bench.py"
$B --steps 200 --warmup 10 > gpurun_out/bench10.log 2>&1; tail -1 gpurun_out/bench10.log | cut -c1-1500
$B --workload cfg3 --requests 16384 --steps 20 --warmup 3 > gpurun_out/cfg3_fp32.log 2>&1; tail -1 gpurun_out/cfg3_fp32.log
$B --workload cfg3 --requests 16384 --steps 20 --warmup 3 --precision tf32 > gpurun_out/cfg3_tf32.log 2>&1; tail -1 gpurun_out/cfg3_tf32.log
LWC_SMALL_IMPL=direct $B --workload cfg3 --requests 16384 --steps 20 --warmup 3 > gpurun_out/cfg3_direct.log 2>&1; tail -1 gpurun_out/cfg3_direct.log
LWC_STREAM_NCW=8 LWC_STREAM_NF=1 $B --workload cfg3 --requests 16384 --steps 20 --warmup 3 > gpurun_out/cfg3_n8f1.log 2>&1; tail -1 gpurun_out/cfg3_n8f1.log
LWC_STREAM_NCW=4 LWC_STREAM_NF=2 $B --workload cfg3 --requests 16384 --steps 20 --warmup 3 > gpurun_out/cfg3_n4f2.log 2>&1; tail -1 gpurun_out/cfg3_n4f2.log
$B --workload cfg4 --requests 2048 --steps 10 --warmup 3 > gpurun_out/cfg4_generic.log 2>&1; tail -1 gpurun_out/cfg4_generic.log
