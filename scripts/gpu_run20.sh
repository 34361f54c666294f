set -x
B="timeout 280 python bench.py"
$B --workload cfg3 --steps 5 --warmup 3 > gpurun_out/cfg3_full.log 2>&1; tail -1 gpurun_out/cfg3_full.log | cut -c1-700
$B --workload cfg3 --steps 5 --warmup 3 --precision tf32 > gpurun_out/cfg3_full_tf32.log 2>&1; tail -1 gpurun_out/cfg3_full_tf32.log | cut -c1-700
$B --workload cfg4 --steps 5 --warmup 3 --precision tf32 > gpurun_out/cfg4_full_tf32.log 2>&1; tail -1 gpurun_out/cfg4_full_tf32.log | cut -c1-700
$B --workload cfg4 --steps 5 --warmup 3 > gpurun_out/cfg4_full_fp32.log 2>&1; tail -1 gpurun_out/cfg4_full_fp32.log | cut -c1-700
for kw in 2 4 8; do LWC_SMALL_IMPL=direct LWC_SMALL_KW=$kw LWC_SMALL_U=2 $B --workload cfg3 --requests 16384 --steps 10 --warmup 3 > gpurun_out/cfg3_d$kw.log 2>&1; echo "direct kw=$kw: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/cfg3_d$kw.log)"; done
for nf in 1 2 4; do LWC_SMALL_IMPL=stream LWC_STREAM_NCW=8 LWC_STREAM_NF=$nf $B --workload cfg3 --requests 16384 --steps 10 --warmup 3 > gpurun_out/cfg3_s$nf.log 2>&1; echo "stream nf=$nf: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/cfg3_s$nf.log)"; done
