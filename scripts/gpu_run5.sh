set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest5.log; tail -15 gpurun_out/pytest5.log
B="timeout 150 python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s5_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s5_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s5_$tag.log | head -1) $(grep -o '"host_device_paths_agree": [a-z]*' gpurun_out/s5_$tag.log)"; }
run stream_8 LWC_STREAM_NCW=8
run stream_4 LWC_STREAM_NCW=4
run stream_12 LWC_STREAM_NCW=12
run stream_8_kc96 LWC_STREAM_NCW=8 LWC_STREAM_KC=96
run stream_8_kc64 LWC_STREAM_NCW=8 LWC_STREAM_KC=64
run stream_4_kc128 LWC_STREAM_NCW=4 LWC_STREAM_KC=128
run stream_8_s1 LWC_STREAM_NCW=8 LWC_STREAM_SLOTS=1
run direct_2_2 LWC_SMALL_IMPL=direct LWC_SMALL_KW=2 LWC_SMALL_U=2
