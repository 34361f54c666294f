set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest7.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest7.log; tail -15 gpurun_out/pytest7.log
B="timeout 150 python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s7_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s7_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s7_$tag.log | head -1) $(grep -o '"host_device_paths_agree": [a-z]*' gpurun_out/s7_$tag.log)"; }
run stream_8 LWC_STREAM_NCW=8
run stream_4 LWC_STREAM_NCW=4
run stream_12 LWC_STREAM_NCW=12
run stream_8_kc512 LWC_STREAM_NCW=8 LWC_STREAM_KC=512
run stream_8_kc1536 LWC_STREAM_NCW=8 LWC_STREAM_KC=1536
run stream_4_kc1536 LWC_STREAM_NCW=4 LWC_STREAM_KC=1536
run stream_12_kc384 LWC_STREAM_NCW=12 LWC_STREAM_KC=384
timeout 150 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain7.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:score_stream -s 3 -c 1 -o gpurun_out/prof7 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full7.log 2>&1
