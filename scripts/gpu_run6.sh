set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest6.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest6.log; tail -15 gpurun_out/pytest6.log
B="timeout 150 python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s6_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s6_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s6_$tag.log | head -1) $(grep -o '"host_device_paths_agree": [a-z]*' gpurun_out/s6_$tag.log)"; }
run stream_8 LWC_STREAM_NCW=8
run stream_4 LWC_STREAM_NCW=4
run stream_12 LWC_STREAM_NCW=12
run stream_8_bps2 LWC_STREAM_NCW=8 LWC_STREAM_BPS=2
run stream_8_kc208 LWC_STREAM_NCW=8 LWC_STREAM_KC=208
run stream_8_kc176 LWC_STREAM_NCW=8 LWC_STREAM_KC=176
run stream_8_s6 LWC_STREAM_NCW=8 LWC_STREAM_SLOTS=6
run stream_12_bps2 LWC_STREAM_NCW=12 LWC_STREAM_BPS=2
timeout 150 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain6.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:score_stream -s 3 -c 1 -o gpurun_out/prof6 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full6.log 2>&1
