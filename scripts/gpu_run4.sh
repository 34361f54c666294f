set -x
B="timeout 150 python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s4_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s4_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s4_$tag.log | head -1) $(grep -o '"host_device_paths_agree": [a-z]*' gpurun_out/s4_$tag.log)"; }
run stream_8_144 LWC_STREAM_NCW=8 LWC_STREAM_KC=144
run stream_12_240 LWC_STREAM_NCW=12 LWC_STREAM_KC=240
run stream_4_144 LWC_STREAM_NCW=4 LWC_STREAM_KC=144
timeout 150 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain4.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:score_stream -s 3 -c 1 -o gpurun_out/prof4 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full4.log 2>&1
