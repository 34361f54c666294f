set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest3.log; tail -15 gpurun_out/pytest3.log
B="timeout 150 python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s3_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s3_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s3_$tag.log | head -1) $(tail -1 gpurun_out/s3_$tag.log | cut -c1-80)"; }
run stream_8_144 LWC_STREAM_NCW=8 LWC_STREAM_KC=144
run stream_8_240 LWC_STREAM_NCW=8 LWC_STREAM_KC=240
run stream_4_144 LWC_STREAM_NCW=4 LWC_STREAM_KC=144
run stream_12_144 LWC_STREAM_NCW=12 LWC_STREAM_KC=144
run stream_12_240 LWC_STREAM_NCW=12 LWC_STREAM_KC=240
run stream_8_144_s8 LWC_STREAM_NCW=8 LWC_STREAM_KC=144 LWC_STREAM_SLOTS=8
run stream_8_144_s12 LWC_STREAM_NCW=8 LWC_STREAM_KC=144 LWC_STREAM_SLOTS=12
