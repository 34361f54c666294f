set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest9.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest9.log; tail -5 gpurun_out/pytest9.log
B="timeout 150 python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s9_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s9_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s9_$tag.log | head -1) $(grep -o '"host_device_paths_agree": [a-z]*' gpurun_out/s9_$tag.log)"; }
run n8_f4 LWC_STREAM_NCW=8 LWC_STREAM_NF=4
run n8_f2 LWC_STREAM_NCW=8 LWC_STREAM_NF=2
run n8_f3 LWC_STREAM_NCW=8 LWC_STREAM_NF=3
run n8_f6 LWC_STREAM_NCW=8 LWC_STREAM_NF=6
run n12_f4 LWC_STREAM_NCW=12 LWC_STREAM_NF=4
run n4_f4 LWC_STREAM_NCW=4 LWC_STREAM_NF=4
run n8_f4_s3 LWC_STREAM_NCW=8 LWC_STREAM_NF=4 LWC_STREAM_SLOTS=3
