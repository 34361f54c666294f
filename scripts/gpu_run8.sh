set -x
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest8.log; tail -5 gpurun_out/pytest8.log
B="timeout 150 python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s8_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s8_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s8_$tag.log | head -1) $(grep -o '"host_device_paths_agree": [a-z]*' gpurun_out/s8_$tag.log)"; }
run nf1_8 LWC_STREAM_NCW=8 LWC_STREAM_NF=1
run nf1_12 LWC_STREAM_NCW=12 LWC_STREAM_NF=1
run nf1_4 LWC_STREAM_NCW=4 LWC_STREAM_NF=1
run nf2_8 LWC_STREAM_NCW=8 LWC_STREAM_NF=2
run nf1_8_s3 LWC_STREAM_NCW=8 LWC_STREAM_NF=1 LWC_STREAM_SLOTS=3
run nf1_8_kc1536 LWC_STREAM_NCW=8 LWC_STREAM_NF=1 LWC_STREAM_KC=1536
