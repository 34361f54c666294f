set -x
B="timeout 150 python bench.py --steps 300 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s18_$tag.log 2>&1; echo "$tag rc=$?: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s18_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s18_$tag.log | head -1)"; }
run default
run n12_f4 LWC_STREAM_NCW=12 LWC_STREAM_NF=4
run n12_f2 LWC_STREAM_NCW=12 LWC_STREAM_NF=2
run n8_f6 LWC_STREAM_NCW=8 LWC_STREAM_NF=6
run n8_f3 LWC_STREAM_NCW=8 LWC_STREAM_NF=3
run n12_f4_kc1536 LWC_STREAM_NCW=12 LWC_STREAM_NF=4 LWC_STREAM_KC=1536
run n12_f4_kc512 LWC_STREAM_NCW=12 LWC_STREAM_NF=4 LWC_STREAM_KC=512
run n12_f4_kc384 LWC_STREAM_NCW=12 LWC_STREAM_NF=4 LWC_STREAM_KC=384
run n12_f4_s3 LWC_STREAM_NCW=12 LWC_STREAM_NF=4 LWC_STREAM_SLOTS=3
run default2
