set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest2.log; tail -15 gpurun_out/pytest2.log
B="python bench.py --steps 200 --warmup 10 --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/s2_$tag.log 2>&1; echo "$tag: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s2_$tag.log | head -1) $(grep -o '"frac": [0-9.]*' gpurun_out/s2_$tag.log | head -1) $(tail -1 gpurun_out/s2_$tag.log | cut -c1-100)"; }
run direct_4_4 LWC_SMALL_IMPL=direct LWC_SMALL_KW=4 LWC_SMALL_U=4
run direct_2_2 LWC_SMALL_IMPL=direct LWC_SMALL_KW=2 LWC_SMALL_U=2
run stream_8_128 LWC_STREAM_NCW=8 LWC_STREAM_KC=128
run stream_8_256 LWC_STREAM_NCW=8 LWC_STREAM_KC=256
run stream_4_128 LWC_STREAM_NCW=4 LWC_STREAM_KC=128
run stream_4_256 LWC_STREAM_NCW=4 LWC_STREAM_KC=256
run stream_12_128 LWC_STREAM_NCW=12 LWC_STREAM_KC=128
run stream_12_256 LWC_STREAM_NCW=12 LWC_STREAM_KC=256
run stream_8_128_s8 LWC_STREAM_NCW=8 LWC_STREAM_KC=128 LWC_STREAM_SLOTS=8
run stream_8_128_s4 LWC_STREAM_NCW=8 LWC_STREAM_KC=128 LWC_STREAM_SLOTS=4
