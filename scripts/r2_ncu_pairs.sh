#!/bin/bash
# ncu --set full captures of the between-shapes forms: two 64-candidate requests per tcgen05 tile (fp32-grade and TF32),
# and the ring kernel's half-tile form at 8 candidates.  The command has exited 0 without ncu first.
mkdir -p gpurun_out
python scripts/r2_pool_sizes.py 8,64 > gpurun_out/pairs_plain.log 2>&1 || exit 1
NCU="ncu --set full --clock-control none --import-source on -f"
$NCU -k regex:score_tc16_kernel -c 1 -o gpurun_out/r2_tc16_pair python scripts/r2_pool_sizes.py 64 > gpurun_out/ncu_tc16_pair.log 2>&1
# the fp32 leg launches the fix-up form of score_tc_kernel 8 times (3 warm-up + 5 timed) before the TF32 leg's first launch
$NCU -k regex:score_tc_kernel --launch-skip 8 -c 1 -o gpurun_out/r2_tc_pair python scripts/r2_pool_sizes.py 64 > gpurun_out/ncu_tc_pair.log 2>&1
$NCU -k regex:score_stream_kernel -c 1 -o gpurun_out/r2_stream_half python scripts/r2_pool_sizes.py 8 > gpurun_out/ncu_stream_half.log 2>&1
for f in r2_tc16_pair r2_tc_pair r2_stream_half; do ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/${f}_raw.csv 2>/dev/null; rm -f gpurun_out/$f.ncu-rep; done  # reports with sources are ~25 MB each
python - <<PY
import csv
for f in ["r2_tc16_pair", "r2_tc_pair", "r2_stream_half"]:
    rows = list(csv.reader(open(f"gpurun_out/{f}_raw.csv"))); h = rows[0]; x = rows[2]
    print(f)
    for k in ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
              "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
              "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread"]:
        if k in h: print("  ", k, x[h.index(k)], rows[1][h.index(k)])
PY
