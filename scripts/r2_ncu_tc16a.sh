#!/bin/bash
# ncu --set full capture of the fp32-grade tcgen05 kernel with its A operands in TMEM (cfg4 slice of 2,048 requests); the
# command has exited 0 without ncu first
mkdir -p gpurun_out
python bench.py --workload cfg4 --requests 2048 --steps 1 --warmup 3 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -f -k regex:score_tc16a_kernel -c 1 -o gpurun_out/r2_tc16a python bench.py --workload cfg4 --requests 2048 --steps 1 --warmup 3 > gpurun_out/ncu_tc16a.log 2>&1
ncu -i gpurun_out/r2_tc16a.ncu-rep --page raw --csv > gpurun_out/r2_tc16a_raw.csv 2>/dev/null
rm -f gpurun_out/r2_tc16a.ncu-rep
python - <<PY
import csv
rows = list(csv.reader(open("gpurun_out/r2_tc16a_raw.csv"))); h = rows[0]; x = rows[2]
for k in ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum.pct_of_peak_sustained_elapsed",
          "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread"]:
    if k in h: print(k, x[h.index(k)], rows[1][h.index(k)])
PY
