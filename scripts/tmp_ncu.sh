python scripts/r2_pool_sizes.py 16 4 384 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -f -k regex:score_stream_kernel -c 1 -o gpurun_out/r2_stream_d384 python scripts/r2_pool_sizes.py 16 4 384 > gpurun_out/ncu_stream_d384.log 2>&1
ncu -i gpurun_out/r2_stream_d384.ncu-rep --page raw --csv > gpurun_out/r2_stream_d384_raw.csv 2>/dev/null
ncu -i gpurun_out/r2_stream_d384.ncu-rep --page source --csv > gpurun_out/r2_stream_d384_source.csv 2>/dev/null
rm -f gpurun_out/r2_stream_d384.ncu-rep
python - <<PY
import csv
rows = list(csv.reader(open("gpurun_out/r2_stream_d384_raw.csv"))); h = rows[0]; x = rows[2]
for i,k in enumerate(h):
    if any(t in k for t in ["gpu__time_duration.sum","dram__bytes_read.sum","issue_active.avg.pct","stalled","pipe_tensor","lsu_wavefronts_mem_shared.sum.pct","warps_active.avg.pct","inst_executed.sum "]):
        print(k, x[i], rows[1][i])
PY
