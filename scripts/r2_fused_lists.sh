#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do
for f in 1 0; do for k in 20 200; do
LWC_LIST_FUSION=$f python bench.py --steps $k --warmup 10 --no-extras --no-cpu-baseline --no-host-leg --no-cfg1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fusion=$f K=$k', round(d['ms_per_step']*1e3,2), round(d['roofline']['frac'],4), 'launches', d['gpu_launches'], 'agree', d['host_device_paths_agree'], d['config']['l2'][:40])"
done; done; done
ncu --set full --clock-control none --import-source on -f -k regex:score_stream_list -c 1 -o gpurun_out/r2_stream_list python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extras --no-host-leg --no-cfg1 > gpurun_out/ncu_stream_list.log 2>&1
ncu -i gpurun_out/r2_stream_list.ncu-rep --page raw --csv > gpurun_out/r2_stream_list_raw.csv 2>/dev/null
python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/r2_stream_list_raw.csv'))); hdr=rows[0]
for r in rows[2:]:
    print(r[hdr.index("Kernel Name")][:50], r[hdr.index("gpu__time_duration.sum")], r[hdr.index("dram__bytes_read.sum")], rows[1][hdr.index("dram__bytes_read.sum")], r[hdr.index("dram__bytes_write.sum")], rows[1][hdr.index("dram__bytes_write.sum")])
PY
