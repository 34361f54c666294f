# Round-1 evidence capture: plain run first, then the ncu launch list and one full capture of the top kernel.
set -x
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/plain_r1.log 2>&1 && \
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1_launches.csv $CMD > gpurun_out/ncu_launch_r1.log 2>&1
timeout 200 $CMD > gpurun_out/plain_r1b.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:score_stream -s 3 -c 2 -o gpurun_out/r1_score_stream $CMD > gpurun_out/ncu_full_r1.log 2>&1
tail -2 gpurun_out/plain_r1.log | cut -c1-300
