set -x
CMD="python bench.py --workload cfg4 --requests 1024 --steps 3 --warmup 3 --precision tf32"
timeout 200 $CMD > gpurun_out/plain_tc.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:score_tc -s 3 -c 1 -o gpurun_out/r1_score_tc $CMD > gpurun_out/ncu_tc.log 2>&1
CMD2="python bench.py --workload pool --requests 16384 --steps 2 --warmup 3"
timeout 200 $CMD2 > gpurun_out/plain_pool.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:pool_gemm -s 3 -c 1 -o gpurun_out/r1_pool_gemm $CMD2 > gpurun_out/ncu_pool.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_pool_launches.csv $CMD2 > gpurun_out/ncu_pool_l.log 2>&1
ls -la gpurun_out/*.ncu-rep
