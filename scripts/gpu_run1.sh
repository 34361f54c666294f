set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python -m pytest tests -m gpu -x -q > gpurun_out/pytest1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest1.log; tail -30 gpurun_out/pytest1.log
python bench.py --steps 100 --warmup 5 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"; tail -5 gpurun_out/bench1.log
for kw in 2 4 8; do for u in 2 4; do
  LWC_SMALL_KW=$kw LWC_SMALL_U=$u python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/sweep_${kw}_${u}.log 2>&1
  echo "kw=$kw u=$u: $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/sweep_${kw}_${u}.log | head -1)"
done; done
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches1.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:score_small -s 3 -c 2 -o gpurun_out/prof1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
