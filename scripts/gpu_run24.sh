set -x
W=$(nvidia-smi -L | wc -l)
for n in 1 2 4 8; do if [ $n -le $W ]; then
if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2970$n"; fi
timeout 300 $L bench.py --gpus $n --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/scale2_$n.log 2>&1; echo "cfg2 n=$n: $(grep -o '"value": [0-9.]*' gpurun_out/scale2_$n.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/scale2_$n.log | head -1)"
timeout 300 $L bench.py --gpus $n --workload cfg3 --steps 10 --warmup 3 > gpurun_out/cfg3_scale_$n.log 2>&1; echo "cfg3 n=$n: $(grep -o '"value": [0-9.]*' gpurun_out/cfg3_scale_$n.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/cfg3_scale_$n.log | head -1)"
fi; done
timeout 200 tests/cpp/test_host --gpu | tail -2
