set -x
W=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29641 tests/mgpu_pool_check.py > gpurun_out/mgpu_check_$W.log 2>&1; echo "check rc=$?"; grep -E "sharded pool|MGPU_POOL_OK|Error|error" gpurun_out/mgpu_check_$W.log | head
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29642 tests/mgpu_pool_bench.py > gpurun_out/cfg5_$W.log 2>&1; echo "bench rc=$?"; grep "^{" gpurun_out/cfg5_$W.log
for n in 2 4 8; do if [ $n -le $W ]; then
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2965$n bench.py --gpus $n --steps 200 --warmup 10 > gpurun_out/scale_$n.log 2>&1; grep -o '"value": [0-9.]*' gpurun_out/scale_$n.log | head -1; fi; done
