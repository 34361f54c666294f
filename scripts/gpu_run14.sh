set -x
nvidia-smi --query-gpu=index,name --format=csv
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest14.log 2>&1; echo "rc=$?" >> gpurun_out/pytest14.log; tail -30 gpurun_out/pytest14.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29620 bench.py --gpus 2 --steps 200 --warmup 10 > gpurun_out/bench_2gpu.log 2>&1; tail -2 gpurun_out/bench_2gpu.log | cut -c1-900
