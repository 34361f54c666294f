set -x
W=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29661 tests/mgpu_pool_check.py > gpurun_out/mgpu_check_p2p_$W.log 2>&1; echo "check rc=$?"; grep -E "sharded pool|MGPU_POOL_OK|Error|error|disagree|failed" gpurun_out/mgpu_check_p2p_$W.log | head
LWC_POOL_N=$((32768*W)) timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29662 tests/mgpu_pool_bench.py > gpurun_out/pool_nccl_$W.log 2>&1; echo "nccl rc=$?"; grep "^{" gpurun_out/pool_nccl_$W.log | cut -c1-330
LWC_POOL_P2P=1 LWC_POOL_N=$((32768*W)) timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29663 tests/mgpu_pool_bench.py > gpurun_out/pool_p2p_$W.log 2>&1; echo "p2p rc=$?"; grep "^{" gpurun_out/pool_p2p_$W.log | cut -c1-330
