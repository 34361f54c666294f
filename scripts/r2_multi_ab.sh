#!/bin/bash
N=$(python -c "import torch; print(torch.cuda.device_count())")
for d in 300 0 300 0; do
LWC_BENCH_SAMPLER_DELAY_US=$d timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus $N --steps 20 --warmup 5 --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('delay=$d', 'ms', round(d['ms_per_step']*1e3,2), 'ranks', [round(x*1e3,1) for x in d['per_rank_ms_per_step']], 'enq', round(d['host_enqueue_ms'],3), d['clocks'])"
done
python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline --no-host-leg 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=1', round(d['ms_per_step']*1e3,2), 'enq', round(d['host_enqueue_ms'],3), d['clocks'])"
