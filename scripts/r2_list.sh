#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_regime.py tests/test_gpu_parity.py -q -m gpu --timeout 300 -k "batch or list or cfg2 or overlapped" 2>&1 | tail -6 | cut -c1-300
for i in 1 2; do
for f in 1 0; do for k in 20 200; do
LWC_LIST_FUSION=$f python bench.py --steps $k --warmup 10 --no-extras --no-cpu-baseline --no-host-leg --no-cfg1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fusion=$f K=$k', round(d['ms_per_step']*1e3,2), round(d['roofline']['frac'],4), 'launches', d['gpu_launches'], 'agree', d['host_device_paths_agree'], d['parity_sample']['tolerance_used'])"
done; done; done
