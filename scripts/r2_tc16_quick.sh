#!/bin/bash
timeout 600 python -m pytest tests -q -m gpu --timeout 300 -k "tcgen05 or fp16_piece or large_pool or cfg4" 2>&1 | tail -4 | cut -c1-200
for i in 1 2; do python bench.py --workload cfg4 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg4', round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['parity_sample']['tolerance_used'])"; done
python bench.py --workload cfg4 --precision tf32 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg4 tf32', round(d['ms_per_step'],4))"
