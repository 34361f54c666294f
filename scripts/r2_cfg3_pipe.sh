#!/bin/bash
timeout 900 python -m pytest tests -q -m gpu --timeout 300 -k "small or cfg3 or random_parity or extremes" 2>&1 | tail -4 | cut -c1-300
for i in 1 2; do for ps in 4 3 0; do
LWC_SMALL_PIPE=$ps python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('pipe=$ps slice', round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['parity_sample']['tolerance_used'])"
done; done
for ps in 4 3 0; do LWC_SMALL_PIPE=$ps python bench.py --workload cfg3 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('pipe=$ps full', round(d['ms_per_step'],4), round(d['roofline']['frac'],4))"; done
