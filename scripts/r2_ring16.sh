#!/bin/bash
timeout 900 python -m pytest tests -q -m gpu --timeout 300 -k "small or cfg3 or random_parity or extremes or batch" 2>&1 | tail -6 | cut -c1-300
for i in 1 2; do
echo "ring f16 NCW=8"; python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['parity_sample']['tolerance_used'], d['config']['kernel'])"
echo "ring f16 NCW=12"; LWC_STREAM_NCW=12 python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['frac'],4))"
echo "direct f16"; LWC_SMALL_IMPL=direct python bench.py --workload cfg3 --requests 16384 --steps 20 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['frac'],4))"
done
echo "full ring f16"; python bench.py --workload cfg3 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['frac'],4))"
echo "full direct f16"; LWC_SMALL_IMPL=direct python bench.py --workload cfg3 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['frac'],4))"
