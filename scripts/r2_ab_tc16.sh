#!/bin/bash
for i in 1 2; do
for lib in "" "$PWD/build/liblwc_bo_120.so" "$PWD/build/liblwc_bo_250.so"; do
LWC_B200_LIB=$lib python bench.py --workload cfg4 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('lib=${lib:-default}', round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['parity_sample']['tolerance_used'])"
done; done
for st in 5 3; do LWC_TC_STAGES=$st python bench.py --workload cfg4 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('stages=$st', round(d['ms_per_step'],4))"; done
