#!/bin/bash
# A/B on one box: cfg2 main kernel with and without the accumulator-chain flush code
for i in 1 2 3; do
for lib in "" "$PWD/build/liblwc_noflush.so"; do
for k in 20 200; do
LWC_B200_LIB=$lib python bench.py --steps $k --warmup 10 --no-extras --no-cpu-baseline --no-host-leg | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('lib=${lib:-default}', 'K=$k', round(d['ms_per_step']*1e3,2), round(d['roofline']['frac'],4))"
done; done; done
