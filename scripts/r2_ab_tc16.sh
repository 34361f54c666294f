#!/bin/bash
# A/B of two builds of the fp32-grade tcgen05 route on one box: cfg4 at full size, and the 64-candidate pair form
for i in 1 2 3; do
for lib in "$PWD/build/liblwc_head.so" "$PWD/build/liblwc_ysplit.so" "$PWD/build/liblwc_fresh.so"; do
LWC_B200_LIB=$lib python bench.py --workload cfg4 --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('lib=$(basename $lib)', round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['parity_sample']['tolerance_used'])"
done; done
for lib in "$PWD/build/liblwc_head.so" "$PWD/build/liblwc_ysplit.so" "$PWD/build/liblwc_fresh.so"; do
LWC_B200_LIB=$lib python scripts/r2_pool_sizes.py 64,96,128 | grep fp32 | cut -c1-200
done
