#!/bin/bash
# cfg3 slice (16,384 x 32 x 3072, fp32-grade) by kernel variant, then the full-size batch with the default
run() { timeout 200 python bench.py --workload cfg3 --requests $1 --steps 10 --warmup 3 2>/dev/null | tail -1 > /tmp/_c.json
  python -c "import json,sys; d=json.loads(open('/tmp/_c.json').read()); print(sys.argv[1], d['config']['kernel'], d['ms_per_step'], round(d['roofline']['frac'],4))" "$2"; }
run 16384 "default"
LWC_SMALL_KW=4 run 16384 "KW=4"
LWC_SMALL_KW=2 LWC_SMALL_U=4 run 16384 "KW=2,U=4"
LWC_SMALL_KW=2 LWC_SMALL_U=2 run 16384 "KW=2,U=2"
LWC_SMALL_IMPL=stream run 16384 "stream"
run 65536 "full default"
