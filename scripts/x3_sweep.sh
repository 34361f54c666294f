#!/bin/bash
# fp32-grade tcgen05 route on cfg4: parity tests, then time by stage geometry, then the TF32 route for calibration
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tc or cfg4 or large_pool" 2>&1 | tail -4
for cfg in "2 8" "2 3" "4 8"; do
  set -- $cfg
  LWC_TC_KO=$1 LWC_TC_STAGES=$2 timeout 200 python bench.py --workload cfg4 --steps 10 --warmup 3 2>/dev/null | tail -1 > /tmp/_x.json
  python - "$1" "$2" <<'PY'
import json, sys
d = json.loads(open('/tmp/_x.json').read())
print("KO", sys.argv[1], "stages", sys.argv[2], d["ms_per_step"], round(d["roofline"]["frac"], 4))
PY
done
timeout 200 python bench.py --workload cfg4 --precision tf32 --steps 10 --warmup 3 2>/dev/null | tail -1 > /tmp/_x.json
python - <<'PY'
import json
d = json.loads(open('/tmp/_x.json').read())
print("tf32", d["ms_per_step"], round(d["roofline"]["frac"], 4))
PY
