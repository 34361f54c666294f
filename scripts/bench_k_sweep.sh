#!/bin/bash
# ms/step and roofline fraction of the main bench line as a function of --steps (how short a run still measures the kernel)
for k in 5 10 20 50 200; do
  timeout 300 python bench.py --steps $k --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > /tmp/_b.json
  python - "$k" <<'PY'
import json, sys
d = json.loads(open('/tmp/_b.json').read())
print(sys.argv[1], d["ms_per_step"], round(d["roofline"]["frac"], 4), round(d["e2e"]["value"]))
PY
done
