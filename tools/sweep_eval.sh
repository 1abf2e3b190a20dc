#!/bin/bash
# usage: tools/sweep_eval.sh "<env settings>" ...   -- each argument is one configuration
for cfg in "$@"; do
  out=$(env $cfg timeout 300 python bench.py --steps 3 --warmup 3 --points ${POINTS:-20000000} --workload ${WORKLOAD:-cfg2} --no-cpu --no-e2e 2>&1 | tail -1)
  echo "$cfg => $(echo "$out" | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print("%.3e pm/s  %.2f ms  frac %.3f" % (d["value"], d["ms_per_step"], d["roofline"]["frac"]))
except Exception as e: print("FAILED", e)')"
done
