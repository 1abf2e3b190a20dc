#!/bin/bash
# usage: tools/dds_sweep.sh "<env>" ... ; runs the dds bench on a few shapes per configuration
for cfg in "$@"; do
  for a in "cfg4 256" "cfg4 1" "cfg2 8" "cfg5 64" "cfg3 64" "cfg2 256" "cfg1 64"; do set -- $a
    env $cfg python tools/bench_dds.py --workload $1 --k $2 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$cfg', d[\"workload\"], d[\"k\"], \"%.4f ms\"%d[\"ms\"], \"%.3e upd/s\"%d[\"updates_per_s\"], \"hbm frac %.3f\"%d[\"roofline\"][\"frac\"])"
  done
done
