#!/bin/bash
# A/B of PRT_TRACE_FLAGS values on short bench runs.  usage: gpu_ab.sh "0 1" [workload:spp ...]
set -u
FLAGS=$1; shift
O=gpurun_out; mkdir -p $O
for fl in $FLAGS; do
  export PRT_TRACE_FLAGS=$fl
  python -m pytest tests -m gpu -x -q 2>&1 | tail -1
  for item in "$@"; do
    w=${item%%:*}; spp=""; [ "$item" != "$w" ] && spp="--spp ${item##*:}"
    steps="--steps 2 --warmup 1"; [ $w = c2 ] && steps="--steps 20 --warmup 3"
    timeout 600 python bench.py --workload $w $spp $steps --no-cpu-baseline > $O/ab_${fl}_$w.json 2> $O/ab_${fl}_$w.err
    python - $O/ab_${fl}_$w.json $fl <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("flags", sys.argv[2], d["config"]["workload"][:3], round(d["value"],1), d["unit"], (d.get("roofline") or {}).get("mrays_per_s"))
except Exception as e:
    print("ERR", e)
PY
  done
done
