#!/bin/bash
# A/B of environment switches on short bench runs, no pytest.  usage: gpu_ab2.sh TAG "ENV=VAL ENV=VAL|ENV=VAL ..." [workload:spp ...]
set -u
T=$1; VARIANTS=$2; shift 2
O=gpurun_out; mkdir -p $O
IFS='|' read -ra VS <<< "$VARIANTS"
i=0
for v in "${VS[@]}"; do
  for item in "$@"; do
    w=${item%%:*}; spp=""; [ "$item" != "$w" ] && spp="--spp ${item##*:}"
    steps="--steps 2 --warmup 1"; [ $w = c2 ] && steps="--steps 20 --warmup 3"
    env $v timeout 600 python bench.py --workload $w $spp $steps --no-cpu-baseline > $O/${T}_${i}_$w.json 2> $O/${T}_${i}_$w.err
    python - $O/${T}_${i}_$w.json "$v" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"[{sys.argv[2]}]", d["config"]["workload"][:3], round(d["value"],1), d["unit"], "e2e", round(d["e2e"]["value"],1), (d.get("roofline") or {}).get("mrays_per_s"))
except Exception as e:
    print("ERR", sys.argv[2], e)
PY
  done
  i=$((i+1))
done
