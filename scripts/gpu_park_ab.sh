#!/bin/bash
# A/B of the two-launch trace (PRT_WF_PARK) on the multi-mesh render workloads.  usage: gpu_park_ab.sh [workload:spp:steps ...]
set -u
O=gpurun_out; mkdir -p $O
for pk in 0 1; do
  export PRT_WF_PARK=$pk
  for item in "$@"; do
    IFS=: read w spp steps <<< "$item"
    sppArg=""; [ -n "${spp:-}" ] && [ "$spp" != 0 ] && sppArg="--spp $spp"
    timeout 600 python bench.py --workload $w $sppArg --steps ${steps:-2} --warmup 3 --no-cpu-baseline > $O/park${pk}_$w.json 2> $O/park${pk}_$w.err
    python - $O/park${pk}_$w.json $pk <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("park", sys.argv[2], d["config"]["workload"][:3], round(d["value"],1), d["unit"], "e2e", round(d["e2e"]["value"],1), "launches", d.get("gpu_launches"))
except Exception as e:
    print("ERR", e)
PY
  done
done
