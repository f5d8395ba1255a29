#!/bin/bash
# A/B of raw PRT_TRACE_FLAGS values on render workloads.  usage: gpu_flags_ab.sh "3075 3091" workload:spp:steps ...
set -u
V=$1; shift
O=gpurun_out; mkdir -p $O
for fl in $V; do
  export PRT_TRACE_FLAGS=$fl
  for item in "$@"; do
    IFS=: read w spp steps <<< "$item"
    sppArg=""; [ -n "${spp:-}" ] && [ "$spp" != 0 ] && sppArg="--spp $spp"
    timeout 600 python bench.py --workload $w $sppArg --steps ${steps:-2} --warmup 3 --no-cpu-baseline > $O/fl${fl}_$w.json 2> $O/fl${fl}_$w.err
    python - $O/fl${fl}_$w.json $fl <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("flags", sys.argv[2], d["config"]["workload"][:3], round(d["value"],1), d["unit"])
except Exception as e:
    print("ERR", e)
PY
  done
done
