#!/bin/bash
# Refill-threshold sweep (traceFlags bits 8-15; bit 1 = front-to-back any-hit order as the default) on render workloads.
# usage: gpu_refill_ab.sh "4 8 12 16" workload:spp:steps ...
set -u
TH=$1; shift
O=gpurun_out; mkdir -p $O
for th in $TH; do
  export PRT_TRACE_FLAGS=$(( (th << 8) | 2 ))
  for item in "$@"; do
    IFS=: read w spp steps <<< "$item"
    sppArg=""; [ -n "${spp:-}" ] && [ "$spp" != 0 ] && sppArg="--spp $spp"
    timeout 600 python bench.py --workload $w $sppArg --steps ${steps:-2} --warmup 3 --no-cpu-baseline > $O/refill${th}_$w.json 2> $O/refill${th}_$w.err
    python - $O/refill${th}_$w.json $th <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("refill", sys.argv[2], d["config"]["workload"][:3], round(d["value"],1), d["unit"])
except Exception as e:
    print("ERR", e)
PY
  done
done
