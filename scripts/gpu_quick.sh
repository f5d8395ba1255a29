#!/bin/bash
# Quick GPU pass: GPU parity tests, then the bench lines of the listed workloads.  usage: gpu_quick.sh TAG [workloads...]
set -u
T=${1:-q}; shift || true
WL=${@:-"c2 c1 c3 c4"}
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; tail -3 $O/${T}_pytest.log
for w in $WL; do
  extra="--steps 2 --warmup 1 --no-cpu-baseline"; [ $w = c2 ] && extra="--no-cpu-baseline"
  timeout 600 python bench.py --workload $w $extra > $O/${T}_bench_$w.json 2> $O/${T}_bench_$w.err
  python - $O/${T}_bench_$w.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["config"]["workload"][:3], round(d["value"],1), d["unit"], "ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1),
          (d.get("roofline") or {}).get("mrays_per_s"), {k:round(v["mrays_per_s"]) for k,v in d.get("north_star_primary",{}).items() if isinstance(v,dict)})
except Exception as e:
    print("ERR", e)
PY
  tail -2 $O/${T}_bench_$w.err
done
