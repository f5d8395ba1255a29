#!/bin/bash
# Multi-GPU pass: bench lines at N ranks (torchrun, one rank per GPU).  usage: gpu_multi.sh TAG N [workload[:spp]...]
set -u
T=$1; N=$2; shift 2
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8
for item in "$@"; do
  w=${item%%:*}; spp=""; [ "$item" != "$w" ] && spp="--spp ${item##*:}"
  steps="--steps 2 --warmup 1"; [ $w = c2 ] && steps="--steps 20 --warmup 3"
  if [ $N = 1 ]; then cmd="python bench.py"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py"; fi
  timeout 900 $cmd --gpus $N --workload $w $spp $steps --no-cpu-baseline > $O/${T}_n${N}_$w.json 2> $O/${T}_n${N}_$w.err
  python - $O/${T}_n${N}_$w.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["config"]["workload"][:3], "N", d["n_gpus"], round(d["value"],1), d["unit"], "ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), d.get("clocks"))
except Exception as e:
    print("ERR", e)
PY
  tail -3 $O/${T}_n${N}_$w.err | cut -c1-300
done
