#!/bin/bash
# A/B of environment-switched variants of the fast traversal mode.  usage: gpu_r2_fast_ab.sh TAG "wl:spp ..." "VAR=val;VAR=val,VAR2=val;..."
set -u
T=$1; WLS=$2; VARIANTS=$3; O=gpurun_out; mkdir -p $O
IFS=';' read -ra VS <<< "$VARIANTS"
for WS in $WLS; do
  W=${WS%%:*}; SPP=${WS##*:}
  for V in "${VS[@]}"; do
    ENVS=$(echo "$V" | tr ',' ' ')
    OUT=$(env $ENVS python bench.py --workload $W --spp $SPP --render-steps 2 --no-cpu-baseline --no-fetched 2>>$O/${T}_err.log | tail -1)
    echo "$OUT" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); v=d['render']['$W']; f=v.get('fast_mode',{})
print('$W', '$SPP spp', '[$V]', 'exact %.1f fast %.1f Msamples/s' % (v['value'], f.get('value',0)), 'pixels differing %.5f' % f.get('pixels_differing_from_exact',-1))
" | tee -a $O/${T}_ab.log
  done
done
