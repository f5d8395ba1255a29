#!/bin/bash
# A/B of environment-switched variants on render workloads.  usage: gpu_r2_ab.sh TAG "wl:spp wl:spp ..." "VAR=val,VAR2=val;VAR=val;..."
set -u
T=$1; WLS=$2; VARIANTS=$3; O=gpurun_out; mkdir -p $O
IFS=';' read -ra VS <<< "$VARIANTS"
for WS in $WLS; do
  W=${WS%%:*}; SPP=${WS##*:}
  for V in "${VS[@]}"; do
    ENVS=$(echo "$V" | tr ',' ' ')
    if [[ $W == c2 ]]; then C="python bench.py --workload c2 --steps 5 --passes 16 --no-cpu-baseline --no-fetched"
    else C="python bench.py --workload $W --spp $SPP --render-steps 3 --no-cpu-baseline --no-fetched"; fi
    OUT=$(env $ENVS $C 2>>$O/${T}_err.log | tail -1)
    echo "$OUT" | python -c "
import json,sys
d=json.loads(sys.stdin.read())
if '$W'=='c2':
    r=d['rules']; print('$W', '[$V]', 'packet %.0f (c %.0f o %.0f) single %.0f (c %.0f o %.0f)' % (r['packet']['value'], r['packet']['roofline']['mrays_per_s']['closest_hit'], r['packet']['roofline']['mrays_per_s']['occlusion'], r['single']['value'], r['single']['roofline']['mrays_per_s']['closest_hit'], r['single']['roofline']['mrays_per_s']['occlusion']))
else:
    v=d['render']['$W']; print('$W', '$SPP spp', '[$V]', '%.1f Msamples/s' % v['value'], 'crc', v['image_crc'])
" | tee -a $O/${T}_ab.log
  done
done
