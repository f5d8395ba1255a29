#!/bin/bash
# A/B of two builds of the library on the fast-mode render numbers.  usage: gpu_r2_lib_ab.sh TAG variant.so "wl:spp ..."
set -u
T=$1; VAR=$2; WLS=$3; O=gpurun_out; mkdir -p $O
cp prt_b200/libprt_cuda.so $O/keep.so
for L in base variant; do
  if [[ $L == variant ]]; then cp $VAR prt_b200/libprt_cuda.so; fi
  for WS in $WLS; do
    W=${WS%%:*}; SPP=${WS##*:}
    OUT=$(python bench.py --workload $W --spp $SPP --render-steps 3 --no-cpu-baseline --no-fetched 2>>$O/${T}_err.log | tail -1)
    echo "$OUT" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); v=d['render']['$W']; f=v.get('fast_mode',{})
print('$L', '$W', '$SPP spp', 'exact %.1f fast %.1f Msamples/s' % (v['value'], f.get('value',0)), 'pixels differing %.5f' % f.get('pixels_differing_from_exact',-1), v['image_crc'])
" | tee -a $O/${T}_ab.log
  done
done
cp $O/keep.so prt_b200/libprt_cuda.so; rm -f $O/keep.so
