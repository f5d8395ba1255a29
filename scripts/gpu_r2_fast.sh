#!/bin/bash
# Fast traversal mode + the no-alloc entry point on the GPU: their tests, then exact vs fast on the render workloads.
set -u
O=gpurun_out; mkdir -p $O
# a library built with -DPRT_WIDE_CAP=16 (parks stack entries in local memory at nearly every node) through the same tests, when present
if [[ -f prt_b200/libprt_cuda_cap16.so ]]; then
  cp prt_b200/libprt_cuda.so $O/libprt_cuda_keep.so && cp prt_b200/libprt_cuda_cap16.so prt_b200/libprt_cuda.so
  timeout 900 python -m pytest tests/test_gpu_fast_mode.py -q -s > $O/f1_fast_tests_cap16.log 2>&1; echo "fast tests (spill build) rc=$?" | tee -a $O/f1_summary.log
  cp $O/libprt_cuda_keep.so prt_b200/libprt_cuda.so; rm -f $O/libprt_cuda_keep.so
fi
timeout 900 python -m pytest tests/test_gpu_fast_mode.py -q -s > $O/f1_fast_tests.log 2>&1; echo "fast tests rc=$?" | tee -a $O/f1_summary.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q > $O/f1_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" | tee -a $O/f1_summary.log
for WS in c3:64 c4:64 c5:32; do
  W=${WS%%:*}; SPP=${WS##*:}
  timeout 600 python bench.py --workload $W --spp $SPP --render-steps 4 --no-cpu-baseline --no-fetched 2>>$O/f1_err.log | tail -1 > $O/f1_bench_$W.json
  python - <<PY | tee -a $O/f1_summary.log
import json
d=json.load(open("$O/f1_bench_$W.json")); v=d["render"]["$W"]; f=v.get("fast_mode",{})
print("$W $SPP spp exact %.1f fast %.1f Msamples/s (x%.3f) pixels differing %.5f rel L1 %.2e" % (v["value"], f.get("value",0), f.get("speedup_over_exact",0), f.get("pixels_differing_from_exact",-1), f.get("relative_l1_difference",-1)))
PY
done
tail -25 $O/f1_fast_tests.log; tail -5 $O/f1_noalloc_alone.log
