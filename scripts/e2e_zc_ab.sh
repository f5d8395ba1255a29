set -x
for zc in 0 1 2; do PRT_ZC=$zc timeout 300 python scripts/e2e_zc_probe.py 2>&1 | tail -8; done > gpurun_out/zc_probe.log 2>&1
for dg in 1 4; do PRT_ZC=1 PRT_D2H_GROUP=$dg timeout 300 python scripts/e2e_zc_probe.py 2>&1 | tail -8; done >> gpurun_out/zc_probe.log 2>&1
cat gpurun_out/zc_probe.log
