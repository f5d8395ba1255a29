#!/bin/bash
# Evidence pass on the current kernels: default bench, reference arm, launch lists and full ncu captures (c2, c3@16spp, c5@8spp).
set -u
T=${1:-f1}; O=gpurun_out; mkdir -p $O
python bench.py --impl reference --steps 5 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err
python bench.py > $O/${T}_bench_c2.json 2> $O/${T}_bench_c2.err; cut -c1-300 $O/${T}_bench_c2.json
C2="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$C2 > $O/${T}_plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${T}_c2_launches.csv $C2 > $O/${T}_ncu1.log 2>&1
$C2 > $O/${T}_plain_c2b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'intersectKernelDyn|occludedKernelDyn' -s 6 -c 2 -f -o $O/${T}_c2_trace $C2 > $O/${T}_ncu2.log 2>&1
C3="python bench.py --workload c3 --spp 16 --steps 1 --warmup 1 --no-cpu-baseline"
$C3 > $O/${T}_plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 135 -c 45 --csv --log-file $O/${T}_c3_launches.csv $C3 > $O/${T}_ncu3.log 2>&1
$C3 > $O/${T}_plain_c3b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'wfTrace|wfShade' -s 32 -c 4 -f -o $O/${T}_c3_wf $C3 > $O/${T}_ncu4.log 2>&1
C5="python bench.py --workload c5 --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
$C5 > $O/${T}_plain_c5.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'wfTrace' -s 1 -c 1 -f -o $O/${T}_c5_wf $C5 > $O/${T}_ncu5.log 2>&1
ls -la $O | grep ${T}_
