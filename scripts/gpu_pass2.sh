#!/bin/bash
# GPU pass 2 of round 1: default bench (with the north-star primary-ray numbers), C5 on one GPU, launch lists and
# full ncu captures of the traversal / wavefront kernels.  Run under gpurun from the repo root.
set -u
O=gpurun_out
mkdir -p $O
python bench.py > $O/p2_bench_c2.json 2> $O/p2_bench_c2.err; cat $O/p2_bench_c2.json
timeout 600 python bench.py --workload c5 --steps 2 --warmup 1 > $O/p2_bench_c5.json 2> $O/p2_bench_c5.err; cat $O/p2_bench_c5.json; tail -3 $O/p2_bench_c5.err

C2="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$C2 > $O/p2_plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/p2_c2_launches.csv $C2 > $O/p2_ncu1.log 2>&1
$C2 > $O/p2_plain_c2b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'intersectKernelDyn|occludedKernelDyn' -s 6 -c 2 -f -o $O/p2_c2_trace $C2 > $O/p2_ncu2.log 2>&1

C3="python bench.py --workload c3 --spp 16 --steps 1 --warmup 1 --no-cpu-baseline"
$C3 > $O/p2_plain_c3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 540 -c 180 --csv --log-file $O/p2_c3_launches.csv $C3 > $O/p2_ncu3.log 2>&1
$C3 > $O/p2_plain_c3b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'wfTrace|wfShade' -s 119 -c 4 -f -o $O/p2_c3_wf $C3 > $O/p2_ncu4.log 2>&1
ls -la $O
