#!/bin/bash
# Full ncu capture of wfTrace/wfShade launches of a render workload.  usage: gpu_ncu_wl.sh TAG WORKLOAD SPP SKIP COUNT
set -u
T=$1; W=$2; SPP=$3; SK=$4; CN=$5; O=gpurun_out; mkdir -p $O
C="python bench.py --workload $W --spp $SPP --steps 1 --warmup 1 --no-cpu-baseline"
$C > $O/${T}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'wfTrace|wfShade' -s $SK -c $CN -f -o $O/${T}_wf $C > $O/${T}_ncu.log 2>&1
tail -1 $O/${T}_plain.log | cut -c1-200; tail -2 $O/${T}_ncu.log
