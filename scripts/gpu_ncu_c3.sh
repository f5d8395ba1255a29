#!/bin/bash
# Full ncu capture of two wfTrace/wfShade pairs of a 16-spp C3 render (+ optional launch list).  usage: gpu_ncu_c3.sh TAG
set -u
T=${1:-x}; O=gpurun_out; mkdir -p $O
C3="python bench.py --workload c3 --spp 16 --steps 1 --warmup 1 --no-cpu-baseline"
$C3 > $O/${T}_plain_c3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'wfTrace|wfShade' -s 119 -c 4 -f -o $O/${T}_c3_wf $C3 > $O/${T}_ncu.log 2>&1
tail -2 $O/${T}_plain_c3.log | cut -c1-300
