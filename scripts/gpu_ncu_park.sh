#!/bin/bash
# ncu evidence for the two-launch trace on C3 (16 spp): launch list of one bench run + full capture of the first two stages'
# wfShade / wfTrace<1> / wfTrace<2> launches of the timed step.  usage: gpu_ncu_park.sh TAG
set -u
T=${1:-x}; O=gpurun_out; mkdir -p $O
C3="python bench.py --workload c3 --spp 16 --steps 1 --warmup 3 --no-cpu-baseline"
$C3 > $O/${T}_plain_c3.log 2>&1 || { tail -5 $O/${T}_plain_c3.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/${T}_c3_16spp_launches.csv $C3 > $O/${T}_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'wfTrace|wfShade' -s 258 -c 6 -f -o $O/${T}_c3_wf $C3 > $O/${T}_ncu_full.log 2>&1
tail -1 $O/${T}_plain_c3.log | cut -c1-200; tail -2 $O/${T}_ncu_full.log; ls -la $O | grep ${T}_
