#!/bin/bash
# Launch list (ncu gpu__time_duration only) of one render workload under an env variant.  usage: gpu_r2_launches.sh TAG WL SPP "VAR=val VAR2=val"
set -u
T=$1; W=$2; SPP=$3; ENVS=${4:-PRT_X=0}; O=gpurun_out; mkdir -p $O
C="python bench.py --workload $W --spp $SPP --render-steps 1 --no-cpu-baseline --no-fetched"
env $ENVS $C > $O/${T}_plain.log 2>&1 && env $ENVS ncu --clock-control none --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/${T}_launches.csv $C > $O/${T}_ncu.log 2>&1
python scripts/launch_shares.py $O/${T}_launches.csv | tee $O/${T}_shares.txt
