#!/bin/bash
# Round 2 ncu evidence pass (one GPU): launch lists + `--set full` captures of the dominant kernels of c2 (both rule sets), c3, c4, c5.
# Every capture follows a plain run of the same command that exited 0.  The .ncu-rep files are turned into the raw-page CSV and
# the per-source-line summary on the box and deleted there (gpurun brings back at most 64 MiB).  usage: gpu_r2_ncu.sh TAG [parts]
set -u
T=${1:-n1}; PARTS=${2:-c2,c3,c4,c5}; O=gpurun_out; mkdir -p $O
NCU="ncu --clock-control none"
digest() {   # rep-basename
  ncu -i $O/$1.ncu-rep --page raw --csv > $O/$1_raw.csv 2>/dev/null
  ncu -i $O/$1.ncu-rep --page source --csv --print-source cuda,sass > $O/$1_source.csv 2>/dev/null
  python scripts/ncu_summary.py $O/$1_raw.csv > $O/$1_summary.txt 2>&1
  python scripts/ncu_source_split.py $O/$1_source.csv $O/$1_lines 30 > /dev/null 2>&1
  rm -f $O/$1.ncu-rep $O/$1_source.csv
}
if [[ $PARTS == *c2* ]]; then
  C="python bench.py --workload c2 --steps 1 --passes 4 --no-cpu-baseline --no-fetched"
  $C > $O/${T}_c2_plain.log 2>&1 && {
    $NCU --metrics gpu__time_duration.sum -c 120 --csv --log-file $O/${T}_c2_launches.csv $C > $O/${T}_c2_ncu1.log 2>&1
    $NCU --set full --import-source on --kernel-name-base mangled -k regex:'KernelDynILi1E' -s 3 -c 2 -f -o $O/${T}_c2_packet $C > $O/${T}_c2_ncu2.log 2>&1; digest ${T}_c2_packet
    $NCU --set full --import-source on --kernel-name-base mangled -k regex:'KernelDynILi0E' -s 3 -c 2 -f -o $O/${T}_c2_single $C > $O/${T}_c2_ncu3.log 2>&1; digest ${T}_c2_single
  }
fi
for W in c3 c4 c5; do
  if [[ $PARTS == *$W* ]]; then
    SPP=16; [[ $W == c5 ]] && SPP=8
    C="python bench.py --workload $W --spp $SPP --render-steps 1 --no-cpu-baseline --no-fetched"
    $C > $O/${T}_${W}_plain.log 2>&1 && {
      $NCU --metrics gpu__time_duration.sum -c 400 --csv --log-file $O/${T}_${W}_launches.csv $C > $O/${T}_${W}_ncu1.log 2>&1
      $NCU --set full --import-source on -k regex:'wfTrace|wfShade|wfPrimary' -s 3 -c 7 -f -o $O/${T}_${W}_wf $C > $O/${T}_${W}_ncu2.log 2>&1; digest ${T}_${W}_wf
    }
    tail -1 $O/${T}_${W}_plain.log | cut -c1-200
  fi
done
du -sh $O; ls -la $O | grep ${T}_ | head -60
