#!/bin/bash
# ncu evidence for the fast traversal mode: the first trace launches of the same frame (seed 100, the warm-up render) in the exact
# kernel (wfTrace) and in the wide one (wfTraceWide), `--set full`, plus the launch list of the whole command.
# usage: gpu_r2_ncu_fast.sh TAG "c4:16 c5:8"
set -u
T=${1:-n2}; WLS=${2:-"c4:16 c5:8"}; O=gpurun_out; mkdir -p $O
NCU="ncu --clock-control none"
digest() {
  ncu -i $O/$1.ncu-rep --page raw --csv > $O/$1_raw.csv 2>/dev/null
  ncu -i $O/$1.ncu-rep --page source --csv --print-source cuda,sass > $O/$1_source.csv 2> $O/$1_source.err
  python scripts/ncu_summary.py $O/$1_raw.csv > $O/$1_summary.txt 2>&1
  python scripts/ncu_source_split.py $O/$1_source.csv $O/$1_lines 40 > $O/$1_split.log 2>&1
  head -c 20000000 $O/$1_source.csv | gzip > $O/$1_source.csv.gz
  rm -f $O/$1.ncu-rep $O/$1_source.csv
}
for WS in $WLS; do
  W=${WS%%:*}; SPP=${WS##*:}
  C="python bench.py --workload $W --spp $SPP --render-steps 1 --no-cpu-baseline --no-fetched"
  $C > $O/${T}_${W}_plain.log 2>&1 && {
    $NCU --metrics gpu__time_duration.sum -c 1200 --csv --log-file $O/${T}_${W}_launches.csv $C > $O/${T}_${W}_ncu1.log 2>&1
    $NCU --set full --import-source on --kernel-name-base function -k regex:'^wfTrace$' -c 3 -f -o $O/${T}_${W}_exact $C > $O/${T}_${W}_ncu2.log 2>&1; digest ${T}_${W}_exact
    $NCU --set full --import-source on --kernel-name-base function -k regex:'^wfTraceWide$' -c 3 -f -o $O/${T}_${W}_wide $C > $O/${T}_${W}_ncu3.log 2>&1; digest ${T}_${W}_wide
  }
  tail -1 $O/${T}_${W}_plain.log | cut -c1-160
  grep -h "Kernel Name\|gpu__time_duration\|data_pipe\|l1tex__throughput\|inst_executed.sum\|registers" $O/${T}_${W}_exact_summary.txt $O/${T}_${W}_wide_summary.txt | cut -c1-160
done
du -sh $O
