#!/bin/bash
# Round 2, first GPU pass: the GPU test suite, smoke(), then the default bench line (N = 1) and the reference arm.
set -u
T=${1:-a1}; O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/${T}_gpus.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_pytest.log; tail -15 $O/${T}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/${T}_smoke.log
timeout 900 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"; cut -c1-600 $O/${T}_bench.json; tail -5 $O/${T}_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref rc=$?"; cut -c1-300 $O/${T}_bench_ref.json
