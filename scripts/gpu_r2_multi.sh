#!/bin/bash
# Round 2, multi-GPU pass (run under gpurun --gpus N): the tests that need >= 2 GPUs, smoke(), then the default bench line at N ranks
# launched exactly as the driver launches it, and the reference arm at the same N.
set -u
N=${1:-2}; T=${2:-m1}; O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/${T}_gpus.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -k "multi_gpu or two_contexts or one_process_per_gpu or threads" > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_pytest.log; tail -8 $O/${T}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${T}_smoke.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 \
    > $O/${T}_bench_n$N.json 2> $O/${T}_bench_n$N.err; echo "bench rc=$?"; cut -c1-400 $O/${T}_bench_n$N.json; tail -5 $O/${T}_bench_n$N.err
