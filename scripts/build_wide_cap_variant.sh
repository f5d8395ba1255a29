#!/bin/bash
# Builds prt_b200/libprt_cuda_cap16.so: the product library with the fast mode's shared-memory traversal stack cut to 16 entries, the minimum
# (-DPRT_WIDE_CAP=16), so that every ray whose stack passes 8 entries parks entries in local memory and takes them back (prt_wide.cuh, wideSpill /
# wideRefill) — a path real scenes take for a handful of rays only.  scripts/gpu_r2_fast.sh runs tests/test_gpu_fast_mode.py
# against it.  The variant also carries bounds checks on the stack pointers (-DPRT_WIDE_CHECK: a violated bound traps the launch, which
# the tests see as a CUDA error) — compute-sanitizer is not available on the GPU pool.  Test aid, git-ignored like every .so.
set -e
cd "$(dirname "$0")/../prt_b200/csrc"
make -j8 > /dev/null
mkdir -p build/cap16
NV="/usr/local/cuda/bin/nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-O3,-ffp-contract=off,-fno-fast-math,-Wall -DPRT_WIDE_CAP=16 -DPRT_WIDE_CHECK=1"
$NV -c trace_kernels.cu -o build/cap16/trace_kernels.o 2> build/cap16/trace_kernels.log &
$NV -c wavefront.cu -o build/cap16/wavefront.o 2> build/cap16/wavefront.log &
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libprt_cuda_cap16.so build/prt_cuda.o build/cap16/trace_kernels.o build/cap16/wavefront.o \
    build/aov_output.o build/probe_kernels.o build/multi_gpu.o build/surface.o build/bvh_build.o -ldl -lpthread
ls -la ../libprt_cuda_cap16.so
