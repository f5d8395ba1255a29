// Roofline denominators measured on the device the context runs on ("additionally against a measured L2 read
// bandwidth (builder micro-benchmark) and FP32 issue rate").  The traversal kernels are cache-resident on C1-C3, so the
// HBM copy peak alone does not bound them; bench.py reports these two beside it.  Diagnostics only: nothing on the ray
// path calls into this file.
#include <cuda_runtime.h>
#include "prt_context.h"
using namespace prt_b200;

namespace {

// Every thread streams the whole buffer (L2-resident after the first pass) with independent 128-bit ld.global.cg loads
// (cached at L2 only, so L1 hits cannot inflate the figure).
__global__ void __launch_bounds__(256) l2ReadKernel(const uint4* __restrict__ buf, uint32_t n16, uint32_t iters, uint32_t* sink) {
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t acc = 0;
    for (uint32_t it = 0; it < iters; it++) {
        uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < n16; i += 4 * stride) {
            const uint4 a = __ldcg(buf + i), b = __ldcg(buf + i + stride), c = __ldcg(buf + i + 2 * stride), d = __ldcg(buf + i + 3 * stride);
            acc ^= a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w ^ c.x ^ c.y ^ c.z ^ c.w ^ d.x ^ d.y ^ d.z ^ d.w;
        }
        for (; i < n16; i += stride) { const uint4 a = __ldcg(buf + i); acc ^= a.x ^ a.y ^ a.z ^ a.w; }
    }
    if (acc == 0x9e3779b9u) *sink = acc;        // never true for the zero-filled buffer; keeps the loads alive
}

// 8 independent FP32 chains per thread, 128 FFMA per loop trip (loop overhead < 3 %): warp-instruction issue rate of
// the FP32 pipe with enough warps resident to cover its latency.
constexpr uint32_t kIssuePerTrip = 128;
__global__ void __launch_bounds__(256) issueKernel(uint32_t trips, float* sink) {
    float a0 = threadIdx.x, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 0.999f, c = 0.001f;
    for (uint32_t t = 0; t < trips; t++) {
#pragma unroll
        for (uint32_t u = 0; u < kIssuePerTrip / 8; u++) {
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a0) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a1) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a2) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a3) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a4) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a5) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a6) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a7) : "f"(m), "f"(c));
        }
    }
    const float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 12345.678f) *sink = s;
}

}  // namespace

extern "C" int prt_cuda_probe_peaks(prt_ctx* ctx, double* l2ReadGBs, double* fp32WarpGinstS) {
    if (!ctx) return PRT_ERR_ARG;
    if (!l2ReadGBs || !fp32WarpGinstS) return setError(ctx, PRT_ERR_ARG, "probe_peaks: null output");
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "probe_peaks: host-only context");
    DeviceGuard guard(ctx->device);
    int sms = 0;
    PRT_CUDA_CHECK(ctx, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
    const size_t bytes = 32u << 20;                 // well inside one L2 partition
    const uint32_t n16 = (uint32_t)(bytes / 16), iters = 200, trips = 4096;
    void* buf = nullptr;
    uint32_t* sink = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = PRT_OK;
    auto fail = [&](cudaError_t e, const char* what) { rc = setError(ctx, PRT_ERR_CUDA, "probe_peaks: %s: %s", what, cudaGetErrorString(e)); };
    cudaError_t e;
    do {
        if ((e = cudaMalloc(&buf, bytes)) != cudaSuccess) { fail(e, "cudaMalloc"); break; }
        if ((e = cudaMalloc((void**)&sink, 16)) != cudaSuccess) { fail(e, "cudaMalloc"); break; }
        if ((e = cudaMemsetAsync(buf, 0, bytes, ctx->stream)) != cudaSuccess) { fail(e, "cudaMemset"); break; }
        if ((e = cudaEventCreate(&e0)) != cudaSuccess || (e = cudaEventCreate(&e1)) != cudaSuccess) { fail(e, "cudaEventCreate"); break; }
        const int grid = sms * 8;
        float ms = 0.f;
        l2ReadKernel<<<grid, 256, 0, ctx->stream>>>((const uint4*)buf, n16, 2, sink);                 // warm L2, clocks
        cudaEventRecord(e0, ctx->stream);
        l2ReadKernel<<<grid, 256, 0, ctx->stream>>>((const uint4*)buf, n16, iters, sink);
        cudaEventRecord(e1, ctx->stream);
        if ((e = cudaEventSynchronize(e1)) != cudaSuccess) { fail(e, "l2ReadKernel"); break; }
        cudaEventElapsedTime(&ms, e0, e1);
        *l2ReadGBs = (double)bytes * iters / (ms * 1e-3) / 1e9;
        issueKernel<<<grid, 256, 0, ctx->stream>>>(64, (float*)sink);
        cudaEventRecord(e0, ctx->stream);
        issueKernel<<<grid, 256, 0, ctx->stream>>>(trips, (float*)sink);
        cudaEventRecord(e1, ctx->stream);
        if ((e = cudaEventSynchronize(e1)) != cudaSuccess) { fail(e, "issueKernel"); break; }
        cudaEventElapsedTime(&ms, e0, e1);
        *fp32WarpGinstS = (double)grid * 8.0 /*warps per block*/ * trips * kIssuePerTrip / (ms * 1e-3) / 1e9;
    } while (0);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (buf) cudaFree(buf);
    if (sink) cudaFree(sink);
    return rc;
}
