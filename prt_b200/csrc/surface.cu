// Scene::getSurfaceProperties (scene.cpp:30-45 -> Mesh::getSurfaceProperties, mesh.cpp:311-372) over a batch of public hit
// records.  A host integrator that traces through prt_cuda_intersect and shades on its own side gets the same
// SurfaceProperties the reference hands to its bounce loop (path_tracer.cpp:99, :281): interpolated or geometric normal,
// uv, material, and the bump frame dp01 / dp02 / duv01 / duv02.
//
// A public hit names its triangle the reference's way — (meshId, primId), the primitive's index in the mesh as it was
// added — while the device works with the scene-wide leaf-ordered triangle index; DScene::primToTri is the inverse of the
// builder's m_primRemapping and takes one to the other.  The arithmetic is surfaceProps() of prt_shade.cuh, the routine the
// wavefront path tracer itself shades with (bit-identical to the reference in the render parity tests).
#include "prt_context.h"
#include "prt_shade.cuh"

using namespace prt_b200;

namespace {

__global__ void surfaceKernel(DScene sc, const prt_hit* __restrict__ hits, prt_surface* __restrict__ out, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2* hp = reinterpret_cast<const float2*>(hits + i);
    const float2 h0 = __ldg(hp), h1 = __ldg(hp + 1), h2 = __ldg(hp + 2);
    float4* o = reinterpret_cast<float4*>(out + i);
    const uint32_t primId = __float_as_uint(h2.x), meshId = __float_as_uint(h2.y);
    if (h0.x == -1.0f || meshId >= sc.meshCount) {                // miss (RayHit::kMissT): all-zero record
        const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        o[0] = z; o[1] = z; o[2] = z; o[3] = z;
        return;
    }
    const DMesh& m = sc.meshes[meshId];
    Hit h;
    h.t = h0.x; h.i = h0.y; h.j = h1.x; h.k = h1.y; h.meshId = meshId;
    h.tri = __ldg(sc.primToTri + m.primBase + primId);
    const Surf s = surfaceProps(sc, h, true);
    o[0] = make_float4(s.normal.x, s.normal.y, s.normal.z, __uint_as_float(s.matRef - m.materialBase));   // Mesh::getPrimToMaterial(primId)
    o[1] = make_float4(s.u, s.v, s.duv01x, s.duv01y);
    o[2] = make_float4(s.duv02x, s.duv02y, s.dp01.x, s.dp01.y);
    o[3] = make_float4(s.dp01.z, s.dp02.x, s.dp02.y, s.dp02.z);
}

}  // namespace

static_assert(sizeof(prt_surface) == 64, "prt_surface is written as four float4");

extern "C" {

int prt_cuda_surface_properties_dev(prt_ctx* ctx, uint32_t n, const prt_hit* d_hits, prt_surface* d_out) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!d_hits || !d_out) return setError(ctx, PRT_ERR_ARG, "surface_properties: null buffer");
    if (ctx->meshes.empty()) return setError(ctx, PRT_ERR_STATE, "surface_properties: the scene has no mesh");
    surfaceKernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->scene, d_hits, d_out, n);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    ctx->counters.kernelLaunches++;
    return PRT_OK;
}

int prt_cuda_surface_properties(prt_ctx* ctx, uint32_t n, const prt_hit* hits, prt_surface* out) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!hits || !out) return setError(ctx, PRT_ERR_ARG, "surface_properties: null buffer");
    if ((r = ensureScratch(ctx, 0, (size_t)n * sizeof(prt_hit)))) return r;
    if ((r = ensureScratch(ctx, 1, (size_t)n * sizeof(prt_surface)))) return r;
    // primId is the caller's: reject indices outside the mesh instead of reading past primToTri
    for (uint32_t i = 0; i < n; i++)
        if (hits[i].t != -1.0f && (hits[i].meshId >= ctx->meshes.size() || hits[i].primId >= ctx->meshes[hits[i].meshId]->primCount))
            return setError(ctx, PRT_ERR_ARG, "surface_properties: hit %u names primitive %u of mesh %u, which does not exist", i, hits[i].primId, hits[i].meshId);
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(ctx->d_scratch[0], hits, (size_t)n * sizeof(prt_hit), cudaMemcpyHostToDevice, ctx->stream));
    if ((r = prt_cuda_surface_properties_dev(ctx, n, (const prt_hit*)ctx->d_scratch[0], (prt_surface*)ctx->d_scratch[1]))) return r;
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(out, ctx->d_scratch[1], (size_t)n * sizeof(prt_surface), cudaMemcpyDeviceToHost, ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return PRT_OK;
}

}  // extern "C"
