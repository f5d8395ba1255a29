// G-buffer pass and output stage.
//
//   gbufferKernel   GbufferVisualizer::Trace (gbuffer_visualizer.cpp:28-51): one jittered single-ray primary per pixel
//                   (Camera::GenerateJitteredRay, camera.cpp:12-33), single-ray closest hit, Mesh::getSurfaceProperties,
//                   then diffuse colour (material.cpp:87-96) or bumped normal * 0.5 + 0.5 (material.cpp:98-114);
//                   Image::writePixel's exposure multiply (image.cpp:44-50).  The visualiser's RNG is reseeded per
//                   pixel with the 1-spp packet hash, so the camera ray equals prt_cuda_primary_dev(SINGLE, 1 spp).
//   ppmKernel       Image::savePpm payload (image.cpp:60-69): c/(c+1) if tonemap, fmin(fmax(c,0),1), powf(c,1/2.2f)*255,
//                   truncation to uint8.
//   halfKernel      the FP32 -> half conversion tinyexr applies in Image::saveExr (tinyexr.h:899-934 float_to_half_full):
//                   round-half-up on the magnitude, denormal inputs flushed, NaN -> 0x200; planes in B, G, R order
//                   (image.cpp:106-118).
//
// All three are streaming kernels: grid-stride, one pixel per thread, coalesced 12-byte pixel reads; bound by HBM.
#include "prt_context.h"
#include "prt_shade.cuh"

using namespace prt_b200;

namespace {

__device__ __forceinline__ uint32_t pixelSeed(uint32_t seed, uint32_t pixel) {          // prt_packet_seed(seed, pixel, 1, 0)
    uint32_t h = seed * 0x9E3779B9u + pixel;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h | 1u;
}

__global__ void __launch_bounds__(128) gbufferKernel(DScene sc, uint32_t x0, uint32_t y0, uint32_t bw, uint32_t nPix, int type, uint32_t seed,
                                                     float exposure, float* __restrict__ rgb) {
    uint32_t stack[kStackSize];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPix) return;
    const DCamera& c = sc.cam;
    const uint32_t px = x0 + i % bw, py = y0 + i / bw;
    uint32_t s = pixelSeed(seed, py * c.width + px);
    s = xs32(s); const float u0 = rngFloat(s);
    s = xs32(s); const float u1 = rngFloat(s);
    const float kScreenScale = 0.6f;
    const float kAspect = (float)c.width / (float)c.height;
    const float jx = (2.0f * u0 - 1.0f) * (0.5f * c.invW);
    const float jy = (2.0f * u1 - 1.0f) * (0.5f * c.invH);
    const float nx = 2.0f * ((float)px * c.invW - 0.5f + jx) * kScreenScale * kAspect;
    const float ny = -2.0f * ((float)py * c.invH - 0.5f + jy) * kScreenScale;
    const F3 d = normalize3(nx * f3(c.right[0], c.right[1], c.right[2]) + ny * f3(c.up[0], c.up[1], c.up[2]) + f3(c.dir[0], c.dir[1], c.dir[2]));
    const RayD r = makeRay(c.pos[0], c.pos[1], c.pos[2], d.x, d.y, d.z, 100000.0f, false);
    const Hit h = sceneIntersectSingle(sc, r, stack);
    F3 color = f3(0.0f, 0.0f, 0.0f);
    if (h.t != -1.0f) {
        const uint32_t matRef = surfaceMaterial(sc, h);
        const DMaterial& mt = sc.materials[matRef];
        const Surf sp = surfaceProps(sc, h, type != 0 && mt.bumpTex >= 0);
        if (type == 0) color = sampleDiffuse(sc, mt, sp.u, sp.v);
        else color = sampleBump(sc, mt, sp) * 0.5f + f3(0.5f, 0.5f, 0.5f);
    }
    float* o = rgb + 3 * ((size_t)py * c.width + px);
    o[0] = exposure * color.x; o[1] = exposure * color.y; o[2] = exposure * color.z;
}

__global__ void __launch_bounds__(256) ppmKernel(const float* __restrict__ rgb, uint8_t* __restrict__ out, size_t n, int tonemap) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float c = __ldg(rgb + i);
        if (tonemap) c = c / (c + 1.0f);
        c = fminf(fmaxf(c, 0.0f), 1.0f);
        c = powRef(c, 1 / 2.2f) * 255.0f;
        out[i] = (uint8_t)c;
    }
}

__device__ __forceinline__ uint16_t halfBitsTinyexr(float f) {
    const uint32_t u = __float_as_uint(f);
    const uint32_t sign = u >> 31, e = (u >> 23) & 0xffu, m = u & 0x7fffffu;
    uint32_t o = 0;
    if (e == 0) o = 0;
    else if (e == 255) o = (31u << 10) | (m ? 0x200u : 0u);
    else {
        const int ne = (int)e - 127 + 15;
        if (ne >= 31) o = 31u << 10;
        else if (ne <= 0) {
            if ((14 - ne) <= 24) {
                const uint32_t mant = m | 0x800000u;
                o = (mant >> (14 - ne)) & 0x3ffu;
                if ((mant >> (13 - ne)) & 1u) o++;
            }
        } else {
            o = ((uint32_t)ne << 10) | (m >> 13);
            if (m & 0x1000u) o++;
        }
    }
    return (uint16_t)((o & 0xffffu) | (sign << 15));
}

// planes[0] = B, planes[1] = G, planes[2] = R, each nPix halfs
__global__ void __launch_bounds__(256) halfKernel(const float* __restrict__ rgb, uint16_t* __restrict__ planes, size_t nPix) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nPix; i += (size_t)gridDim.x * blockDim.x) {
        const float r = __ldg(rgb + 3 * i), g = __ldg(rgb + 3 * i + 1), b = __ldg(rgb + 3 * i + 2);
        planes[i] = halfBitsTinyexr(b);
        planes[nPix + i] = halfBitsTinyexr(g);
        planes[2 * nPix + i] = halfBitsTinyexr(r);
    }
}

int streamGrid(prt_ctx* ctx, size_t n, int block) {
    size_t need = (n + block - 1) / block, cap = (size_t)ctx->smCount * 8;
    return (int)(need < cap ? (need ? need : 1) : cap);
}

}  // namespace

extern "C" {

int prt_cuda_gbuffer_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, int type, uint32_t seed, float exposure, float* d_rgb) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (!ctx->hasCamera) return setError(ctx, PRT_ERR_STATE, "gbuffer: no camera set");
    if (type < PRT_GBUFFER_DIFFUSE || type > PRT_GBUFFER_NORMAL) return setError(ctx, PRT_ERR_ARG, "gbuffer: unknown type %d", type);
    if (!d_rgb || x1 <= x0 || y1 <= y0 || x1 > ctx->scene.cam.width || y1 > ctx->scene.cam.height) return setError(ctx, PRT_ERR_ARG, "gbuffer: bad region");
    const uint32_t nPix = (x1 - x0) * (y1 - y0);
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    gbufferKernel<<<(nPix + 127) / 128, 128, 0, ctx->stream>>>(ctx->scene, x0, y0, x1 - x0, nPix, type, seed, exposure, d_rgb);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->counters.kernelLaunches++;
    ctx->counters.primaryRays += nPix;
    return PRT_OK;
}

int prt_cuda_gbuffer(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, int type, uint32_t seed, float exposure, float* rgb) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (!rgb || !ctx->hasCamera) return setError(ctx, PRT_ERR_ARG, "gbuffer: bad argument");
    const uint32_t W = ctx->scene.cam.width, H = ctx->scene.cam.height;
    if (x1 <= x0 || y1 <= y0 || x1 > W || y1 > H) return setError(ctx, PRT_ERR_ARG, "gbuffer: bad region");
    if ((r = ensureScratch(ctx, 3, (size_t)W * H * 12))) return r;
    if ((r = prt_cuda_gbuffer_dev(ctx, x0, y0, x1, y1, type, seed, exposure, (float*)ctx->d_scratch[3]))) return r;
    // only the region's pixels cross PCIe; the rest of the caller's image is left alone
    const size_t pitch = (size_t)W * 12, off = ((size_t)y0 * W + x0) * 12;
    PRT_CUDA_CHECK(ctx, cudaMemcpy2DAsync((char*)rgb + off, pitch, (const char*)ctx->d_scratch[3] + off, pitch, (size_t)(x1 - x0) * 12, y1 - y0,
                                          cudaMemcpyDeviceToHost, ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return PRT_OK;
}

int prt_cuda_image_to_ppm_dev(prt_ctx* ctx, const float* d_rgb, uint32_t width, uint32_t height, int tonemap, uint8_t* d_out) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "image_to_ppm: host-only context — no CPU fallback exists");
    DeviceGuard g(ctx->device);
    if (!d_rgb || !d_out || !width || !height) return setError(ctx, PRT_ERR_ARG, "image_to_ppm: bad argument");
    const size_t n = 3 * (size_t)width * height;
    ppmKernel<<<streamGrid(ctx, n, 256), 256, 0, ctx->stream>>>(d_rgb, d_out, n, tonemap);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    ctx->counters.kernelLaunches++;
    return PRT_OK;
}

int prt_cuda_image_to_half_bgr_dev(prt_ctx* ctx, const float* d_rgb, uint32_t width, uint32_t height, uint16_t* d_planes) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "image_to_half_bgr: host-only context — no CPU fallback exists");
    DeviceGuard g(ctx->device);
    if (!d_rgb || !d_planes || !width || !height) return setError(ctx, PRT_ERR_ARG, "image_to_half_bgr: bad argument");
    const size_t n = (size_t)width * height;
    halfKernel<<<streamGrid(ctx, n, 256), 256, 0, ctx->stream>>>(d_rgb, d_planes, n);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    ctx->counters.kernelLaunches++;
    return PRT_OK;
}

int prt_cuda_image_to_ppm(prt_ctx* ctx, const float* rgb, uint32_t width, uint32_t height, int tonemap, uint8_t* out) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "image_to_ppm: host-only context — no CPU fallback exists");
    DeviceGuard g(ctx->device);
    if (!rgb || !out || !width || !height) return setError(ctx, PRT_ERR_ARG, "image_to_ppm: bad argument");
    const size_t n = 3 * (size_t)width * height;
    int r;
    if ((r = ensureScratch(ctx, 3, n * 4))) return r;
    if ((r = ensureScratch(ctx, 1, n))) return r;
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(ctx->d_scratch[3], rgb, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    if ((r = prt_cuda_image_to_ppm_dev(ctx, (const float*)ctx->d_scratch[3], width, height, tonemap, (uint8_t*)ctx->d_scratch[1]))) return r;
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(out, ctx->d_scratch[1], n, cudaMemcpyDeviceToHost, ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return PRT_OK;
}

int prt_cuda_image_to_half_bgr(prt_ctx* ctx, const float* rgb, uint32_t width, uint32_t height, uint16_t* planes) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "image_to_half_bgr: host-only context — no CPU fallback exists");
    DeviceGuard g(ctx->device);
    if (!rgb || !planes || !width || !height) return setError(ctx, PRT_ERR_ARG, "image_to_half_bgr: bad argument");
    const size_t n = 3 * (size_t)width * height;
    int r;
    if ((r = ensureScratch(ctx, 3, n * 4))) return r;
    if ((r = ensureScratch(ctx, 1, n * 2))) return r;
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(ctx->d_scratch[3], rgb, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    if ((r = prt_cuda_image_to_half_bgr_dev(ctx, (const float*)ctx->d_scratch[3], width, height, (uint16_t*)ctx->d_scratch[1]))) return r;
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(planes, ctx->d_scratch[1], n * 2, cudaMemcpyDeviceToHost, ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return PRT_OK;
}

}  // extern "C"
