// Batch traversal kernels behind prt_cuda_intersect / prt_cuda_occluded / prt_cuda_primary_dev.
//
// Persistent threads: the grid is sized to the machine (SMs x resident blocks), and each warp pulls the next
// 32 rays with one atomicAdd, so long rays in one warp do not leave the rest of the machine idle.  One ray per
// thread, traversal stack in (L1-resident) local memory, node/triangle fetches as 128-bit read-only loads.
#include <cstdlib>
#include <vector>
#include "prt_context.h"
#include "prt_trace.cuh"

using namespace prt_b200;

namespace {

constexpr int kBlock = 128;

__device__ __forceinline__ void storeHit(const DScene& sc, prt_hit* out, size_t i, const Hit& h) {
    float2* o = reinterpret_cast<float2*>(out + i);
    o[0] = make_float2(h.t, h.i);
    o[1] = make_float2(h.j, h.k);
    o[2] = make_float2(__uint_as_float(hitPrimId(sc, h)), __uint_as_float(h.meshId));
}

__device__ __forceinline__ uint32_t revMaskOf(float ax, float ay, float az) {            // bvh.cpp:444-445
    return (ax < 0.0f ? 1u : 0u) | (ay < 0.0f ? 2u : 0u) | (az < 0.0f ? 4u : 0u);
}

// MODE 0: single-ray rules, MODE 1: packet-lane rules (8 consecutive rays share avgDir)
template <int MODE>
__global__ void __launch_bounds__(kBlock) intersectKernel(DScene sc, const float4* __restrict__ rays, const float* __restrict__ avgDir,
                                                          prt_hit* __restrict__ hits, uint32_t n, uint32_t* counter) {
    uint32_t stack[kStackSize];
    float stackT[MODE == 1 ? kStackSize : 1];
    const unsigned lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(counter, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        uint32_t i = base + lane;
        if (i < n) {
            float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
            RayD r = makeRay(a.x, a.y, a.z, a.w, b.x, b.y, b.z, MODE == 1);
            Hit h;
            if (MODE == 0) h = sceneIntersectSingle(sc, r, stack);
            else {
                const float* av = avgDir + 3 * (size_t)(i >> 3);
                h = sceneIntersectPacket(sc, r, revMaskOf(__ldg(av), __ldg(av + 1), __ldg(av + 2)), stack, stackT);
            }
            storeHit(sc, hits, i, h);
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) occludedKernel(DScene sc, const float4* __restrict__ rays, uint8_t* __restrict__ out, uint32_t n, uint32_t* counter) {
    uint32_t stack[kStackSize];
    const unsigned lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(counter, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        uint32_t i = base + lane;
        if (i < n) {
            float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
            RayD r = makeRay(a.x, a.y, a.z, a.w, b.x, b.y, b.z, MODE == 1);
            out[i] = sceneOccluded<MODE == 1>(sc, r, stack) ? 1 : 0;
        }
    }
}

// ---- dynamic-fetch variants (prt_trace.cuh) ----
template <int MODE>
__global__ void __launch_bounds__(kBlock) intersectKernelDyn(DScene sc, const float4* __restrict__ rays, const float* __restrict__ avgDir,
                                                             prt_hit* __restrict__ hits, uint32_t n, uint32_t* counter) {
    uint32_t stack[kStackSize];
    float stackT[MODE == 1 ? kStackSize : 1];
    auto fetch = [&](uint32_t i, RayD& r, uint32_t& rev, bool& pkt) {
        float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
        r = makeRay(a.x, a.y, a.z, a.w, b.x, b.y, b.z, MODE == 1);
        pkt = MODE == 1;
        if (MODE == 1) { const float* av = avgDir + 3 * (size_t)(i >> 3); rev = revMaskOf(__ldg(av), __ldg(av + 1), __ldg(av + 2)); }
    };
    auto commit = [&](uint32_t i, const Hit& h, bool, uint32_t) { storeHit(sc, hits, i, h); };
    traceQueue<MODE == 1 ? kClosestPacket : kClosestSingle>(sc, n, counter, stack, stackT, fetch, commit);
}

template <int MODE>
__global__ void __launch_bounds__(kBlock) occludedKernelDyn(DScene sc, const float4* __restrict__ rays, uint8_t* __restrict__ out, uint32_t n, uint32_t* counter) {
    uint32_t stack[kStackSize];
    float dummyT[1];
    auto fetch = [&](uint32_t i, RayD& r, uint32_t& rev, bool& pkt) {
        float4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
        r = makeRay(a.x, a.y, a.z, a.w, b.x, b.y, b.z, MODE == 1);
        pkt = MODE == 1; rev = 0;
    };
    auto commit = [&](uint32_t i, const Hit&, bool occ, uint32_t) { out[i] = occ ? 1 : 0; };
    traceQueue<kOcclude>(sc, n, counter, stack, dummyT, fetch, commit);
}

// xorshift32 (random.h:41-48) and the [0,1) float trick (random.h:23-33)
__device__ __forceinline__ uint32_t xs32(uint32_t x) { x ^= x << 13; x ^= x >> 17; x ^= x << 5; return x; }
__device__ __forceinline__ float rngFloat(uint32_t x) { return __uint_as_float((x & 0x007fffffu) | 0x3f800000u) - 1.0f; }

__device__ __forceinline__ uint32_t packetSeedDev(uint32_t seed, uint32_t pixel, uint32_t ppp, uint32_t packet) {
    uint32_t h = seed * 0x9E3779B9u + pixel * ppp + packet;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h | 1u;
}

// Camera ray for jitter draws (u0,u1) (camera.cpp:20-27 / 50-55)
__device__ __forceinline__ void cameraDir(const DCamera& c, uint32_t x, uint32_t y, float u0, float u1, float& dx, float& dy, float& dz) {
    const float kScreenScale = 0.6f;
    const float kAspect = (float)c.width / (float)c.height;
    const float kScaleX = 0.5f * c.invW, kScaleY = 0.5f * c.invH;
    float jx = (2.0f * u0 - 1.0f) * kScaleX;
    float jy = (2.0f * u1 - 1.0f) * kScaleY;
    float nx = 2.0f * ((float)x * c.invW - 0.5f + jx) * kScreenScale * kAspect;
    float ny = -2.0f * ((float)y * c.invH - 0.5f + jy) * kScreenScale;
    float vx = nx * c.right[0] + ny * c.up[0] + c.dir[0];
    float vy = nx * c.right[1] + ny * c.up[1] + c.dir[1];
    float vz = nx * c.right[2] + ny * c.up[2] + c.dir[2];
    float xx = vx * vx, yy = vy * vy, zz = vz * vz;
    float inv = 1.0f / sqrtf(xx + yy + zz);
    dx = inv * vx; dy = inv * vy; dz = inv * vz;
}

// One thread per primary ray; 8 consecutive lanes are the 8 samples of one packet (camera.cpp:35-73): lane l
// uses draws 2l and 2l+1 of the packet's xorshift stream, avgDir = (sequential sum of the 8 dirs)/8.
template <int MODE>
__global__ void __launch_bounds__(kBlock) primaryKernel(DScene sc, uint32_t x0, uint32_t y0, uint32_t bw, uint32_t bh, uint32_t ppp, uint32_t seed,
                                                        prt_ray* __restrict__ raysOut, float* __restrict__ avgOut, prt_hit* __restrict__ hitsOut,
                                                        uint32_t n, uint32_t* counter) {
    uint32_t stack[kStackSize];
    float stackT[MODE == 1 ? kStackSize : 1];
    const unsigned lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(counter, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        uint32_t i = base + lane;
        bool active = i < n;
        uint32_t ii = active ? i : n - 1;
        Hit h; RayD r;
        if (MODE == 1) {
            uint32_t pk = ii >> 3, l = ii & 7u;
            uint32_t pix = pk / ppp, p = pk - pix * ppp;
            uint32_t px = x0 + pix % bw, py = y0 + pix / bw;
            uint32_t s = packetSeedDev(seed, py * sc.cam.width + px, ppp, p);
            for (uint32_t k = 0; k < 2 * l; k++) s = xs32(s);
            s = xs32(s); float u0 = rngFloat(s);
            s = xs32(s); float u1 = rngFloat(s);
            float dx, dy, dz;
            cameraDir(sc.cam, px, py, u0, u1, dx, dy, dz);
            // avgDir: ((((((d0+d1)+d2)+d3)+d4)+d5)+d6)+d7 starting from 0 (camera.cpp:40,56), then /8
            float ax = 0.0f, ay = 0.0f, az = 0.0f;
            const unsigned gbase = lane & ~7u;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                ax = ax + __shfl_sync(0xffffffffu, dx, gbase + k);
                ay = ay + __shfl_sync(0xffffffffu, dy, gbase + k);
                az = az + __shfl_sync(0xffffffffu, dz, gbase + k);
            }
            ax = ax / 8.0f; ay = ay / 8.0f; az = az / 8.0f;
            r = makeRay(sc.cam.pos[0], sc.cam.pos[1], sc.cam.pos[2], dx, dy, dz, 100000.0f, true);
            if (active && avgOut && l == 0) { avgOut[3 * (size_t)pk] = ax; avgOut[3 * (size_t)pk + 1] = ay; avgOut[3 * (size_t)pk + 2] = az; }
            if (active && hitsOut) h = sceneIntersectPacket(sc, r, revMaskOf(ax, ay, az), stack, stackT);
        } else {
            uint32_t pix = ii / ppp, p = ii - pix * ppp;
            uint32_t px = x0 + pix % bw, py = y0 + pix / bw;
            uint32_t s = packetSeedDev(seed, py * sc.cam.width + px, ppp, p);
            s = xs32(s); float u0 = rngFloat(s);
            s = xs32(s); float u1 = rngFloat(s);
            float dx, dy, dz;
            cameraDir(sc.cam, px, py, u0, u1, dx, dy, dz);
            r = makeRay(sc.cam.pos[0], sc.cam.pos[1], sc.cam.pos[2], dx, dy, dz, 100000.0f, false);
            if (active && hitsOut) h = sceneIntersectSingle(sc, r, stack);
        }
        if (active) {
            if (raysOut) {
                float4* o = reinterpret_cast<float4*>(raysOut + i);
                o[0] = make_float4(r.ox, r.oy, r.oz, r.dx);
                o[1] = make_float4(r.dy, r.dz, r.maxT, 0.0f);
            }
            if (hitsOut) storeHit(sc, hitsOut, i, h);
        }
    }
}

int gridFor(prt_ctx* ctx, const void* kernel) {
    auto it = ctx->gridCache.find(kernel);
    if (it != ctx->gridCache.end()) return it->second;
    int perSm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, kBlock, 0) != cudaSuccess || perSm <= 0) perSm = 4;
    return ctx->gridCache[kernel] = ctx->smCount * perSm;
}

template <typename K, typename... Args>
int launchTimed(prt_ctx* ctx, K kernel, uint32_t n, Args... args) {
    if (!ctx->inPipe) {
        PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counter + ctx->counterSlot, 0, sizeof(uint32_t), ctx->stream));
        PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    }
    int grid = gridFor(ctx, (const void*)kernel);
    uint32_t need = (n + kBlock - 1) / kBlock;
    if ((uint32_t)grid > need) grid = (int)(need ? need : 1);
    kernel<<<grid, kBlock, 0, ctx->stream>>>(args...);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    if (!ctx->inPipe) PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->counters.kernelLaunches++;
    return PRT_OK;
}

int finishTimed(prt_ctx* ctx) {
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0.0f;
    PRT_CUDA_CHECK(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->counters.lastKernelMs = ms;
    return PRT_OK;
}

}  // namespace

extern "C" {

int prt_cuda_intersect_dev(prt_ctx* ctx, uint32_t n, const prt_ray* d_rays, const float* d_avgDir, int mode, prt_hit* d_hits) {
    if (!ctx) return PRT_ERR_ARG;
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!d_rays || !d_hits) return setError(ctx, PRT_ERR_ARG, "intersect: null buffer");
    if (mode == PRT_MODE_PACKET) {
        if ((n & 7u) || !d_avgDir) return setError(ctx, PRT_ERR_ARG, "intersect: packet mode needs n %% 8 == 0 and avgDir");
        r = ctx->traceImpl ? launchTimed(ctx, intersectKernelDyn<1>, n, ctx->scene, (const float4*)d_rays, d_avgDir, d_hits, n, ctx->d_counter + ctx->counterSlot)
                           : launchTimed(ctx, intersectKernel<1>, n, ctx->scene, (const float4*)d_rays, d_avgDir, d_hits, n, ctx->d_counter + ctx->counterSlot);
    } else if (mode == PRT_MODE_SINGLE) {
        r = ctx->traceImpl ? launchTimed(ctx, intersectKernelDyn<0>, n, ctx->scene, (const float4*)d_rays, d_avgDir, d_hits, n, ctx->d_counter + ctx->counterSlot)
                           : launchTimed(ctx, intersectKernel<0>, n, ctx->scene, (const float4*)d_rays, d_avgDir, d_hits, n, ctx->d_counter + ctx->counterSlot);
    } else return setError(ctx, PRT_ERR_ARG, "intersect: unknown mode %d", mode);
    if (r) return r;
    ctx->counters.extendRays += n;
    return PRT_OK;
}

int prt_cuda_occluded_dev(prt_ctx* ctx, uint32_t n, const prt_ray* d_rays, int mode, uint8_t* d_out) {
    if (!ctx) return PRT_ERR_ARG;
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!d_rays || !d_out) return setError(ctx, PRT_ERR_ARG, "occluded: null buffer");
    if (mode == PRT_MODE_PACKET) r = ctx->traceImpl ? launchTimed(ctx, occludedKernelDyn<1>, n, ctx->scene, (const float4*)d_rays, d_out, n, ctx->d_counter + ctx->counterSlot)
                                                      : launchTimed(ctx, occludedKernel<1>, n, ctx->scene, (const float4*)d_rays, d_out, n, ctx->d_counter + ctx->counterSlot);
    else if (mode == PRT_MODE_SINGLE) r = ctx->traceImpl ? launchTimed(ctx, occludedKernelDyn<0>, n, ctx->scene, (const float4*)d_rays, d_out, n, ctx->d_counter + ctx->counterSlot)
                                                         : launchTimed(ctx, occludedKernel<0>, n, ctx->scene, (const float4*)d_rays, d_out, n, ctx->d_counter + ctx->counterSlot);
    else return setError(ctx, PRT_ERR_ARG, "occluded: unknown mode %d", mode);
    if (r) return r;
    ctx->counters.shadowRays += n;
    return PRT_OK;
}

// Host-buffer entry points.  The batch is cut into chunks; chunk k's upload (copyIn stream), traversal (context
// stream) and download (copyOut stream) overlap with the neighbouring chunks', so a call costs about
// max(H2D, D2H) over PCIe instead of H2D + kernel + D2H.  Results are identical to one big launch (rays are independent).
struct PipeScope {                    // chunk launches skip the per-launch counter memset and timing events; restored on every exit path
    prt_ctx* c;
    explicit PipeScope(prt_ctx* ctx) : c(ctx) { c->inPipe = true; }
    ~PipeScope() { c->inPipe = false; c->counterSlot = 0; }
};

static bool pipeTrace() {             // PRT_PIPE_TRACE=1: timing-enabled pipeline events, prt_cuda_trace_jobs prints its schedule (profiling only)
    static const bool on = getenv("PRT_PIPE_TRACE") && atoi(getenv("PRT_PIPE_TRACE")) != 0;
    return on;
}
static int ensurePipe(prt_ctx* ctx, size_t chunks) {
    if (!ctx->copyIn) PRT_CUDA_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->copyIn, cudaStreamNonBlocking));
    if (!ctx->copyOut) PRT_CUDA_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->copyOut, cudaStreamNonBlocking));
    while (ctx->pipeEv.size() < 2 * chunks) {
        cudaEvent_t e;
        PRT_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&e, pipeTrace() ? cudaEventDefault : cudaEventDisableTiming));
        ctx->pipeEv.push_back(e);
    }
    return PRT_OK;
}

// Zero-copy input: when the caller's buffer is page-locked host memory (cudaHostAlloc / cudaHostRegister, which UVA maps into
// the device's address space), the traversal kernels read the rays straight out of it over PCIe — the fetch of a ray is one
// coalesced 32-byte read per lane, issued by whichever lanes just finished, and thousands of them are in flight — instead of
// waiting for a staging copy.  PRT_ZC=1: rays (and avgDir) zero-copy, results staged; 2: results written to the caller's pinned
// buffer directly as well.  Measured on B200 and OFF by default (profiles/r1/zc_probe.log): SM-issued reads of host memory reach
// 42 - 45 GB/s where the copy engine reaches 54.5, so a 2 M-ray call takes 2.04 - 2.25 ms instead of 1.77, and direct 24-byte
// result writes over PCIe take it to 3.1 - 3.9 ms.  Kept as a profiling knob only.
static int zeroCopyMode() {
    static const int m = getenv("PRT_ZC") ? atoi(getenv("PRT_ZC")) : 0;
    return m;
}
static void* mappedDevPtrV(const void* host) {
    if (!host) return nullptr;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
    return a.devicePointer;
}
#define mappedDevPtr(p) ((decltype(p))mappedDevPtrV((const void*)(p)))

static uint32_t chunkRays(uint32_t n) {
    // ~256 Ki rays per chunk (8 MB up, 6 MB down): measured best on B200 (scripts/e2e_probe.py) — smaller chunks are
    // bound by the host's enqueue rate, larger ones expose the first upload / last download.  At most 48 chunks (the
    // context has 64 work counters, zeroed once per call); multiple of 8 (packets).
    uint32_t c = 1u << 18;
    if (const char* e = getenv("PRT_CHUNK_RAYS")) { uint32_t v = (uint32_t)atoll(e) & ~7u; if (v >= 8) c = v; }   // profiling override
    if ((n + c - 1) / c > 48u) c = ((n / 48u) + 8u) & ~7u;
    return c;
}

int prt_cuda_intersect(prt_ctx* ctx, uint32_t n, const prt_ray* rays, const float* avgDir, int mode, prt_hit* hits) {
    if (!ctx) return PRT_ERR_ARG;
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!rays || !hits) return setError(ctx, PRT_ERR_ARG, "intersect: null buffer");
    if (mode != PRT_MODE_PACKET && mode != PRT_MODE_SINGLE) return setError(ctx, PRT_ERR_ARG, "intersect: unknown mode %d", mode);
    if (mode == PRT_MODE_PACKET && ((n & 7u) || !avgDir)) return setError(ctx, PRT_ERR_ARG, "intersect: packet mode needs n %% 8 == 0 and avgDir");
    const int zc = zeroCopyMode();
    const prt_ray* mR = zc >= 1 ? mappedDevPtr(rays) : nullptr;
    const float* mA = (mR && mode == PRT_MODE_PACKET) ? mappedDevPtr(avgDir) : nullptr;
    if (mode == PRT_MODE_PACKET && !mA) mR = nullptr;
    prt_hit* mH = (mR && zc >= 2) ? mappedDevPtr(hits) : nullptr;
    if (!mR && (r = ensureScratch(ctx, 0, (size_t)n * sizeof(prt_ray)))) return r;
    if (!mH && (r = ensureScratch(ctx, 1, (size_t)n * sizeof(prt_hit)))) return r;
    if (!mR && mode == PRT_MODE_PACKET && (r = ensureScratch(ctx, 2, (size_t)(n / 8) * 12))) return r;
    const uint32_t chunk = chunkRays(n);
    const size_t chunks = (n + chunk - 1) / chunk;
    if ((r = ensurePipe(ctx, chunks))) return r;
    const prt_ray* dR = mR ? mR : (const prt_ray*)ctx->d_scratch[0];
    prt_hit* dH = mH ? mH : (prt_hit*)ctx->d_scratch[1];
    const float* dA = mR ? mA : (const float*)ctx->d_scratch[2];
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));      // scene uploads / earlier work on the context stream
    PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counter, 0, 64 * sizeof(uint32_t), ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    PipeScope scope(ctx);
    size_t k = 0;
    for (uint32_t o = 0; o < n; o += chunk, k++) {
        ctx->counterSlot = (uint32_t)k;
        const uint32_t m = n - o < chunk ? n - o : chunk;
        if (!mR) {
            PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((prt_ray*)dR + o, rays + o, (size_t)m * sizeof(prt_ray), cudaMemcpyHostToDevice, ctx->copyIn));
            if (mode == PRT_MODE_PACKET)
                PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((float*)dA + 3 * (size_t)(o / 8), avgDir + 3 * (size_t)(o / 8), (size_t)(m / 8) * 12, cudaMemcpyHostToDevice, ctx->copyIn));
            PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k], ctx->copyIn));
            PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipeEv[2 * k], 0));
        }
        if ((r = prt_cuda_intersect_dev(ctx, m, dR + o, mode == PRT_MODE_PACKET ? dA + 3 * (size_t)(o / 8) : nullptr, mode, dH + o))) return r;
        if (mH) continue;
        // results go back in groups of `dg` chunks: fewer, larger downloads disturb the upload stream less (measured on B200,
        // scripts/e2e_group_ab.sh: 1.73 ms per 2 M-ray call with one download per chunk, 1.69 with one per two, 1.79 per three)
        static const uint32_t dg = getenv("PRT_D2H_GROUP") ? (uint32_t)atoi(getenv("PRT_D2H_GROUP")) : 2u;
        const bool last = o + m >= n;
        if (dg <= 1u || (k + 1) % dg == 0 || last) {
            const uint32_t g0 = dg <= 1u ? o : (uint32_t)((k / dg) * dg) * chunk;
            PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k + 1], ctx->stream));
            PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->copyOut, ctx->pipeEv[2 * k + 1], 0));
            PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(hits + g0, dH + g0, (size_t)(o + m - g0) * sizeof(prt_hit), cudaMemcpyDeviceToHost, ctx->copyOut));
        }
    }
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    if (!mH) PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->copyOut));
    return finishTimed(ctx);
}

int prt_cuda_occluded(prt_ctx* ctx, uint32_t n, const prt_ray* rays, int mode, uint8_t* out) {
    if (!ctx) return PRT_ERR_ARG;
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!rays || !out) return setError(ctx, PRT_ERR_ARG, "occluded: null buffer");
    if (mode != PRT_MODE_PACKET && mode != PRT_MODE_SINGLE) return setError(ctx, PRT_ERR_ARG, "occluded: unknown mode %d", mode);
    const prt_ray* mR = zeroCopyMode() >= 1 ? mappedDevPtr(rays) : nullptr;
    if (!mR && (r = ensureScratch(ctx, 0, (size_t)n * sizeof(prt_ray)))) return r;
    if ((r = ensureScratch(ctx, 1, (size_t)n))) return r;
    const uint32_t chunk = mR ? n : chunkRays(n);       // nothing to overlap without staging copies: one launch
    const size_t chunks = (n + chunk - 1) / chunk;
    if ((r = ensurePipe(ctx, chunks))) return r;
    const prt_ray* dR = mR ? mR : (const prt_ray*)ctx->d_scratch[0];
    uint8_t* dO = (uint8_t*)ctx->d_scratch[1];
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counter, 0, 64 * sizeof(uint32_t), ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    PipeScope scope(ctx);
    size_t k = 0;
    for (uint32_t o = 0; o < n; o += chunk, k++) {
        ctx->counterSlot = (uint32_t)k;
        const uint32_t m = n - o < chunk ? n - o : chunk;
        if (!mR) {
            PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((prt_ray*)dR + o, rays + o, (size_t)m * sizeof(prt_ray), cudaMemcpyHostToDevice, ctx->copyIn));
            PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k], ctx->copyIn));
            PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipeEv[2 * k], 0));
        }
        if ((r = prt_cuda_occluded_dev(ctx, m, dR + o, mode, dO + o))) return r;
    }
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    // the result is one byte per ray: a single download at the end
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(out, dO, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    return finishTimed(ctx);
}

// Several host-buffer batches through ONE copy pipeline (see include/prt_cuda.h): every job's rays are staged at their own
// offset of the scratch buffers, chunks of consecutive jobs follow each other on the three streams without a drain in between,
// and the host waits once, at the end.
int prt_cuda_trace_jobs(prt_ctx* ctx, uint32_t count, const prt_trace_job* jobs) {
    if (!ctx) return PRT_ERR_ARG;
    int r = commitScene(ctx);
    if (r) return r;
    if (count == 0) return PRT_OK;
    if (!jobs) return setError(ctx, PRT_ERR_ARG, "trace_jobs: null job list");
    uint64_t totalRays = 0;
    for (uint32_t j = 0; j < count; j++) {
        const prt_trace_job& J = jobs[j];
        if (J.kind != PRT_JOB_INTERSECT && J.kind != PRT_JOB_OCCLUDED) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: unknown kind %u", j, J.kind);
        if (J.mode != PRT_MODE_PACKET && J.mode != PRT_MODE_SINGLE) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: unknown mode %d", j, J.mode);
        if (J.n == 0) continue;
        if (!J.rays || !J.out) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: null buffer", j);
        if (J.kind == PRT_JOB_INTERSECT && J.mode == PRT_MODE_PACKET && ((J.n & 7u) || !J.avgDir))
            return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: packet mode needs n %% 8 == 0 and avgDir", j);
        totalRays += J.n;
    }
    if (totalRays == 0) return PRT_OK;
    // Chunk size per job.  What a call costs is the uploads back to back plus the tail behind the last one (its traversal and
    // its download), so chunks are twice the single-batch size (fewer stream operations between the uploads: 1.37 -> 1.43
    // Grays/s on the c2 frame, profiles/r1/z3_jobs_probe.log) except in a closing closest-hit job, whose last 24 B/ray
    // download is the tail.  Grown until all jobs together fit the context's work counters.
    uint32_t lastJob = 0;
    for (uint32_t j = 0; j < count; j++) if (jobs[j].n) lastJob = j;
    std::vector<uint32_t> chunkOf(count);
    uint32_t base = chunkRays(0);
    const bool fixedChunk = getenv("PRT_CHUNK_RAYS") != nullptr;         // profiling override: the same size everywhere
    for (;;) {
        size_t c = 0;
        for (uint32_t j = 0; j < count; j++) {
            chunkOf[j] = (fixedChunk || (j == lastJob && jobs[j].kind == PRT_JOB_INTERSECT)) ? base : 2u * base;
            c += (jobs[j].n + chunkOf[j] - 1) / chunkOf[j];
        }
        if (c <= 48) break;
        base *= 2;
    }
    // staging layout: rays back to back; results 256-byte aligned per job; avgDir back to back
    std::vector<size_t> rayOff(count), outOff(count), avgOff(count);
    size_t rayBytes = 0, outBytes = 0, avgBytes = 0, chunks = 0;
    for (uint32_t j = 0; j < count; j++) {
        const prt_trace_job& J = jobs[j];
        rayOff[j] = rayBytes; rayBytes += (size_t)J.n * sizeof(prt_ray);
        outOff[j] = outBytes; outBytes += (((size_t)J.n * (J.kind == PRT_JOB_INTERSECT ? sizeof(prt_hit) : 1)) + 255) & ~(size_t)255;
        avgOff[j] = avgBytes; if (J.kind == PRT_JOB_INTERSECT && J.mode == PRT_MODE_PACKET) avgBytes += (size_t)(J.n / 8) * 12;
        chunks += (J.n + chunkOf[j] - 1) / chunkOf[j];
    }
    if ((r = ensureScratch(ctx, 0, rayBytes))) return r;
    if ((r = ensureScratch(ctx, 1, outBytes))) return r;
    if (avgBytes && (r = ensureScratch(ctx, 2, avgBytes))) return r;
    if ((r = ensurePipe(ctx, chunks))) return r;
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));      // scene uploads / earlier work on the context stream
    PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counter, 0, 64 * sizeof(uint32_t), ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    PipeScope scope(ctx);
    static const uint32_t dg = getenv("PRT_D2H_GROUP") ? (uint32_t)atoi(getenv("PRT_D2H_GROUP")) : 2u;
    size_t k = 0;
    for (uint32_t j = 0; j < count; j++) {
        const prt_trace_job& J = jobs[j];
        if (J.n == 0) continue;
        const bool closest = J.kind == PRT_JOB_INTERSECT, packet = closest && J.mode == PRT_MODE_PACKET;
        prt_ray* dR = (prt_ray*)((char*)ctx->d_scratch[0] + rayOff[j]);
        float* dA = packet ? (float*)((char*)ctx->d_scratch[2] + avgOff[j]) : nullptr;
        char* dO = (char*)ctx->d_scratch[1] + outOff[j];
        uint32_t kj = 0;
        const uint32_t chunk = chunkOf[j];
        for (uint32_t o = 0; o < J.n; o += chunk, k++, kj++) {
            ctx->counterSlot = (uint32_t)k;
            const uint32_t m = J.n - o < chunk ? J.n - o : chunk;
            PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(dR + o, J.rays + o, (size_t)m * sizeof(prt_ray), cudaMemcpyHostToDevice, ctx->copyIn));
            if (packet)
                PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(dA + 3 * (size_t)(o / 8), J.avgDir + 3 * (size_t)(o / 8), (size_t)(m / 8) * 12, cudaMemcpyHostToDevice, ctx->copyIn));
            PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k], ctx->copyIn));
            PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipeEv[2 * k], 0));
            const bool last = o + m >= J.n;
            if (closest) {
                if ((r = prt_cuda_intersect_dev(ctx, m, dR + o, packet ? dA + 3 * (size_t)(o / 8) : nullptr, J.mode, (prt_hit*)dO + o))) return r;
                if (dg <= 1u || (kj + 1) % dg == 0 || last) {             // hit records go back in groups of dg chunks (see prt_cuda_intersect)
                    const uint32_t g0 = dg <= 1u ? o : (uint32_t)((kj / dg) * dg) * chunk;
                    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k + 1], ctx->stream));
                    PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->copyOut, ctx->pipeEv[2 * k + 1], 0));
                    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((prt_hit*)J.out + g0, (prt_hit*)dO + g0, (size_t)(o + m - g0) * sizeof(prt_hit), cudaMemcpyDeviceToHost, ctx->copyOut));
                }
            } else {
                if ((r = prt_cuda_occluded_dev(ctx, m, dR + o, J.mode, (uint8_t*)dO + o))) return r;
                if (last) {                                               // one byte per ray: a single download per job
                    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k + 1], ctx->stream));
                    PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->copyOut, ctx->pipeEv[2 * k + 1], 0));
                    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(J.out, dO, (size_t)J.n, cudaMemcpyDeviceToHost, ctx->copyOut));
                }
            }
        }
    }
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    if (pipeTrace()) PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * (k - 1)], ctx->copyOut));   // end of the last download (slot re-used)
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->copyOut));
    if (pipeTrace()) {
        PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->copyIn));
        PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        fprintf(stderr, "[prt pipe] %zu chunks of <= %u rays; ms since the call's first stream op: ", k, 2u * base);
        for (size_t c = 0; c < k; c++) {
            float up = -1.0f, tr = -1.0f;
            if (c + 1 < k) cudaEventElapsedTime(&up, ctx->ev0, ctx->pipeEv[2 * c]);
            if (cudaEventElapsedTime(&tr, ctx->ev0, ctx->pipeEv[2 * c + 1]) != cudaSuccess) { tr = -1.0f; cudaGetLastError(); }
            fprintf(stderr, "[%zu up %.3f traced %.3f] ", c, up, tr);
        }
        float end = -1.0f, kend = -1.0f;
        cudaEventElapsedTime(&end, ctx->ev0, ctx->pipeEv[2 * (k - 1)]);
        cudaEventElapsedTime(&kend, ctx->ev0, ctx->ev1);
        fprintf(stderr, "last kernel %.3f last download %.3f\n", kend, end);
    }
    return finishTimed(ctx);
}

int prt_cuda_primary_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t ppp, uint32_t seed, int mode,
                         prt_ray* d_rays, float* d_avgDir, prt_hit* d_hits) {
    if (!ctx) return PRT_ERR_ARG;
    int r = commitScene(ctx);
    if (r) return r;
    if (!ctx->hasCamera) return setError(ctx, PRT_ERR_STATE, "primary: no camera set");
    if (x1 <= x0 || y1 <= y0 || x1 > ctx->scene.cam.width || y1 > ctx->scene.cam.height || ppp == 0) return setError(ctx, PRT_ERR_ARG, "primary: bad region");
    uint64_t n64 = (uint64_t)(x1 - x0) * (y1 - y0) * ppp * (mode == PRT_MODE_PACKET ? 8u : 1u);
    if (n64 >= 0xffffff00ull) return setError(ctx, PRT_ERR_ARG, "primary: too many rays for one call");
    uint32_t n = (uint32_t)n64;
    if (mode == PRT_MODE_PACKET)
        r = launchTimed(ctx, primaryKernel<1>, n, ctx->scene, x0, y0, x1 - x0, y1 - y0, ppp, seed, d_rays, d_avgDir, d_hits, n, ctx->d_counter + ctx->counterSlot);
    else if (mode == PRT_MODE_SINGLE)
        r = launchTimed(ctx, primaryKernel<0>, n, ctx->scene, x0, y0, x1 - x0, y1 - y0, ppp, seed, d_rays, d_avgDir, d_hits, n, ctx->d_counter + ctx->counterSlot);
    else return setError(ctx, PRT_ERR_ARG, "primary: unknown mode %d", mode);
    if (r) return r;
    ctx->counters.primaryRays += n;
    return PRT_OK;
}

}  // extern "C"
