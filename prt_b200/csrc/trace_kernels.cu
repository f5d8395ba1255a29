// Batch traversal kernels behind prt_cuda_intersect / prt_cuda_occluded / prt_cuda_trace_jobs / prt_cuda_primary_dev.
//
// Persistent threads with dynamic ray fetch (prt_trace.cuh): the grid is sized to the machine (SMs x resident blocks), a lane
// whose ray ended pulls the next one from the batch's queue.  One ray per thread; traversal stacks in shared memory, one
// column per thread and as deep as the scene's deepest tree (round 1 kept them in local memory: 200 LDL/STL in the SASS and
// stack lines competing with node / triangle records for L1); node fetches are 256-bit, triangle fetches 128-bit read-only loads.
//
// Where a ray comes from is a launch parameter (RaySrc): 32-byte records, 12-byte directions / origins with the shared half in
// the parameter block, camera rays generated in place, or shadow rays rebuilt from an earlier batch's hits — the host-buffer
// entry points use this to move 12 or 0 bytes per ray over PCIe instead of 32 (include/prt_cuda.h, prt_ray_layout).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "prt_context.h"
#define PRT_STACK_STRIDE 128   // shared-memory stacks, entry k of thread t at stack[k * 128 + t]
#include "prt_trace.cuh"
#include "prt_wide.cuh"

using namespace prt_b200;

namespace {

constexpr int kBlock = 128;
static_assert(kBlock == PRT_STACK_STRIDE, "one stack column per thread");

struct RaySrc {
    uint32_t layout;          // prt_ray_layout
    uint32_t first;           // PRT_RAYS_PRIMARY: index of this launch's ray 0 in the job's enumeration
    const void* base;         // prt_ray[n] (FULL) or float[3n] (SHARED_*), device
    float shared[4];
    uint32_t x0, y0, bw, perPixel, seed, pad;   // PRT_RAYS_PRIMARY
};

__device__ __forceinline__ void storeHit(const DScene& sc, prt_hit* out, size_t i, const Hit& h) {
    float2* o = reinterpret_cast<float2*>(out + i);
    o[0] = make_float2(h.t, h.i);
    o[1] = make_float2(h.j, h.k);
    o[2] = make_float2(__uint_as_float(hitPrimId(sc, h)), __uint_as_float(h.meshId));
}

__device__ __forceinline__ uint32_t revMaskOf(float ax, float ay, float az) {            // bvh.cpp:444-445
    return (ax < 0.0f ? 1u : 0u) | (ay < 0.0f ? 2u : 0u) | (az < 0.0f ? 4u : 0u);
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

struct RawRay { float ox, oy, oz, dx, dy, dz, maxT; };

// Ray i of a batch.  PACKET (primary rays only): also the sign mask of the packet's avgDir.  The layout is uniform over the
// launch, so the switch does not diverge.
template <bool PACKET>
__device__ __forceinline__ RawRay loadRay(const DScene& sc, const RaySrc& src, uint32_t i, uint32_t& rev) {
    RawRay r;
    if (src.layout == PRT_RAYS_FULL) {
        const float4* p = reinterpret_cast<const float4*>(src.base) + 2 * (size_t)i;
        const float4 a = __ldg(p), b = __ldg(p + 1);
        r.ox = a.x; r.oy = a.y; r.oz = a.z; r.dx = a.w; r.dy = b.x; r.dz = b.y; r.maxT = b.z;
    } else if (src.layout == PRT_RAYS_SHARED_ORIGIN) {
        const float* p = reinterpret_cast<const float*>(src.base) + 3 * (size_t)i;
        r.ox = src.shared[0]; r.oy = src.shared[1]; r.oz = src.shared[2]; r.maxT = src.shared[3];
        r.dx = __ldg(p); r.dy = __ldg(p + 1); r.dz = __ldg(p + 2);
    } else if (src.layout == PRT_RAYS_SHARED_DIR) {
        const float* p = reinterpret_cast<const float*>(src.base) + 3 * (size_t)i;
        r.dx = src.shared[0]; r.dy = src.shared[1]; r.dz = src.shared[2]; r.maxT = src.shared[3];
        r.ox = __ldg(p); r.oy = __ldg(p + 1); r.oz = __ldg(p + 2);
    } else {                                                      // PRT_RAYS_PRIMARY (camera.cpp:12-73)
        const uint32_t ii = src.first + i;
        r.ox = sc.cam.pos[0]; r.oy = sc.cam.pos[1]; r.oz = sc.cam.pos[2]; r.maxT = 100000.0f;
        if (PACKET) {
            // lane l of packet pk uses draws 2l and 2l+1 of the packet's xorshift stream; avgDir = (sequential sum of the 8
            // dirs) / 8 (camera.cpp:40,56).  Every ray recomputes its packet's 8 directions: 16 xorshifts and 8 normalisations
            // against hundreds of instructions of traversal, and the rays of a packet need not sit in one warp.
            const uint32_t pk = ii >> 3, l = ii & 7u;
            const uint32_t pix = pk / src.perPixel, p = pk - pix * src.perPixel;
            const uint32_t px = src.x0 + pix % src.bw, py = src.y0 + pix / src.bw;
            uint32_t s = packetSeedDev(src.seed, py * sc.cam.width + px, src.perPixel, p);
            float ax = 0.0f, ay = 0.0f, az = 0.0f;
            r.dx = r.dy = r.dz = 0.0f;
#pragma unroll
            for (uint32_t k = 0; k < 8; k++) {
                s = xs32(s); const float u0 = rngFloat(s);
                s = xs32(s); const float u1 = rngFloat(s);
                float dx, dy, dz;
                cameraDir(sc.cam, px, py, u0, u1, dx, dy, dz);
                ax = ax + dx; ay = ay + dy; az = az + dz;
                if (k == l) { r.dx = dx; r.dy = dy; r.dz = dz; }
            }
            rev = revMaskOf(ax / 8.0f, ay / 8.0f, az / 8.0f);
        } else {
            const uint32_t pix = ii / src.perPixel, p = ii - pix * src.perPixel;
            const uint32_t px = src.x0 + pix % src.bw, py = src.y0 + pix / src.bw;
            uint32_t s = packetSeedDev(src.seed, py * sc.cam.width + px, src.perPixel, p);
            s = xs32(s); const float u0 = rngFloat(s);
            s = xs32(s); const float u1 = rngFloat(s);
            cameraDir(sc.cam, px, py, u0, u1, r.dx, r.dy, r.dz);
        }
    }
    return r;
}

// The packet's avgDir itself (prt_cuda_primary_dev's optional output; same sums as loadRay<true>)
__device__ __forceinline__ void primaryAvgDir(const DScene& sc, const RaySrc& src, uint32_t pk, float& ax, float& ay, float& az) {
    const uint32_t pix = pk / src.perPixel, p = pk - pix * src.perPixel;
    const uint32_t px = src.x0 + pix % src.bw, py = src.y0 + pix / src.bw;
    uint32_t s = packetSeedDev(src.seed, py * sc.cam.width + px, src.perPixel, p);
    ax = ay = az = 0.0f;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) {
        s = xs32(s); const float u0 = rngFloat(s);
        s = xs32(s); const float u1 = rngFloat(s);
        float dx, dy, dz;
        cameraDir(sc.cam, px, py, u0, u1, dx, dy, dz);
        ax = ax + dx; ay = ay + dy; az = az + dz;
    }
    ax = ax / 8.0f; ay = ay / 8.0f; az = az / 8.0f;
}

// MODE 0: single-ray rules, MODE 1: packet-lane rules (8 consecutive rays share avgDir).  hits may be null (generate only:
// raysOut / avgOut of prt_cuda_primary_dev).
template <int MODE>
__global__ void __launch_bounds__(kBlock, MODE == 0 ? 9 : 8) intersectKernelDyn(DScene sc, RaySrc src, const float* __restrict__ avgDir, prt_hit* __restrict__ hits,
                                                             prt_ray* __restrict__ raysOut, float* __restrict__ avgOut, uint32_t n, uint32_t depthCap, uint32_t* counter) {
    extern __shared__ uint32_t smemStack[];
    uint32_t* stack = smemStack + threadIdx.x;
    float* stackT = reinterpret_cast<float*>(smemStack + (MODE == 1 ? depthCap * kBlock : 0u)) + threadIdx.x;
    auto fetch = [&](uint32_t i, RayD& r, uint32_t& rev, bool& pkt) -> bool {
        const RawRay q = loadRay<MODE == 1>(sc, src, i, rev);
        r = makeRay(q.ox, q.oy, q.oz, q.dx, q.dy, q.dz, q.maxT, MODE == 1);
        pkt = MODE == 1;
        if (MODE == 1 && src.layout != PRT_RAYS_PRIMARY) { const float* av = avgDir + 3 * (size_t)(i >> 3); rev = revMaskOf(__ldg(av), __ldg(av + 1), __ldg(av + 2)); }
        if (raysOut) {
            float4* o = reinterpret_cast<float4*>(raysOut + i);
            o[0] = make_float4(q.ox, q.oy, q.oz, q.dx);
            o[1] = make_float4(q.dy, q.dz, q.maxT, 0.0f);
        }
        if (MODE == 1 && avgOut && (i & 7u) == 0u) {
            float ax, ay, az;
            primaryAvgDir(sc, src, (src.first + i) >> 3, ax, ay, az);
            avgOut[3 * (size_t)(i >> 3)] = ax; avgOut[3 * (size_t)(i >> 3) + 1] = ay; avgOut[3 * (size_t)(i >> 3) + 2] = az;
        }
        return hits != nullptr;
    };
    auto commit = [&](uint32_t i, const Hit& h, bool, uint32_t) { storeHit(sc, hits, i, h); };
    traceQueue<MODE == 1 ? kClosestPacket : kClosestSingle>(sc, n, counter, stack, stackT, fetch, commit);
}

// parentHits != null: PRT_RAYS_FROM_HITS — ray i is the shadow ray of parent ray i (src describes the PARENT job's rays,
// shared[] of `light` the light direction); a parent ray that missed has no shadow ray: out[i] = 2.
template <int MODE, bool FROM_HITS>
__global__ void __launch_bounds__(kBlock) occludedKernelDyn(DScene sc, RaySrc src, const prt_hit* __restrict__ parentHits, float lx, float ly, float lz,
                                                            int parentPacket, uint8_t* __restrict__ out, uint32_t n, uint32_t* counter) {
    extern __shared__ uint32_t smemStack[];
    uint32_t* stack = smemStack + threadIdx.x;
    float dummyT[1];
    const float kFar = 2.0f * sc.radius;                                                                                // path_tracer.cpp:192
    auto fetch = [&](uint32_t i, RayD& r, uint32_t& rev, bool& pkt) -> bool {
        pkt = MODE == 1; rev = 0;
        uint32_t dummyRev;
        if (!FROM_HITS) {
            const RawRay q = loadRay<false>(sc, src, i, dummyRev);
            r = makeRay(q.ox, q.oy, q.oz, q.dx, q.dy, q.dz, q.maxT, MODE == 1);
            return true;
        }
        const float t = __ldg(reinterpret_cast<const float*>(parentHits + i));
        if (t == -1.0f) { out[i] = 2; return false; }
        const RawRay q = parentPacket ? loadRay<true>(sc, src, i, dummyRev) : loadRay<false>(sc, src, i, dummyRev);
        // pos = hit.t*ray.dir + ray.org (path_tracer.cpp:101 / :283); org = pos + kFar*lightDir, dir = -lightDir,
        // maxT = kFar - 0.0008 (path_tracer.cpp:209-212 / :236-238)
        const float px = t * q.dx + q.ox, py = t * q.dy + q.oy, pz = t * q.dz + q.oz;
        r = makeRay(px + kFar * lx, py + kFar * ly, pz + kFar * lz, -lx, -ly, -lz, kFar - 0.0008f, MODE == 1);
        return true;
    };
    auto commit = [&](uint32_t i, const Hit&, bool occ, uint32_t) { out[i] = occ ? 1 : 0; };
    traceQueue<kOcclude>(sc, n, counter, stack, dummyT, fetch, commit);
}

// ---- fast traversal mode (prt_wide.cuh): single-ray-rule batches over the Wide8 trees ----
// MODE as in the exact kernels: which of the reference's two rule sets the RAY is prepared with (dominant axis signed / abs,
// ray.h:26-40 / 58-71) and which root-box test it gets; the wide traversal itself has one rule.
template <int MODE>
__global__ void __launch_bounds__(kBlock) intersectKernelWide(DScene sc, RaySrc src, prt_hit* __restrict__ hits, uint32_t n, uint32_t* counter) {
    extern __shared__ uint2 smemWide[];
    uint2 deep[kWideDeep];
    const WideStack stack{smemWide + threadIdx.x, deep};
    auto fetch = [&](uint32_t i, RayD& r, uint32_t& rev, bool& pkt) -> bool {
        const RawRay q = loadRay<MODE == 1>(sc, src, i, rev);
        r = makeRay(q.ox, q.oy, q.oz, q.dx, q.dy, q.dz, q.maxT, MODE == 1);
        pkt = MODE == 1;
        return true;
    };
    auto commit = [&](uint32_t i, const Hit& h, bool, uint32_t) { storeHit(sc, hits, i, h); };
    traceQueueWide<false, MODE>(sc, n, counter, stack, fetch, commit);
}
template <int MODE>
__global__ void __launch_bounds__(kBlock) occludedKernelWide(DScene sc, RaySrc src, const prt_hit* __restrict__ parentHits, float lx, float ly, float lz,
                                                             int parentPacket, uint8_t* __restrict__ out, uint32_t n, uint32_t* counter) {
    extern __shared__ uint2 smemWide[];
    uint2 deep[kWideDeep];
    const WideStack stack{smemWide + threadIdx.x, deep};
    const float kFar = 2.0f * sc.radius;
    auto fetch = [&](uint32_t i, RayD& r, uint32_t& rev, bool& pkt) -> bool {
        pkt = MODE == 1; rev = 0;
        uint32_t dummyRev;
        if (!parentHits) {
            const RawRay q = loadRay<false>(sc, src, i, dummyRev);
            r = makeRay(q.ox, q.oy, q.oz, q.dx, q.dy, q.dz, q.maxT, MODE == 1);
            return true;
        }
        const float t = __ldg(reinterpret_cast<const float*>(parentHits + i));
        if (t == -1.0f) { out[i] = 2; return false; }
        const RawRay q = parentPacket ? loadRay<true>(sc, src, i, dummyRev) : loadRay<false>(sc, src, i, dummyRev);
        const float px = t * q.dx + q.ox, py = t * q.dy + q.oy, pz = t * q.dz + q.oz;
        r = makeRay(px + kFar * lx, py + kFar * ly, pz + kFar * lz, -lx, -ly, -lz, kFar - 0.0008f, MODE == 1);
        return true;
    };
    auto commit = [&](uint32_t i, const Hit&, bool occ, uint32_t) { out[i] = occ ? 1 : 0; };
    traceQueueWide<true, MODE>(sc, n, counter, stack, fetch, commit);
}
int wideReady(prt_ctx* ctx) {
    uint32_t wd = 0;
    for (auto m : ctx->meshes) wd = std::max(wd, m->fastDepth);
    if (7u * wd + 1u > (uint32_t)(kWideCap + kWideDeep)) return setError(ctx, PRT_ERR_STATE, "fast-mode tree of %u levels exceeds the traversal stack", wd);
    return PRT_OK;
}

uint32_t depthCapOf(prt_ctx* ctx) {
    uint32_t d = 1;
    for (auto m : ctx->meshes) d = std::max(d, m->build->bvh.maxDepth + 1u);
    return d;
}
int gridFor(prt_ctx* ctx, const void* kernel, size_t smemBytes) {
    const void* key = (const void*)((uintptr_t)kernel ^ (smemBytes << 1));
    auto it = ctx->gridCache.find(key);
    if (it != ctx->gridCache.end()) return it->second;
    int perSm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, kBlock, smemBytes) != cudaSuccess || perSm <= 0) perSm = 4;
    return ctx->gridCache[key] = ctx->smCount * perSm;
}

template <typename K, typename... Args>
int launchTimed(prt_ctx* ctx, K kernel, uint32_t n, size_t smemBytes, Args... args) {
    if (!ctx->inPipe) {
        PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counter + ctx->counterSlot, 0, sizeof(uint32_t), ctx->stream));
        PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    }
    int grid = gridFor(ctx, (const void*)kernel, smemBytes);
    uint32_t need = (n + kBlock - 1) / kBlock;
    if ((uint32_t)grid > need) grid = (int)(need ? need : 1);
    if (smemBytes > 48 * 1024) PRT_CUDA_CHECK(ctx, cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
    kernel<<<grid, kBlock, smemBytes, ctx->stream>>>(args...);
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

int launchIntersect(prt_ctx* ctx, uint32_t n, const RaySrc& src, const float* d_avgDir, int mode, prt_hit* d_hits, prt_ray* d_raysOut, float* d_avgOut) {
    const uint32_t depthCap = depthCapOf(ctx);
    const size_t smem = (size_t)depthCap * kBlock * 4u * (mode == PRT_MODE_PACKET ? 2u : 1u);
    uint32_t* counter = ctx->d_counter + ctx->counterSlot;
    if (ctx->traversal == PRT_TRAVERSAL_FAST && d_hits && !d_raysOut) {                                   // opt-in fast mode (prt_wide.cuh)
        int rw = wideReady(ctx);
        if (rw) return rw;
        const size_t wsm = (size_t)kWideCap * kBlock * sizeof(uint2);
        rw = mode == PRT_MODE_PACKET ? launchTimed(ctx, intersectKernelWide<1>, n, wsm, ctx->scene, src, d_hits, n, counter)
                                     : launchTimed(ctx, intersectKernelWide<0>, n, wsm, ctx->scene, src, d_hits, n, counter);
        if (rw) return rw;
        if (src.layout == PRT_RAYS_PRIMARY) ctx->counters.primaryRays += n; else ctx->counters.extendRays += n;
        return PRT_OK;
    }
    int r = mode == PRT_MODE_PACKET ? launchTimed(ctx, intersectKernelDyn<1>, n, smem, ctx->scene, src, d_avgDir, d_hits, d_raysOut, d_avgOut, n, depthCap, counter)
                                    : launchTimed(ctx, intersectKernelDyn<0>, n, smem, ctx->scene, src, d_avgDir, d_hits, d_raysOut, d_avgOut, n, depthCap, counter);
    if (r) return r;
    if (d_hits) { if (src.layout == PRT_RAYS_PRIMARY) ctx->counters.primaryRays += n; else ctx->counters.extendRays += n; }
    return PRT_OK;
}

int launchOccluded(prt_ctx* ctx, uint32_t n, const RaySrc& src, const prt_hit* d_parentHits, const float* light, int parentPacket, int mode, uint8_t* d_out) {
    const size_t smem = (size_t)depthCapOf(ctx) * kBlock * 4u;
    uint32_t* counter = ctx->d_counter + ctx->counterSlot;
    const float lx = light ? light[0] : 0.0f, ly = light ? light[1] : 0.0f, lz = light ? light[2] : 0.0f;
    if (ctx->traversal == PRT_TRAVERSAL_FAST) {                                                          // opt-in fast mode (prt_wide.cuh)
        int rw = wideReady(ctx);
        if (rw) return rw;
        const size_t wsm = (size_t)kWideCap * kBlock * sizeof(uint2);
        rw = mode == PRT_MODE_PACKET ? launchTimed(ctx, occludedKernelWide<1>, n, wsm, ctx->scene, src, d_parentHits, lx, ly, lz, parentPacket, d_out, n, counter)
                                     : launchTimed(ctx, occludedKernelWide<0>, n, wsm, ctx->scene, src, d_parentHits, lx, ly, lz, parentPacket, d_out, n, counter);
        if (rw) return rw;
        ctx->counters.shadowRays += n;
        return PRT_OK;
    }
    int r;
    if (d_parentHits) r = mode == PRT_MODE_PACKET ? launchTimed(ctx, occludedKernelDyn<1, true>, n, smem, ctx->scene, src, d_parentHits, lx, ly, lz, parentPacket, d_out, n, counter)
                                                  : launchTimed(ctx, occludedKernelDyn<0, true>, n, smem, ctx->scene, src, d_parentHits, lx, ly, lz, parentPacket, d_out, n, counter);
    else r = mode == PRT_MODE_PACKET ? launchTimed(ctx, occludedKernelDyn<1, false>, n, smem, ctx->scene, src, d_parentHits, lx, ly, lz, parentPacket, d_out, n, counter)
                                     : launchTimed(ctx, occludedKernelDyn<0, false>, n, smem, ctx->scene, src, d_parentHits, lx, ly, lz, parentPacket, d_out, n, counter);
    if (r) return r;
    ctx->counters.shadowRays += n;
    return PRT_OK;
}

RaySrc fullSrc(const void* d_rays) {
    RaySrc s{};
    s.layout = PRT_RAYS_FULL; s.base = d_rays;
    return s;
}

}  // namespace

extern "C" {

int prt_cuda_intersect_dev(prt_ctx* ctx, uint32_t n, const prt_ray* d_rays, const float* d_avgDir, int mode, prt_hit* d_hits) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!d_rays || !d_hits) return setError(ctx, PRT_ERR_ARG, "intersect: null buffer");
    if (mode != PRT_MODE_PACKET && mode != PRT_MODE_SINGLE) return setError(ctx, PRT_ERR_ARG, "intersect: unknown mode %d", mode);
    if (mode == PRT_MODE_PACKET && ((n & 7u) || !d_avgDir)) return setError(ctx, PRT_ERR_ARG, "intersect: packet mode needs n %% 8 == 0 and avgDir");
    return launchIntersect(ctx, n, fullSrc(d_rays), d_avgDir, mode, d_hits, nullptr, nullptr);
}

int prt_cuda_occluded_dev(prt_ctx* ctx, uint32_t n, const prt_ray* d_rays, int mode, uint8_t* d_out) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (n == 0) return PRT_OK;
    if (!d_rays || !d_out) return setError(ctx, PRT_ERR_ARG, "occluded: null buffer");
    if (mode != PRT_MODE_PACKET && mode != PRT_MODE_SINGLE) return setError(ctx, PRT_ERR_ARG, "occluded: unknown mode %d", mode);
    return launchOccluded(ctx, n, fullSrc(d_rays), nullptr, nullptr, 0, mode, d_out);
}

int prt_cuda_primary_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t ppp, uint32_t seed, int mode,
                         prt_ray* d_rays, float* d_avgDir, prt_hit* d_hits) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (!ctx->hasCamera) return setError(ctx, PRT_ERR_STATE, "primary: no camera set");
    if (x1 <= x0 || y1 <= y0 || x1 > ctx->scene.cam.width || y1 > ctx->scene.cam.height || ppp == 0) return setError(ctx, PRT_ERR_ARG, "primary: bad region");
    if (mode != PRT_MODE_PACKET && mode != PRT_MODE_SINGLE) return setError(ctx, PRT_ERR_ARG, "primary: unknown mode %d", mode);
    uint64_t n64 = (uint64_t)(x1 - x0) * (y1 - y0) * ppp * (mode == PRT_MODE_PACKET ? 8u : 1u);
    if (n64 >= 0xffffff00ull) return setError(ctx, PRT_ERR_ARG, "primary: too many rays for one call");
    RaySrc s{};
    s.layout = PRT_RAYS_PRIMARY; s.x0 = x0; s.y0 = y0; s.bw = x1 - x0; s.perPixel = ppp; s.seed = seed;
    return launchIntersect(ctx, (uint32_t)n64, s, nullptr, mode, d_hits, d_rays, mode == PRT_MODE_PACKET ? d_avgDir : nullptr);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------------
// Host-buffer entry points.  Every batch is cut into chunks; chunk k's upload (copyIn stream), traversal (context stream) and
// download (copyOut stream) overlap with the neighbouring chunks', so a call costs about max(H2D, D2H) over PCIe instead of
// H2D + kernel + D2H.  Results are identical to one big launch (rays are independent).
namespace {

struct PipeScope {                    // chunk launches skip the per-launch counter memset and timing events; restored on every exit path
    prt_ctx* c;
    explicit PipeScope(prt_ctx* ctx) : c(ctx) { c->inPipe = true; }
    ~PipeScope() { c->inPipe = false; c->counterSlot = 0; }
};

bool pipeTrace() {                    // PRT_PIPE_TRACE=1: timing-enabled pipeline events, the call prints its schedule (profiling only)
    static const bool on = getenv("PRT_PIPE_TRACE") && atoi(getenv("PRT_PIPE_TRACE")) != 0;
    return on;
}
int ensurePipe(prt_ctx* ctx, size_t chunks) {
    if (!ctx->copyIn) PRT_CUDA_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->copyIn, cudaStreamNonBlocking));
    if (!ctx->copyOut) PRT_CUDA_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->copyOut, cudaStreamNonBlocking));
    while (ctx->pipeEv.size() < 2 * chunks) {
        cudaEvent_t e;
        PRT_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&e, pipeTrace() ? cudaEventDefault : cudaEventDisableTiming));
        ctx->pipeEv.push_back(e);
    }
    return PRT_OK;
}

uint32_t chunkRays() {
    // ~256 Ki rays per chunk (8 MB up, 6 MB down): measured best on B200 (scripts/e2e_probe.py) — smaller chunks are
    // bound by the host's enqueue rate, larger ones expose the first upload / last download.  Multiple of 8 (packets).
    uint32_t c = 1u << 18;
    if (const char* e = getenv("PRT_CHUNK_RAYS")) { uint32_t v = (uint32_t)atoll(e) & ~7u; if (v >= 8) c = v; }   // profiling override
    return c;
}

uint32_t rayBytes(uint32_t layout) { return layout == PRT_RAYS_FULL ? 32u : (layout == PRT_RAYS_SHARED_ORIGIN || layout == PRT_RAYS_SHARED_DIR) ? 12u : 0u; }

int validateJobs(prt_ctx* ctx, uint32_t count, const prt_trace_job* jobs, uint64_t& totalRays) {
    totalRays = 0;
    for (uint32_t j = 0; j < count; j++) {
        const prt_trace_job& J = jobs[j];
        if (J.kind != PRT_JOB_INTERSECT && J.kind != PRT_JOB_OCCLUDED) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: unknown kind %u", j, J.kind);
        if (J.mode != PRT_MODE_PACKET && J.mode != PRT_MODE_SINGLE) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: unknown mode %d", j, J.mode);
        if (J.layout > PRT_RAYS_FROM_HITS) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: unknown ray layout %u", j, J.layout);
        if (J.n == 0) continue;
        if (!J.out) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: null output buffer", j);
        const bool closest = J.kind == PRT_JOB_INTERSECT;
        if (rayBytes(J.layout) && !J.rays) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: null ray buffer", j);
        if (closest && J.mode == PRT_MODE_PACKET && (J.n & 7u)) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: packet mode needs n %% 8 == 0", j);
        if (closest && J.mode == PRT_MODE_PACKET && rayBytes(J.layout) && !J.avgDir) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: packet mode needs avgDir", j);
        if (J.layout == PRT_RAYS_PRIMARY) {
            if (!closest) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: PRT_RAYS_PRIMARY is for closest-hit jobs", j);
            if (!ctx->hasCamera) return setError(ctx, PRT_ERR_STATE, "trace_jobs: job %u: no camera set", j);
            const uint32_t* R = J.region;
            if (R[2] <= R[0] || R[3] <= R[1] || R[2] > ctx->scene.cam.width || R[3] > ctx->scene.cam.height || J.perPixel == 0) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: bad region", j);
            const uint64_t want = (uint64_t)(R[2] - R[0]) * (R[3] - R[1]) * J.perPixel * (J.mode == PRT_MODE_PACKET ? 8u : 1u);
            if (want != J.n) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: n = %u but the region holds %llu rays", j, J.n, (unsigned long long)want);
        }
        if (J.layout == PRT_RAYS_FROM_HITS) {
            if (closest) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: PRT_RAYS_FROM_HITS is for any-hit jobs", j);
            if (J.fromJob < 0 || (uint32_t)J.fromJob >= j) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: fromJob must name an earlier job", j);
            const prt_trace_job& S = jobs[J.fromJob];
            if (S.kind != PRT_JOB_INTERSECT || S.layout == PRT_RAYS_FROM_HITS || S.n != J.n) return setError(ctx, PRT_ERR_ARG, "trace_jobs: job %u: source job %d must be a closest-hit job with the same n", j, J.fromJob);
        }
        totalRays += J.n;
    }
    return PRT_OK;
}

// One GPU's share of a call: rays [first[j], first[j] + cnt[j]) of every job.
int traceJobsOne(prt_ctx* ctx, uint32_t count, const prt_trace_job* jobs, const uint32_t* first, const uint32_t* cnt) {
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    // Chunk size per job.  What a call costs is the copies back to back plus the tail behind the last one (its traversal and
    // its download), so chunks are twice the single-batch size (fewer stream operations between the uploads: 1.37 -> 1.43
    // Grays/s on the c2 frame, profiles/r1/z3_jobs_probe.log) except in a closing closest-hit job, whose last 24 B/ray
    // download is the tail.  Grown until all jobs together fit the context's 64 work counters.
    uint32_t lastJob = 0;
    for (uint32_t j = 0; j < count; j++) if (cnt[j]) lastJob = j;
    std::vector<uint32_t> chunkOf(count);
    uint32_t base = chunkRays();
    const bool fixedChunk = getenv("PRT_CHUNK_RAYS") != nullptr;         // profiling override: the same size everywhere
    for (;;) {
        size_t c = 0;
        for (uint32_t j = 0; j < count; j++) {
            chunkOf[j] = (fixedChunk || (j == lastJob && jobs[j].kind == PRT_JOB_INTERSECT)) ? base : 2u * base;
            c += (cnt[j] + chunkOf[j] - 1) / chunkOf[j];
        }
        if (c <= 48) break;
        base *= 2;
    }
    // staging layout: rays back to back (16-byte aligned per job); results 256-byte aligned per job; avgDir back to back
    std::vector<size_t> rayOff(count), outOff(count), avgOff(count);
    size_t rayTotal = 0, outTotal = 0, avgTotal = 0, chunks = 0;
    for (uint32_t j = 0; j < count; j++) {
        const prt_trace_job& J = jobs[j];
        rayOff[j] = rayTotal; rayTotal += (((size_t)cnt[j] * rayBytes(J.layout)) + 15) & ~(size_t)15;
        outOff[j] = outTotal; outTotal += (((size_t)cnt[j] * (J.kind == PRT_JOB_INTERSECT ? sizeof(prt_hit) : 1)) + 255) & ~(size_t)255;
        avgOff[j] = avgTotal; if (J.kind == PRT_JOB_INTERSECT && J.mode == PRT_MODE_PACKET && rayBytes(J.layout)) avgTotal += (size_t)(cnt[j] / 8) * 12;
        chunks += (cnt[j] + chunkOf[j] - 1) / chunkOf[j];
    }
    if (chunks == 0) return PRT_OK;
    if (rayTotal && (r = ensureScratch(ctx, 0, rayTotal))) return r;
    if ((r = ensureScratch(ctx, 1, outTotal))) return r;
    if (avgTotal && (r = ensureScratch(ctx, 2, avgTotal))) return r;
    if ((r = ensurePipe(ctx, chunks))) return r;
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));      // scene uploads / earlier work on the context stream
    PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ctx->d_counter, 0, 64 * sizeof(uint32_t), ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    PipeScope scope(ctx);
    // results go back in groups of `dg` chunks: fewer, larger downloads disturb the upload stream less (measured on B200,
    // scripts/e2e_group_ab.sh: 1.73 ms per 2 M-ray call with one download per chunk, 1.69 with one per two, 1.79 per three)
    static const uint32_t dg = getenv("PRT_D2H_GROUP") ? (uint32_t)atoi(getenv("PRT_D2H_GROUP")) : 2u;
    auto srcOf = [&](uint32_t j, uint32_t o) {                   // RaySrc of job j's rays from local index o on
        const prt_trace_job& J = jobs[j];
        RaySrc s{};
        s.layout = J.layout; s.first = first[j] + o;
        s.base = (const char*)ctx->d_scratch[0] + rayOff[j] + (size_t)o * rayBytes(J.layout);
        memcpy(s.shared, J.shared, 16);
        s.x0 = J.region[0]; s.y0 = J.region[1]; s.bw = J.region[2] - J.region[0]; s.perPixel = J.perPixel; s.seed = J.seed;
        return s;
    };
    size_t k = 0;
    for (uint32_t j = 0; j < count; j++) {
        const prt_trace_job& J = jobs[j];
        if (cnt[j] == 0) continue;
        const uint32_t N = cnt[j], rb = rayBytes(J.layout);
        const bool closest = J.kind == PRT_JOB_INTERSECT, packet = closest && J.mode == PRT_MODE_PACKET;
        const char* hRays = rb ? (const char*)J.rays + (size_t)first[j] * rb : nullptr;
        const float* hAvg = (packet && rb) ? J.avgDir + 3 * (size_t)(first[j] / 8) : nullptr;
        char* dR = (char*)ctx->d_scratch[0] + rayOff[j];
        float* dA = hAvg ? (float*)((char*)ctx->d_scratch[2] + avgOff[j]) : nullptr;
        char* dO = (char*)ctx->d_scratch[1] + outOff[j];
        char* hOut = (char*)J.out + (size_t)first[j] * (closest ? sizeof(prt_hit) : 1);
        uint32_t kj = 0;
        const uint32_t chunk = chunkOf[j];
        for (uint32_t o = 0; o < N; o += chunk, k++, kj++) {
            ctx->counterSlot = (uint32_t)k;
            const uint32_t m = N - o < chunk ? N - o : chunk;
            if (rb) {
                PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(dR + (size_t)o * rb, hRays + (size_t)o * rb, (size_t)m * rb, cudaMemcpyHostToDevice, ctx->copyIn));
                if (hAvg) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(dA + 3 * (size_t)(o / 8), hAvg + 3 * (size_t)(o / 8), (size_t)(m / 8) * 12, cudaMemcpyHostToDevice, ctx->copyIn));
                PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k], ctx->copyIn));
                PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pipeEv[2 * k], 0));
            }
            const bool last = o + m >= N;
            if (closest) {
                if ((r = launchIntersect(ctx, m, srcOf(j, o), dA ? dA + 3 * (size_t)(o / 8) : nullptr, J.mode, (prt_hit*)dO + o, nullptr, nullptr))) return r;
                if (dg <= 1u || (kj + 1) % dg == 0 || last) {             // hit records go back in groups of dg chunks
                    const uint32_t g0 = dg <= 1u ? o : (uint32_t)((kj / dg) * dg) * chunk;
                    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k + 1], ctx->stream));
                    PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->copyOut, ctx->pipeEv[2 * k + 1], 0));
                    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((prt_hit*)hOut + g0, (prt_hit*)dO + g0, (size_t)(o + m - g0) * sizeof(prt_hit), cudaMemcpyDeviceToHost, ctx->copyOut));
                }
            } else {
                if (J.layout == PRT_RAYS_FROM_HITS) {
                    const uint32_t sj = (uint32_t)J.fromJob;
                    const prt_hit* pH = (const prt_hit*)((char*)ctx->d_scratch[1] + outOff[sj]) + o;
                    r = launchOccluded(ctx, m, srcOf(sj, o), pH, J.shared, jobs[sj].mode == PRT_MODE_PACKET ? 1 : 0, J.mode, (uint8_t*)dO + o);
                } else r = launchOccluded(ctx, m, srcOf(j, o), nullptr, nullptr, 0, J.mode, (uint8_t*)dO + o);
                if (r) return r;
                if (last) {                                               // one byte per ray: a single download per job
                    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->pipeEv[2 * k + 1], ctx->stream));
                    PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->copyOut, ctx->pipeEv[2 * k + 1], 0));
                    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(hOut, dO, (size_t)N, cudaMemcpyDeviceToHost, ctx->copyOut));
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
        fprintf(stderr, "[prt pipe dev %d] %zu chunks of <= %u rays; ms since the call's first stream op: ", ctx->device, k, 2u * base);
        for (size_t c = 0; c < k; c++) {
            float up = -1.0f, tr = -1.0f;
            if (c + 1 < k && cudaEventElapsedTime(&up, ctx->ev0, ctx->pipeEv[2 * c]) != cudaSuccess) { up = -1.0f; cudaGetLastError(); }
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

}  // namespace

extern "C" {

// Several host-buffer batches through ONE copy pipeline per GPU (see include/prt_cuda.h): every job's rays are staged at
// their own offset of the scratch buffers, chunks of consecutive jobs follow each other on the three streams without a drain
// in between, and the host waits once, at the end.  On a multi-GPU context every job is cut into one contiguous slice per GPU
// (multiples of 8 rays, so packets stay whole and a FROM_HITS job's slice lines up with its source's): each GPU's worker
// thread runs the pipeline on its slice over its own PCIe link, and the results land in the caller's buffers directly.
int prt_cuda_trace_jobs(prt_ctx* ctx, uint32_t count, const prt_trace_job* jobs) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->device < 0) return commitScene(ctx);                // reports the missing device
    if (count == 0) return PRT_OK;
    if (!jobs) return setError(ctx, PRT_ERR_ARG, "trace_jobs: null job list");
    uint64_t totalRays = 0;
    int r = validateJobs(ctx, count, jobs, totalRays);
    if (r) return r;
    if (totalRays == 0) return PRT_OK;
    const uint32_t nd = 1u + (uint32_t)ctx->peers.size();
    std::vector<std::vector<uint32_t>> first(nd, std::vector<uint32_t>(count)), cnt(nd, std::vector<uint32_t>(count));
    for (uint32_t j = 0; j < count; j++) {
        const uint64_t n = jobs[j].n, per = ((n + nd - 1) / nd + 7u) & ~(uint64_t)7u;
        for (uint32_t d = 0; d < nd; d++) {
            const uint64_t a = std::min<uint64_t>(n, per * d), b = std::min<uint64_t>(n, per * (d + 1));
            first[d][j] = (uint32_t)a; cnt[d][j] = (uint32_t)(b - a);
        }
    }
    return forEachDevice(ctx, [&](prt_ctx* c, int d) -> int { return traceJobsOne(c, count, jobs, first[d].data(), cnt[d].data()); });
}

int prt_cuda_intersect(prt_ctx* ctx, uint32_t n, const prt_ray* rays, const float* avgDir, int mode, prt_hit* hits) {
    if (!ctx) return PRT_ERR_ARG;
    if (n && (!rays || !hits)) return setError(ctx, PRT_ERR_ARG, "intersect: null buffer");
    if (mode == PRT_MODE_PACKET && n && ((n & 7u) || !avgDir)) return setError(ctx, PRT_ERR_ARG, "intersect: packet mode needs n %% 8 == 0 and avgDir");
    prt_trace_job J{};
    J.kind = PRT_JOB_INTERSECT; J.n = n; J.mode = mode; J.layout = PRT_RAYS_FULL; J.rays = rays; J.avgDir = mode == PRT_MODE_PACKET ? avgDir : nullptr; J.out = hits;
    if (ctx->device >= 0 && n == 0) { DeviceGuard g(ctx->device); return commitScene(ctx); }
    return prt_cuda_trace_jobs(ctx, 1, &J);
}

int prt_cuda_occluded(prt_ctx* ctx, uint32_t n, const prt_ray* rays, int mode, uint8_t* out) {
    if (!ctx) return PRT_ERR_ARG;
    if (n && (!rays || !out)) return setError(ctx, PRT_ERR_ARG, "occluded: null buffer");
    prt_trace_job J{};
    J.kind = PRT_JOB_OCCLUDED; J.n = n; J.mode = mode; J.layout = PRT_RAYS_FULL; J.rays = rays; J.out = out;
    if (ctx->device >= 0 && n == 0) { DeviceGuard g(ctx->device); return commitScene(ctx); }
    return prt_cuda_trace_jobs(ctx, 1, &J);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------------
// prt_cuda_primary_shadow_noalloc: an entry point that allocates nothing (include/prt_cuda.h; no counterpart in the reference,
// added for callers that cannot call cudaMalloc: CUDA graph capture, serving loops).
//
// Same two kernels as a PRT_RAYS_PRIMARY + PRT_RAYS_FROM_HITS pair of prt_cuda_trace_jobs (intersectKernelDyn<0>,
// occludedKernelDyn<0, true>).  Their scratch — the dynamic-fetch work counters and the closest-hit records that only feed the
// shadow kernel — lives in a file-scope __device__ array instead of the context's cudaMalloc'ed buffers: 256 B of counters +
// 24 B per ray for at most PRT_NOALLOC_MAX_RAYS rays (6 MiB; every GPU that loads the library holds it until exit).
//
// g_noallocScratch is a host-side stand-in in host code, not a device address: the address kernels may use comes from
// cudaGetSymbolAddress, asked once per device with that device current (noallocScratchFor) and passed to the kernels.
// Every call on a device shares the array (whatever the context): calls that overlap on the GPU race.  The counters are
// cleared by every call; the hit records need no clearing (the first kernel writes all n before the second reads them).
namespace {

constexpr size_t kNoallocCounterBytes = 256;
constexpr size_t kNoallocBytes = kNoallocCounterBytes + (size_t)PRT_NOALLOC_MAX_RAYS * sizeof(prt_hit);
static __device__ __align__(256) unsigned char g_noallocScratch[kNoallocBytes];

constexpr int kNoallocMaxDevices = 64;
std::mutex g_noallocLock;
void* g_noallocAddr[kNoallocMaxDevices] = {};

// device address of g_noallocScratch on `device`, which the caller has made current (DeviceGuard)
int noallocScratchFor(prt_ctx* ctx, int device, unsigned char** out) {
    if (device < 0 || device >= kNoallocMaxDevices) return setError(ctx, PRT_ERR_ARG, "noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_noallocLock);
    if (!g_noallocAddr[device]) {
        void* p = nullptr;
        PRT_CUDA_CHECK(ctx, cudaGetSymbolAddress(&p, g_noallocScratch));
        g_noallocAddr[device] = p;
    }
    *out = static_cast<unsigned char*>(g_noallocAddr[device]);
    return PRT_OK;
}

}  // namespace

extern "C" int prt_cuda_primary_shadow_noalloc(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t perPixel, uint32_t seed,
                                               const float lightDir[3], uint8_t* d_out) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "host-only context (device < 0): no CUDA device, and there is no CPU fallback");
    DeviceGuard g(ctx->device);
    // committing a changed scene uploads its tables with cudaMalloc: not here
    if (ctx->dirty) return setError(ctx, PRT_ERR_STATE, "noalloc: the scene has uncommitted changes; run any tracing call once (it uploads the tables) before this one");
    if (!ctx->hasCamera) return setError(ctx, PRT_ERR_STATE, "noalloc: no camera set");
    if (!lightDir || !d_out) return setError(ctx, PRT_ERR_ARG, "noalloc: null buffer");
    if (x1 <= x0 || y1 <= y0 || x1 > ctx->scene.cam.width || y1 > ctx->scene.cam.height || perPixel == 0) return setError(ctx, PRT_ERR_ARG, "noalloc: bad region");
    const uint64_t n64 = (uint64_t)(x1 - x0) * (y1 - y0) * perPixel;
    if (n64 > PRT_NOALLOC_MAX_RAYS)
        return setError(ctx, PRT_ERR_ARG, "noalloc: %llu rays exceed the static scratch of PRT_NOALLOC_MAX_RAYS = %u rays; split the region", (unsigned long long)n64, (unsigned)PRT_NOALLOC_MAX_RAYS);
    const uint32_t n = (uint32_t)n64;
    unsigned char* scratch = nullptr;
    if (int r = noallocScratchFor(ctx, ctx->device, &scratch)) return r;
    uint32_t* counters = reinterpret_cast<uint32_t*>(scratch);
    prt_hit* hits = reinterpret_cast<prt_hit*>(scratch + kNoallocCounterBytes);

    RaySrc s{};
    s.layout = PRT_RAYS_PRIMARY; s.x0 = x0; s.y0 = y0; s.bw = x1 - x0; s.perPixel = perPixel; s.seed = seed;
    const uint32_t depthCap = depthCapOf(ctx);
    const size_t smem = (size_t)depthCap * kBlock * 4u;
    const uint32_t need = (n + kBlock - 1) / kBlock;
    auto gridOf = [&](const void* k) { const int gmax = gridFor(ctx, k, smem); return (uint32_t)gmax > need ? (int)need : gmax; };
    auto kClosest = intersectKernelDyn<0>;
    auto kShadow = occludedKernelDyn<0, true>;
    if (smem > 48 * 1024) {
        PRT_CUDA_CHECK(ctx, cudaFuncSetAttribute((const void*)kClosest, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        PRT_CUDA_CHECK(ctx, cudaFuncSetAttribute((const void*)kShadow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    PRT_CUDA_CHECK(ctx, cudaMemsetAsync(counters, 0, kNoallocCounterBytes, ctx->stream));        // counter 0: closest hit, counter 1: shadow
    kClosest<<<gridOf((const void*)kClosest), kBlock, smem, ctx->stream>>>(ctx->scene, s, nullptr, hits, nullptr, nullptr, n, depthCap, counters);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    kShadow<<<gridOf((const void*)kShadow), kBlock, smem, ctx->stream>>>(ctx->scene, s, hits, lightDir[0], lightDir[1], lightDir[2], 0, d_out, n, counters + 1);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    ctx->counters.kernelLaunches += 2;
    ctx->counters.primaryRays += n;
    ctx->counters.shadowRays += n;
    return PRT_OK;
}
