// C ABI (include/prt_cuda.h): context lifetime, scene upload (replicated over the context's GPUs), counters, plumbing.
// Compiled for sm_100a with -fmad=false (see prt_device.cuh for the arithmetic contract).
#include "prt_context.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <cstdlib>

using namespace prt_b200;

static std::string g_createError;

namespace prt_b200 {

int setError(prt_ctx* ctx, int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) {
        ctx->lastError = buf;
        if (ctx->leader) ctx->leader->lastError = buf;           // a peer's failure is reported through the context the caller holds
    } else g_createError = buf;
    return code;
}

int ensureScratch(prt_ctx* ctx, int slot, size_t bytes) {
    if (ctx->scratchBytes[slot] >= bytes) return PRT_OK;
    if (ctx->d_scratch[slot]) cudaFree(ctx->d_scratch[slot]);
    ctx->d_scratch[slot] = nullptr; ctx->scratchBytes[slot] = 0;
    size_t cap = bytes + bytes / 4;
    PRT_CUDA_CHECK(ctx, cudaMalloc(&ctx->d_scratch[slot], cap));
    ctx->scratchBytes[slot] = cap;
    return PRT_OK;
}

// ---- worker thread of one additional GPU ----
DeviceWorker::DeviceWorker(int device) { th_ = std::thread([this, device] { loop(device); }); }
DeviceWorker::~DeviceWorker() {
    { std::lock_guard<std::mutex> l(mu_); quit_ = true; }
    cv_.notify_all();
    if (th_.joinable()) th_.join();
}
void DeviceWorker::post(std::function<int()> fn) {
    { std::lock_guard<std::mutex> l(mu_); task_ = std::move(fn); hasTask_ = true; busy_ = true; }
    cv_.notify_all();
}
int DeviceWorker::wait() {
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [this] { return !busy_; });
    return result_;
}
void DeviceWorker::loop(int device) {
    cudaSetDevice(device);
    for (;;) {
        std::function<int()> fn;
        {
            std::unique_lock<std::mutex> l(mu_);
            cv_.wait(l, [this] { return hasTask_ || quit_; });
            if (quit_ && !hasTask_) return;
            fn = std::move(task_); hasTask_ = false;
        }
        int r = fn();
        { std::lock_guard<std::mutex> l(mu_); result_ = r; busy_ = false; }
        cv_.notify_all();
    }
}

int forEachDevice(prt_ctx* ctx, const std::function<int(prt_ctx*, int)>& fn) {
    for (size_t k = 0; k < ctx->peers.size(); k++) {
        prt_ctx* p = ctx->peers[k];
        ctx->workers[k]->post([&fn, p, k] { return fn(p, (int)k + 1); });
    }
    int r = fn(ctx, 0);
    for (size_t k = 0; k < ctx->peers.size(); k++) { int rk = ctx->workers[k]->wait(); if (!r) r = rk; }
    return r;
}

template <typename T>
static int uploadVec(prt_ctx* ctx, std::vector<void*>& owner, const T* src, size_t count, const T** out) {
    *out = nullptr;
    if (count == 0) return PRT_OK;
    void* d = nullptr;
    PRT_CUDA_CHECK(ctx, cudaMalloc(&d, count * sizeof(T)));
    owner.push_back(d);
    PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(d, src, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    *out = (const T*)d;
    return PRT_OK;
}

// Node records arrive with mesh-local references: interior = node index, leaf = kLeafBit | (count-1) << 28 | first triangle.
__global__ void rebaseRefs(float4* nodes, uint32_t count, uint32_t nodeBase, uint32_t triBase) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float4 w = nodes[4 * (size_t)i + 3];
    uint32_t r0 = __float_as_uint(w.x), r1 = __float_as_uint(w.y);
    r0 += (r0 & kLeafBit) ? triBase : nodeBase;
    r1 += (r1 & kLeafBit) ? triBase : nodeBase;
    w.x = __uint_as_float(r0); w.y = __uint_as_float(r1);
    nodes[4 * (size_t)i + 3] = w;
}
__global__ void rebaseIndices(uint32_t* v, uint32_t count, uint32_t base) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) v[i] += base;
}
// Wide8 records (24 words): words 16..23 are child references (unused slots repeat child 0's)
__global__ void rebaseWideRefs(uint32_t* nodes, uint32_t count, uint32_t nodeBase, uint32_t triBase) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count * 8u) return;
    uint32_t* r = nodes + 24 * (size_t)(i >> 3) + 16 + (i & 7u);
    const uint32_t v = *r;
    if (v != 0xffffffffu) *r = v + ((v & kLeafBit) ? triBase : nodeBase);
}
// fast-mode triangle records: vertex 0's w = slot of the same triangle in the exact array
__global__ void rebaseFastTris(float4* tris, uint32_t count, uint32_t triBase) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float4 a = tris[3 * (size_t)i];
    a.w = __uint_as_float(__float_as_uint(a.w) + triBase);
    tris[3 * (size_t)i] = a;
}

// Appends the meshes added since the last commit to the scene-wide node / triangle arrays.
static int mergeGeometry(prt_ctx* ctx) {
    if (ctx->mergedMeshes == ctx->meshes.size()) return PRT_OK;
    size_t nodes = ctx->mergedNodes, tris = ctx->mergedTris;
    bool uv = ctx->d_allTriUv != nullptr, shade = ctx->d_allTriShade != nullptr;
    for (size_t k = ctx->mergedMeshes; k < ctx->meshes.size(); k++) {
        nodes += ctx->meshes[k]->nodeCount; tris += ctx->meshes[k]->primCount; uv |= ctx->meshes[k]->d_triUv != nullptr; shade |= ctx->meshes[k]->d_triShade != nullptr;
    }
    if (tris >= (1u << 28) || nodes >= (1u << 31)) return setError(ctx, PRT_ERR_STATE, "scene too large for 28-bit triangle / 31-bit node references (%zu triangles, %zu nodes)", tris, nodes);
    void *nn = nullptr, *nt = nullptr, *nu = nullptr, *ns = nullptr, *np = nullptr;
    if (nodes) PRT_CUDA_CHECK(ctx, cudaMalloc(&nn, nodes * 64));
    if (tris) PRT_CUDA_CHECK(ctx, cudaMalloc(&nt, tris * 48));
    if (tris) PRT_CUDA_CHECK(ctx, cudaMalloc(&np, tris * 4));
    if (uv && tris) { PRT_CUDA_CHECK(ctx, cudaMalloc(&nu, tris * 24)); PRT_CUDA_CHECK(ctx, cudaMemsetAsync(nu, 0, tris * 24, ctx->stream)); }
    if (shade && tris) { PRT_CUDA_CHECK(ctx, cudaMalloc(&ns, tris * 64)); PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ns, 0, tris * 64, ctx->stream)); }   // zero flags: geometric normal, default uv
    if (ctx->mergedTris && ctx->d_allTriShade) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(ns, ctx->d_allTriShade, ctx->mergedTris * 64, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->mergedNodes) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nn, ctx->d_allNodes, ctx->mergedNodes * 64, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->mergedTris) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nt, ctx->d_allTris, ctx->mergedTris * 48, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->mergedTris) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(np, ctx->d_allPrimToTri, ctx->mergedTris * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->mergedTris && ctx->d_allTriUv) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nu, ctx->d_allTriUv, ctx->mergedTris * 24, cudaMemcpyDeviceToDevice, ctx->stream));
    // fast traversal mode: the Wide8 records, the fast trees' triangle records and their uv side records, merged the same way
    void *nfn = nullptr, *nft = nullptr, *nfu = nullptr;
    size_t fastNodes = ctx->mergedFastNodes;
    if (ctx->buildFast) {
        for (size_t k = ctx->mergedMeshes; k < ctx->meshes.size(); k++) fastNodes += ctx->meshes[k]->fastNodeCount;
        if (fastNodes) PRT_CUDA_CHECK(ctx, cudaMalloc(&nfn, fastNodes * 96));
        if (tris) PRT_CUDA_CHECK(ctx, cudaMalloc(&nft, tris * 48));
        if (uv && tris) { PRT_CUDA_CHECK(ctx, cudaMalloc(&nfu, tris * 24)); PRT_CUDA_CHECK(ctx, cudaMemsetAsync(nfu, 0, tris * 24, ctx->stream)); }
        if (ctx->mergedFastNodes) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nfn, ctx->d_allFastNodes, ctx->mergedFastNodes * 96, cudaMemcpyDeviceToDevice, ctx->stream));
        if (ctx->mergedTris) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nft, ctx->d_allFastTris, ctx->mergedTris * 48, cudaMemcpyDeviceToDevice, ctx->stream));
        if (ctx->mergedTris && ctx->d_allFastTriUv) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nfu, ctx->d_allFastTriUv, ctx->mergedTris * 24, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    size_t nb = ctx->mergedNodes, tb = ctx->mergedTris, fnb = ctx->mergedFastNodes;
    for (size_t k = ctx->mergedMeshes; k < ctx->meshes.size(); k++) {
        HostMesh* hm = ctx->meshes[k];
        const size_t nc = hm->nodeCount, tc = hm->primCount;
        if (ctx->buildFast) {
            const size_t fc = hm->fastNodeCount;
            if (fc) {
                PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nfn + fnb * 96, hm->d_fastNodes, fc * 96, cudaMemcpyDeviceToDevice, ctx->stream));
                rebaseWideRefs<<<(unsigned)((fc * 8 + 255) / 256), 256, 0, ctx->stream>>>((uint32_t*)nfn + 24 * fnb, (uint32_t)fc, (uint32_t)fnb, (uint32_t)tb);
                PRT_CUDA_CHECK(ctx, cudaGetLastError());
            }
            if (tc) {
                PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nft + tb * 48, hm->d_fastTris, tc * 48, cudaMemcpyDeviceToDevice, ctx->stream));
                rebaseFastTris<<<(unsigned)((tc + 255) / 256), 256, 0, ctx->stream>>>((float4*)nft + 3 * tb, (uint32_t)tc, (uint32_t)tb);
                PRT_CUDA_CHECK(ctx, cudaGetLastError());
            }
            if (tc && hm->d_fastTriUv) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nfu + tb * 24, hm->d_fastTriUv, tc * 24, cudaMemcpyDeviceToDevice, ctx->stream));
            hm->d.fastRootRef = hm->fastRootRef + (uint32_t)((hm->fastRootRef & kLeafBit) ? tb : fnb);
            fnb += fc;
        }
        if (nc) {
            PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nn + nb * 64, hm->d_nodes, nc * 64, cudaMemcpyDeviceToDevice, ctx->stream));
            rebaseRefs<<<(unsigned)((nc + 255) / 256), 256, 0, ctx->stream>>>((float4*)nn + 4 * nb, (uint32_t)nc, (uint32_t)nb, (uint32_t)tb);
            PRT_CUDA_CHECK(ctx, cudaGetLastError());
        }
        if (tc) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nt + tb * 48, hm->d_tris, tc * 48, cudaMemcpyDeviceToDevice, ctx->stream));
        if (tc) {
            PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)np + tb * 4, hm->d_primToTri, tc * 4, cudaMemcpyDeviceToDevice, ctx->stream));
            rebaseIndices<<<(unsigned)((tc + 255) / 256), 256, 0, ctx->stream>>>((uint32_t*)np + tb, (uint32_t)tc, (uint32_t)tb);
            PRT_CUDA_CHECK(ctx, cudaGetLastError());
        }
        if (tc && hm->d_triUv) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nu + tb * 24, hm->d_triUv, tc * 24, cudaMemcpyDeviceToDevice, ctx->stream));
        if (tc && hm->d_triShade) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)ns + tb * 64, hm->d_triShade, tc * 64, cudaMemcpyDeviceToDevice, ctx->stream));
        hm->d.rootRef = hm->build->bvh.rootRef + (uint32_t)((hm->build->bvh.rootRef & kLeafBit) ? tb : nb);
        hm->d.primBase = (uint32_t)tb;
        nb += nc; tb += tc;
    }
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t k = ctx->mergedMeshes; k < ctx->meshes.size(); k++) {
        HostMesh* hm = ctx->meshes[k];
        cudaFree(hm->d_nodes); cudaFree(hm->d_tris); cudaFree(hm->d_triUv); cudaFree(hm->d_triShade); cudaFree(hm->d_primToTri);
        cudaFree(hm->d_fastNodes); cudaFree(hm->d_fastTris); cudaFree(hm->d_fastTriUv);
        hm->d_nodes = hm->d_tris = hm->d_triUv = hm->d_triShade = hm->d_primToTri = nullptr;
        hm->d_fastNodes = hm->d_fastTris = hm->d_fastTriUv = nullptr;
    }
    if (ctx->buildFast) {
        cudaFree(ctx->d_allFastNodes); cudaFree(ctx->d_allFastTris); cudaFree(ctx->d_allFastTriUv);
        ctx->d_allFastNodes = nfn; ctx->d_allFastTris = nft; ctx->d_allFastTriUv = nfu;
        ctx->mergedFastNodes = fastNodes;
        ctx->scene.fastNodes = (const float4*)nfn; ctx->scene.fastTris = (const float4*)nft; ctx->scene.fastTriUv = (const float2*)nfu;
    }
    cudaFree(ctx->d_allNodes); cudaFree(ctx->d_allTris); cudaFree(ctx->d_allTriUv); cudaFree(ctx->d_allTriShade); cudaFree(ctx->d_allPrimToTri);
    ctx->d_allNodes = nn; ctx->d_allTris = nt; ctx->d_allTriUv = nu; ctx->d_allTriShade = ns; ctx->d_allPrimToTri = np;
    ctx->mergedNodes = nodes; ctx->mergedTris = tris; ctx->mergedMeshes = ctx->meshes.size();
    ctx->scene.nodes = (const float4*)nn; ctx->scene.tris = (const float4*)nt; ctx->scene.triUv = (const float2*)nu; ctx->scene.triShade = (const float4*)ns;
    ctx->scene.primToTri = (const uint32_t*)np;
    return PRT_OK;
}

// profiles/r2/b3_wide_yield_ab.log: 0 (never) / 1 / 8 / 12 / 16 / 20 -> C4 812 / 1183 / 1187 / 1186 / 1175 / 1137, C5 889 / 1088 / 1039 / 1089 / 1078 / 1030, C3 1125 / 1335 / 1351 / 1349 / 1328 / 1292 Msamples/s
constexpr uint32_t kWideYieldDefault = 12u;

int commitScene(prt_ctx* ctx) {
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "host-only context (device < 0): no CUDA device, and there is no CPU fallback");
    if (!ctx->dirty) return PRT_OK;
    if (int mr = mergeGeometry(ctx)) return mr;
    std::vector<DMesh> dm;
    for (auto m : ctx->meshes) dm.push_back(m->d);
    auto reup = [&](void** d, const void* src, size_t bytes) -> int {
        if (*d) { cudaFree(*d); *d = nullptr; }
        if (bytes == 0) return PRT_OK;
        PRT_CUDA_CHECK(ctx, cudaMalloc(d, bytes));
        PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(*d, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        return PRT_OK;
    };
    int r;
    if ((r = reup(&ctx->d_meshes, dm.data(), dm.size() * sizeof(DMesh)))) return r;
    if ((r = reup(&ctx->d_materials, ctx->materials.data(), ctx->materials.size() * sizeof(DMaterial)))) return r;
    if ((r = reup(&ctx->d_textures, ctx->textures.data(), ctx->textures.size() * sizeof(DTexture)))) return r;
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));   // dm is a stack temporary
    ctx->scene.meshes = (const DMesh*)ctx->d_meshes;
    ctx->scene.materials = (const DMaterial*)ctx->d_materials;
    ctx->scene.textures = (const DTexture*)ctx->d_textures;
    ctx->scene.meshCount = (uint32_t)ctx->meshes.size();
    ctx->scene.stats = ctx->d_stats;
    if (!ctx->traceFlagsForced) {
        // node loop yields to the leaf phase (prt_trace.cuh) on big scenes: measured +8 % on C4 (262 k triangles), +12 % on
        // C5 (4.9 M), -2 % on C3 (16 k), 0 on C2 (82 k) — gpurun A/B, profiles/README.md
        size_t prims = 0;
        for (auto m : ctx->meshes) prims += m->primCount;
        ctx->scene.traceFlags = prims >= 200000 ? 1u : 0u;
        // any-hit rays visit the children front to back along the split axis (the boolean does not depend on the order; measured
        // against the reference's child0-first: C2 occlusion 5.7 -> 6.1 Grays/s, C3 +2.6 %, back to front -15 %)
        ctx->scene.traceFlags |= 1u << 1;
    }
    if (!((ctx->scene.traceFlags >> 8) & 0xffu)) ctx->scene.traceFlags |= 12u << 8;      // bits 8-15: idle lanes before a warp refills from the queue
    // bits 16-23: when the fast mode's node loop hands over to the leaf phase (prt_wide.cuh); PRT_WIDE_YIELD overrides for A/B runs
    if (!((ctx->scene.traceFlags >> 16) & 0xffu)) {
        const char* wy = getenv("PRT_WIDE_YIELD");
        ctx->scene.traceFlags |= ((wy ? (uint32_t)atoi(wy) : kWideYieldDefault) & 0xffu) << 16;
    }
    ctx->dirty = false;
    return PRT_OK;
}

int foldStatCounters(prt_ctx* ctx) {
    if (!ctx->d_stats) return PRT_OK;
    unsigned long long h[15];
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaMemcpy(h, ctx->d_stats, sizeof(h), cudaMemcpyDeviceToHost));
    PRT_CUDA_CHECK(ctx, cudaMemset(ctx->d_stats, 0, sizeof(h)));
    for (int c = 0; c < 3; c++) {
        ctx->counters.statRays[c] += h[0 * 3 + c]; ctx->counters.statNodeSteps[c] += h[1 * 3 + c]; ctx->counters.statLeafVisits[c] += h[2 * 3 + c];
        ctx->counters.statTriTests[c] += h[3 * 3 + c]; ctx->counters.statUvLoads[c] += h[4 * 3 + c];
    }
    return PRT_OK;
}

// One GPU's half of prt_cuda_create: stream, events, work counters.
static int initDevice(prt_ctx* ctx) {
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1);
    if (e == cudaSuccess) e = cudaMalloc((void**)&ctx->d_counter, 64 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&ctx->smCount, cudaDevAttrMultiProcessorCount, ctx->device);
#ifdef PRT_STATS
    if (e == cudaSuccess) e = cudaMalloc((void**)&ctx->d_stats, 15 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(ctx->d_stats, 0, 15 * sizeof(unsigned long long));
#endif
    if (e != cudaSuccess) {
        setError(nullptr, PRT_ERR_CUDA, "prt_cuda_create(device=%d): %s — no CPU fallback exists", ctx->device, cudaGetErrorString(e));
        cudaGetLastError();
        return PRT_ERR_CUDA;
    }
    return PRT_OK;
}

static prt_ctx* newContext(int device) {
    auto ctx = new prt_ctx;
    ctx->device = device;
    for (int k = 0; k < 3; k++) { ctx->bboxLo[k] = FLT_MAX; ctx->bboxHi[k] = -FLT_MAX; }
    ctx->scene.radius = FLT_MAX;                                                              // scene.cpp:16
    if (const char* e = getenv("PRT_TRACE_IMPL")) ctx->traceImpl = atoi(e);                   // A/B switch for profiling only
    ctx->scene.traceFlags = 0u;
    if (const char* e = getenv("PRT_TRACE_FLAGS")) { ctx->scene.traceFlags = (uint32_t)atoi(e); ctx->traceFlagsForced = true; }  // A/B switch for profiling only
    return ctx;
}

static void destroyOne(prt_ctx* ctx) {
    if (ctx->device >= 0) {
        DeviceGuard g(ctx->device);
        if (ctx->stream) cudaStreamSynchronize(ctx->stream);
        destroyWavefront(ctx);
        for (auto m : ctx->meshes) { cudaFree(m->d_nodes); cudaFree(m->d_tris); cudaFree(m->d_triUv); cudaFree(m->d_triShade); cudaFree(m->d_primToTri);
                                     cudaFree(m->d_fastNodes); cudaFree(m->d_fastTris); cudaFree(m->d_fastTriUv); }
        cudaFree(ctx->d_allFastNodes); cudaFree(ctx->d_allFastTris); cudaFree(ctx->d_allFastTriUv);
        cudaFree(ctx->d_allNodes); cudaFree(ctx->d_allTris); cudaFree(ctx->d_allTriUv); cudaFree(ctx->d_allTriShade); cudaFree(ctx->d_allPrimToTri);
        for (auto p : ctx->deviceAllocs) cudaFree(p);
        for (auto p : ctx->lightAllocs) cudaFree(p);
        for (auto p : ctx->d_scratch) if (p) cudaFree(p);
        if (ctx->d_meshes) cudaFree(ctx->d_meshes);
        if (ctx->d_materials) cudaFree(ctx->d_materials);
        if (ctx->d_textures) cudaFree(ctx->d_textures);
        if (ctx->d_counter) cudaFree(ctx->d_counter);
        if (ctx->d_stats) cudaFree(ctx->d_stats);
        if (ctx->d_bandLocal) cudaFree(ctx->d_bandLocal);
        if (ctx->d_bandAll) cudaFree(ctx->d_bandAll);
        if (ctx->ev0) cudaEventDestroy(ctx->ev0);
        if (ctx->ev1) cudaEventDestroy(ctx->ev1);
        if (ctx->evG0) cudaEventDestroy(ctx->evG0);
        if (ctx->evG1) cudaEventDestroy(ctx->evG1);
        for (auto e : ctx->pipeEv) cudaEventDestroy(e);
        for (auto& ls : ctx->laneStreams) if (ls) cudaStreamDestroy(ls);
        if (ctx->evLane) cudaEventDestroy(ctx->evLane);
        if (ctx->copyIn) cudaStreamDestroy(ctx->copyIn);
        if (ctx->copyOut) cudaStreamDestroy(ctx->copyOut);
        if (ctx->stream) cudaStreamDestroy(ctx->stream);
    }
    for (auto m : ctx->meshes) delete m;
    delete ctx;
}

}  // namespace prt_b200

// ------------------------------------------------------------------------------------------------------------
extern "C" {

int prt_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int prt_cuda_create(int device, prt_ctx** out) {
    if (!out) return PRT_ERR_ARG;
    *out = nullptr;
    prt_ctx* ctx = newContext(device);
    if (device >= 0) {
        DeviceGuard g(device);
        if (initDevice(ctx)) { delete ctx; return PRT_ERR_CUDA; }
    }
    *out = ctx;
    return PRT_OK;
}

int prt_cuda_create_multi(const int* devices, int ndev, prt_ctx** out) {
    if (!out) return PRT_ERR_ARG;
    *out = nullptr;
    if (!devices || ndev < 1) { setError(nullptr, PRT_ERR_ARG, "prt_cuda_create_multi: need at least one device"); return PRT_ERR_ARG; }
    for (int a = 0; a < ndev; a++) for (int b = 0; b < a; b++) if (devices[a] == devices[b] || devices[a] < 0) { setError(nullptr, PRT_ERR_ARG, "prt_cuda_create_multi: devices must be distinct and >= 0"); return PRT_ERR_ARG; }
    prt_ctx* ctx = nullptr;
    int r = prt_cuda_create(devices[0], &ctx);
    if (r) return r;
    for (int k = 1; k < ndev; k++) {
        prt_ctx* p = newContext(devices[k]);
        p->leader = ctx; p->rank = k;
        { DeviceGuard g(devices[k]); r = initDevice(p); }
        if (r) { delete p; prt_cuda_destroy(ctx); return r; }
        ctx->peers.push_back(p);
        ctx->workers.push_back(new DeviceWorker(devices[k]));
    }
    if (ndev > 1) {
        if ((r = initCommAll(ctx, devices, ndev))) { g_createError = ctx->lastError; prt_cuda_destroy(ctx); return r; }
    }
    *out = ctx;
    return PRT_OK;
}

void prt_cuda_destroy(prt_ctx* ctx) {
    if (!ctx) return;
    destroyMulti(ctx);                                            // communicators first (they hold streams), then workers
    for (auto p : ctx->peers) destroyOne(p);
    ctx->peers.clear();
    destroyOne(ctx);
}

const char* prt_cuda_last_error(const prt_ctx* ctx) { return ctx ? ctx->lastError.c_str() : g_createError.c_str(); }

int prt_cuda_host_alloc(size_t bytes, void** out) {
    if (!out) return PRT_ERR_ARG;
    *out = nullptr;
    if (cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { setError(nullptr, PRT_ERR_CUDA, "prt_cuda_host_alloc(%zu): %s", bytes, cudaGetErrorString(cudaGetLastError())); return PRT_ERR_CUDA; }
    return PRT_OK;
}
int prt_cuda_host_free(void* p) {
    if (!p) return PRT_OK;
    if (cudaFreeHost(p) != cudaSuccess) { setError(nullptr, PRT_ERR_CUDA, "prt_cuda_host_free: %s", cudaGetErrorString(cudaGetLastError())); return PRT_ERR_CUDA; }
    return PRT_OK;
}
int prt_cuda_host_register(void* p, size_t bytes) {
    if (!p || !bytes) return PRT_ERR_ARG;
    if (cudaHostRegister(p, bytes, cudaHostRegisterPortable) != cudaSuccess) { setError(nullptr, PRT_ERR_CUDA, "prt_cuda_host_register: %s", cudaGetErrorString(cudaGetLastError())); return PRT_ERR_CUDA; }
    return PRT_OK;
}
int prt_cuda_host_unregister(void* p) {
    if (!p) return PRT_ERR_ARG;
    if (cudaHostUnregister(p) != cudaSuccess) { setError(nullptr, PRT_ERR_CUDA, "prt_cuda_host_unregister: %s", cudaGetErrorString(cudaGetLastError())); return PRT_ERR_CUDA; }
    return PRT_OK;
}

int prt_cuda_set_option(prt_ctx* ctx, int option, int value) {
    if (!ctx) return PRT_ERR_ARG;
    if (option == PRT_OPTION_BUILD_FAST_TRAVERSAL) {
        if (!ctx->meshes.empty() && (value != 0) != ctx->buildFast) return setError(ctx, PRT_ERR_STATE, "set_option: PRT_OPTION_BUILD_FAST_TRAVERSAL must be set before the first mesh is added");
        ctx->buildFast = value != 0;
        for (auto p : ctx->peers) p->buildFast = ctx->buildFast;
        if (!ctx->buildFast) { ctx->traversal = PRT_TRAVERSAL_EXACT; for (auto p : ctx->peers) p->traversal = PRT_TRAVERSAL_EXACT; }
        return PRT_OK;
    }
    if (option == PRT_OPTION_TRAVERSAL) {
        if (value != PRT_TRAVERSAL_EXACT && value != PRT_TRAVERSAL_FAST) return setError(ctx, PRT_ERR_ARG, "set_option: unknown traversal mode %d", value);
        if (value == PRT_TRAVERSAL_FAST && !ctx->buildFast) return setError(ctx, PRT_ERR_STATE, "set_option: PRT_TRAVERSAL_FAST needs PRT_OPTION_BUILD_FAST_TRAVERSAL set before the meshes were added");
        ctx->traversal = value;
        for (auto p : ctx->peers) p->traversal = value;
        return PRT_OK;
    }
    return setError(ctx, PRT_ERR_ARG, "set_option: unknown option %d", option);
}
int prt_cuda_get_option(const prt_ctx* ctx, int option, int* value) {
    if (!ctx || !value) return PRT_ERR_ARG;
    if (option == PRT_OPTION_BUILD_FAST_TRAVERSAL) { *value = ctx->buildFast ? 1 : 0; return PRT_OK; }
    if (option == PRT_OPTION_TRAVERSAL) { *value = ctx->traversal; return PRT_OK; }
    return PRT_ERR_ARG;
}

int prt_cuda_scene_add_texture(prt_ctx* ctx, int width, int height, int components, int channel, const void* texels) {
    if (!ctx || !texels || width <= 0 || height <= 0 || components <= 0) return setError(ctx, PRT_ERR_ARG, "add_texture: bad argument");
    const size_t bytes = (size_t)width * height * components * (channel ? 4 : 1);
    // alpha-test requirement, as Texture::isAlphaTestRequired (texture.cpp:338-350): any alpha < 255 in an RGBA8 map
    bool alpha = false;
    if (channel == 0 && components == 4) {
        const uint8_t* p = (const uint8_t*)texels;
        for (size_t i = 0; i < (size_t)width * height; i++) if (p[4 * i + 3] < 255) { alpha = true; break; }
    }
    int r = forEachDevice(ctx, [&](prt_ctx* c, int) -> int {
        DeviceGuard g(c->device);
        DTexture t{};
        t.width = width; t.height = height; t.comp = components; t.channel = channel;
        if (c->device >= 0) {
            const uint8_t* d = nullptr;
            int ru = uploadVec(c, c->deviceAllocs, (const uint8_t*)texels, bytes, &d);
            if (ru) return ru;
            PRT_CUDA_CHECK(c, cudaStreamSynchronize(c->stream));
            t.data = d;
        }
        c->textures.push_back(t);
        c->textureNeedsAlpha.push_back(alpha ? 1 : 0);
        c->dirty = true;
        return PRT_OK;
    });
    if (r) return r;
    return (int)ctx->textures.size() - 1;
}

int prt_cuda_scene_add_mesh(prt_ctx* ctx, prt_mesh_desc* md) {
    if (!ctx || !md || !md->indices || !md->positions || !md->primMaterial || !md->materials || md->primCount == 0 || md->vertexCount == 0 || md->materialCount == 0)
        return setError(ctx, PRT_ERR_ARG, "add_mesh: bad argument");
    // ---- validate everything before any context state changes (a rejected mesh must not move the scene's bounds / radius) ----
    const uint32_t P = md->primCount;
    for (size_t i = 0; i < 3 * (size_t)P; i++)
        if (md->indices[i] >= md->vertexCount) return setError(ctx, PRT_ERR_ARG, "add_mesh: indices[%zu] = %u is outside the %u vertices", i, md->indices[i], md->vertexCount);
    for (uint32_t i = 0; i < P; i++)      // the reference indexes its material table unchecked (mesh.cpp:362); e.g. OBJ faces before any usemtl carry id -1
        if (md->primMaterial[i] >= md->materialCount) return setError(ctx, PRT_ERR_ARG, "add_mesh: primMaterial[%u] = %u is outside the %u materials", i, md->primMaterial[i], md->materialCount);
    auto mb = std::make_shared<MeshBuild>();
    mb->primCount = P; mb->vertexCount = md->vertexCount;
    const uint32_t materialBase = (uint32_t)ctx->materials.size();
    for (uint32_t i = 0; i < md->materialCount; i++) {
        const prt_material& s = md->materials[i];
        DMaterial m{};
        memcpy(m.diffuse, s.diffuse, 12); memcpy(m.emissive, s.emissive, 12);
        m.refl = s.reflType; m.diffuseTex = s.diffuseTex; m.bumpTex = s.bumpTex;
        if (s.diffuseTex >= (int)ctx->textures.size() || s.bumpTex >= (int)ctx->textures.size()) return setError(ctx, PRT_ERR_ARG, "add_mesh: texture id out of range");
        m.alphaTest = (s.diffuseTex >= 0 && ctx->textureNeedsAlpha[s.diffuseTex]) ? 1u : 0u;   // material.cpp:77
        mb->anyAlpha |= m.alphaTest != 0;
        mb->materials.push_back(m);
    }
    if (mb->anyAlpha && !md->texcoords) return setError(ctx, PRT_ERR_ARG, "add_mesh: alpha-tested material needs texcoords");

    buildBvh(mb->bvh, md->indices, P, md->positions, 0);
    if (mb->bvh.maxDepth >= (uint32_t)kStackSize) return setError(ctx, PRT_ERR_STATE, "add_mesh: BVH depth %u exceeds the traversal stack (%d), as the reference asserts (bvh.cpp:552)", mb->bvh.maxDepth, kStackSize);

    // Mesh::calculateBounds (mesh.cpp:300-309)
    for (int k = 0; k < 3; k++) { mb->lower[k] = FLT_MAX; mb->upper[k] = -FLT_MAX; }
    for (uint32_t v = 0; v < md->vertexCount; v++)
        for (int k = 0; k < 3; k++) { mb->lower[k] = std::fmin(mb->lower[k], md->positions[3 * v + k]); mb->upper[k] = std::fmax(mb->upper[k], md->positions[3 * v + k]); }

    // Leaf-ordered triangle records: tri[i] = primitive remap[i] (the reference copies the same positions into
    // its TriangleVector SoA, bvh.cpp:256-270).
    mb->tris.resize(3 * (size_t)P);             // 48-byte records; 64-byte records read with two 256-bit loads measured 1-2 % slower (bigger footprint)
    if (mb->anyAlpha) mb->triUv.resize(3 * (size_t)P);
    mb->primToTri.resize(P);
    // De-indexed shading records (DScene::triShade): the vertex normals / texcoords Mesh::getSurfaceProperties gathers through
    // indices[3*prim + k] (mesh.cpp:318-356), copied per leaf-ordered triangle.
    if (md->normals || md->texcoords) mb->shade.resize(4 * (size_t)P);
    for (uint32_t i = 0; i < P; i++) {
        uint32_t prim = mb->bvh.remap[i];
        mb->primToTri[prim] = i;
        uint32_t mat = md->primMaterial[prim];
        const DMaterial& dmat = mb->materials[mat];
        for (int j = 0; j < 3; j++) {
            uint32_t v = md->indices[3 * (size_t)prim + j];
            float4 f;
            f.x = md->positions[3 * (size_t)v]; f.y = md->positions[3 * (size_t)v + 1]; f.z = md->positions[3 * (size_t)v + 2];
            uint32_t w = j == 0 ? prim : (j == 1 ? (dmat.alphaTest ? (uint32_t)dmat.diffuseTex + 1u : 0u) : materialBase + mat);
            memcpy(&f.w, &w, 4);
            mb->tris[3 * (size_t)i + j] = f;
            if (mb->anyAlpha) mb->triUv[3 * (size_t)i + j] = make_float2(md->texcoords[2 * (size_t)v], md->texcoords[2 * (size_t)v + 1]);
        }
        if (!mb->shade.empty()) {
            float n[3][3] = {}, t[3][2] = {};
            for (int j = 0; j < 3; j++) {
                uint32_t v = md->indices[3 * (size_t)prim + j];
                if (md->normals) memcpy(n[j], md->normals + 3 * (size_t)v, 12);
                if (md->texcoords) memcpy(t[j], md->texcoords + 2 * (size_t)v, 8);
            }
            const uint32_t flags = (md->normals ? 1u : 0u) | (md->texcoords ? 2u : 0u);
            float ff; memcpy(&ff, &flags, 4);
            float4* o = &mb->shade[4 * (size_t)i];
            o[0] = make_float4(n[0][0], n[0][1], n[0][2], t[0][0]);
            o[1] = make_float4(n[1][0], n[1][1], n[1][2], t[0][1]);
            o[2] = make_float4(n[2][0], n[2][1], n[2][2], t[1][0]);
            o[3] = make_float4(t[1][1], t[2][0], t[2][1], ff);
        }
    }

    // ---- fast traversal mode: a second tree with leaves of <= 4 triangles, collapsed to Wide8 records; its triangle records are
    // the same 48 bytes in ITS leaf order, vertex 0's w naming the triangle's slot in the exact array (what a hit carries) ----
    if (ctx->buildFast) {
        BvhHost fb;
        buildBvh(fb, md->indices, P, md->positions, 0, 4);
        buildWide(mb->wide, fb);
        // Triangles the reference's single-ray any-hit traversal can never reach: Bvh::occluded tests every node's box with the
        // strict bool test (bvh.cpp:576-657, vecmath.h:1449-1466: "minT1 > maxT0"), which a box of zero extent along an axis
        // fails for every ray (t0 == t1 on that axis).  Axis-aligned walls hang under such boxes, so this is not a grazing
        // corner case but whole surfaces that cast no shadow through that path; the fast tree has other boxes, so the
        // triangles carry the verdict themselves (bit 31 of vertex 1's w, read by triTest<true, true>).
        std::vector<uint8_t> hiddenSlot(P, 0);
        {
            const std::vector<Node32>& nd = mb->bvh.nodes32;
            std::vector<std::pair<uint32_t, bool>> todo;
            if (!nd.empty()) todo.emplace_back(0u, false);
            while (!todo.empty()) {
                const auto [ni, inherited] = todo.back();
                todo.pop_back();
                const Node32& n = nd[ni];
                const bool flat = inherited || n.lower[0] == n.upper[0] || n.lower[1] == n.upper[1] || n.lower[2] == n.upper[2];
                if (n.primCount() == 0xfu) { todo.emplace_back(ni + 1u, flat); todo.emplace_back(n.primOrSecond, flat); }
                else if (flat) for (uint32_t k = 0; k < n.primCount(); k++) hiddenSlot[n.primOrSecond + k] = 1;
            }
        }
        mb->fastTris.resize(3 * (size_t)P);
        if (mb->anyAlpha) mb->fastTriUv.resize(3 * (size_t)P);
        for (uint32_t i = 0; i < P; i++) {
            const uint32_t prim = fb.remap[i], slot = mb->primToTri[prim];
            for (int j = 0; j < 3; j++) {
                mb->fastTris[3 * (size_t)i + j] = mb->tris[3 * (size_t)slot + j];
                if (mb->anyAlpha) mb->fastTriUv[3 * (size_t)i + j] = mb->triUv[3 * (size_t)slot + j];
            }
            memcpy(&mb->fastTris[3 * (size_t)i].w, &slot, 4);
            if (hiddenSlot[slot]) {
                uint32_t w1; memcpy(&w1, &mb->fastTris[3 * (size_t)i + 1].w, 4);
                w1 |= 0x80000000u;
                memcpy(&mb->fastTris[3 * (size_t)i + 1].w, &w1, 4);
            }
        }
    }

    // ---- per GPU: upload first, then commit the context state (nothing is left half-changed when an upload fails) ----
    const uint32_t meshId = (uint32_t)ctx->meshes.size();
    int r = forEachDevice(ctx, [&](prt_ctx* c, int) -> int {
        DeviceGuard g(c->device);
        auto hm = new HostMesh;
        hm->build = mb;
        hm->primCount = P; hm->vertexCount = md->vertexCount; hm->nodeCount = (uint32_t)mb->bvh.nodes64.size();
        hm->fastNodeCount = (uint32_t)mb->wide.nodes.size(); hm->fastRootRef = mb->wide.rootRef; hm->fastDepth = mb->wide.maxDepth;
        DMesh& d = hm->d;
        d.rootRef = mb->bvh.rootRef; d.primBase = 0; d.materialBase = materialBase;
        memcpy(d.rootLo, mb->bvh.rootLower, 12); memcpy(d.rootHi, mb->bvh.rootUpper, 12);
        if (c->device >= 0) {
            int ru = 0;
            std::vector<void*> staging;                  // moved into the scene-wide arrays by the next commitScene
            const Node64* dn = nullptr; const float4* dt = nullptr; const float2* du = nullptr; const float4* ds = nullptr; const uint32_t* dp = nullptr;
            ru |= uploadVec(c, staging, mb->bvh.nodes64.data(), mb->bvh.nodes64.size(), &dn); hm->d_nodes = (void*)dn;
            ru |= uploadVec(c, staging, mb->tris.data(), mb->tris.size(), &dt); hm->d_tris = (void*)dt;
            ru |= uploadVec(c, staging, mb->triUv.data(), mb->triUv.size(), &du); hm->d_triUv = (void*)du;
            ru |= uploadVec(c, staging, mb->shade.data(), mb->shade.size(), &ds); hm->d_triShade = (void*)ds;
            ru |= uploadVec(c, staging, mb->primToTri.data(), mb->primToTri.size(), &dp); hm->d_primToTri = (void*)dp;
            if (c->buildFast) {
                const Wide8* dw = nullptr; const float4* dft = nullptr; const float2* dfu = nullptr;
                ru |= uploadVec(c, staging, mb->wide.nodes.data(), mb->wide.nodes.size(), &dw); hm->d_fastNodes = (void*)dw;
                ru |= uploadVec(c, staging, mb->fastTris.data(), mb->fastTris.size(), &dft); hm->d_fastTris = (void*)dft;
                ru |= uploadVec(c, staging, mb->fastTriUv.data(), mb->fastTriUv.size(), &dfu); hm->d_fastTriUv = (void*)dfu;
            }
            cudaError_t e = ru ? cudaSuccess : cudaStreamSynchronize(c->stream);
            if (ru || e != cudaSuccess) {
                for (auto p : staging) cudaFree(p);
                delete hm;
                return ru ? ru : setError(c, PRT_ERR_CUDA, "add_mesh upload: %s", cudaGetErrorString(e));
            }
        }
        for (const DMaterial& m : mb->materials) c->materials.push_back(m);
        // Scene::add (scene.cpp:19-27)
        for (int k = 0; k < 3; k++) { c->bboxLo[k] = std::fmin(c->bboxLo[k], mb->lower[k]); c->bboxHi[k] = std::fmax(c->bboxHi[k], mb->upper[k]); }
        float cc[3], dd[3];
        for (int k = 0; k < 3; k++) { cc[k] = 0.5f * (c->bboxHi[k] + c->bboxLo[k]); dd[k] = c->bboxHi[k] - cc[k]; }   // vecmath.h:1108
        float d2[3] = {dd[0] * dd[0], dd[1] * dd[1], dd[2] * dd[2]};
        c->scene.radius = sqrtf(d2[0] + d2[1] + d2[2]);
        c->meshes.push_back(hm);
        c->dirty = true;
        return PRT_OK;
    });
    if (ctx->device >= 0) mb->releaseStaging();
    if (r) return r;
    for (uint32_t i = 0; i < md->materialCount; i++) md->materials[i].alphaTest = mb->materials[i].alphaTest;
    return (int)meshId;
}

int prt_cuda_scene_set_directional_light(prt_ctx* ctx, const float dir[3], const float intensity[3]) {   // scene.h:30-35
    if (!ctx || !dir || !intensity) return setError(ctx, PRT_ERR_ARG, "set_directional_light: bad argument");
    return forEachDevice(ctx, [&](prt_ctx* c, int) -> int {
        memcpy(c->scene.lightDir, dir, 12); memcpy(c->scene.lightIntensity, intensity, 12);
        c->scene.lightType = 1;
        return PRT_OK;
    });
}

int prt_cuda_scene_set_infinite_light(prt_ctx* ctx, int w, int h, const float* rgba) {                    // light.cpp:30-82
    if (!ctx || !rgba || w <= 1 || h <= 1 || w > 65535 || h > 65535) return setError(ctx, PRT_ERR_ARG, "set_infinite_light: bad argument");
    HostLight l;
    l.width = w; l.height = h;
    l.vertP.assign(h, 0.0f); l.horiP.assign((size_t)w * h, 0.0f);
    const float kPi = 3.14159265358979323846f;
    float vsum = 0.0f;
    for (int y = 0; y < h; y++) {
        float hsum = 0.0f;
        for (int x = 0; x < w; x++) {
            size_t b = x + (size_t)y * w;
            float cx = rgba[4 * b], cy = rgba[4 * b + 1], cz = rgba[4 * b + 2];
            float xx = cx * cx, yy = cy * cy, zz = cz * cz;
            float len = sqrtf(xx + yy + zz);
            l.horiP[b] = len;
            hsum += len;
        }
        float sinPhi = std::sin(kPi * (y + 0.5f) / h);
        float vy = hsum * sinPhi;
        l.vertP[y] = vy;
        vsum += vy;
        float invH = 1.0f / hsum, acc = 0.0f;
        for (int x = 0; x < w; x++) { size_t b = x + (size_t)y * w; float ph = acc + invH * l.horiP[b]; l.horiP[b] = ph; acc = ph; }
    }
    float invV = 1.0f / vsum, acc = 0.0f;
    for (int y = 0; y < h; y++) { float pv = acc + invV * l.vertP[y]; l.vertP[y] = pv; acc = pv; }
    // guide tables for the CDF searches: guide[g] = first index i in [1, n) with cdf[i] > g/256 (or n).  For u in bucket g
    // the reference's linear scan result lies in [guide[g], guide[g+1]], so the device searches only that range
    // (typically 0-2 probes instead of log2 n dependent loads); the result is the same index.
    auto guide = [](const float* cdf, int n, uint16_t* out) {
        int i = 1;
        for (int g = 0; g <= 256; g++) {
            const float u = (float)g / 256.0f;
            while (i < n && !(cdf[i] > u)) i++;
            out[g] = (uint16_t)i;
        }
    };
    std::vector<uint16_t> gv(257), gh((size_t)h * 257);
    guide(l.vertP.data(), h, gv.data());
    for (int y = 0; y < h; y++) guide(l.horiP.data() + (size_t)y * w, w, gh.data() + (size_t)y * 257);
    return forEachDevice(ctx, [&](prt_ctx* c, int) -> int {
        DeviceGuard g(c->device);
        c->light = l;
        c->scene.infW = w; c->scene.infH = h;
        c->scene.lightType = 2;
        if (c->device >= 0) {
            PRT_CUDA_CHECK(c, cudaStreamSynchronize(c->stream));  // nothing in flight reads the tables being replaced
            for (auto p : c->lightAllocs) cudaFree(p);
            c->lightAllocs.clear();
            int r = 0;
            r |= uploadVec(c, c->lightAllocs, rgba, 4 * (size_t)w * h, &c->scene.infImage);
            r |= uploadVec(c, c->lightAllocs, l.vertP.data(), l.vertP.size(), &c->scene.infVertP);
            r |= uploadVec(c, c->lightAllocs, l.horiP.data(), l.horiP.size(), &c->scene.infHoriP);
            r |= uploadVec(c, c->lightAllocs, gv.data(), gv.size(), &c->scene.infGuideV);
            r |= uploadVec(c, c->lightAllocs, gh.data(), gh.size(), &c->scene.infGuideH);
            if (r) return PRT_ERR_CUDA;
            PRT_CUDA_CHECK(c, cudaStreamSynchronize(c->stream));
        }
        return PRT_OK;
    });
}

int prt_cuda_scene_get_radius(const prt_ctx* ctx, float* radius) {
    if (!ctx || !radius) return PRT_ERR_ARG;
    *radius = ctx->scene.radius;
    return PRT_OK;
}

int prt_cuda_set_camera(prt_ctx* ctx, const float pos[3], const float dir[3], uint32_t width, uint32_t height) {   // camera.h:17-36
    if (!ctx || !pos || !dir || width == 0 || height == 0) return setError(ctx, PRT_ERR_ARG, "set_camera: bad argument");
    auto cross = [](const float* a, const float* b, float* o) { o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0]; };
    auto length = [](const float* a) { float x = a[0] * a[0], y = a[1] * a[1], z = a[2] * a[2]; return sqrtf(x + y + z); };
    auto normalize = [&](const float* a, float* o) { float inv = 1.0f / length(a); o[0] = inv * a[0]; o[1] = inv * a[1]; o[2] = inv * a[2]; };
    DCamera c{};
    memcpy(c.pos, pos, 12);
    normalize(dir, c.dir);
    c.width = width; c.height = height; c.invW = 1.0f / width; c.invH = 1.0f / height;
    float up[3] = {0, 1.0f, 0}, right[3], x1[3] = {1, 0, 0};
    cross(dir, up, right);                                   // un-normalised dir, as the reference
    if (length(right) < 0.00001f) cross(dir, x1, right);
    float rn[3]; normalize(right, rn);
    float u2[3]; cross(rn, dir, u2);
    normalize(u2, c.up);
    memcpy(c.right, rn, 12);
    ctx->scene.cam = c; ctx->hasCamera = true;
    for (auto p : ctx->peers) { p->scene.cam = c; p->hasCamera = true; }   // plain host state: no worker round trip
    return PRT_OK;
}

int prt_cuda_bvh_node_count(const prt_ctx* ctx, int mesh, uint32_t* nodeCount) {
    if (!ctx || !nodeCount || mesh < 0 || mesh >= (int)ctx->meshes.size()) return PRT_ERR_ARG;
    *nodeCount = (uint32_t)ctx->meshes[mesh]->build->bvh.nodes32.size();
    return PRT_OK;
}
int prt_cuda_bvh_get(const prt_ctx* ctx, int mesh, uint32_t* nodes32, uint32_t* remap) {
    if (!ctx || mesh < 0 || mesh >= (int)ctx->meshes.size()) return PRT_ERR_ARG;
    const BvhHost& b = ctx->meshes[mesh]->build->bvh;
    if (nodes32) memcpy(nodes32, b.nodes32.data(), b.nodes32.size() * sizeof(Node32));
    if (remap) memcpy(remap, b.remap.data(), b.remap.size() * 4);
    return PRT_OK;
}
int prt_cuda_bvh_get_fast(const prt_ctx* ctx, int mesh, uint32_t* nodeCount, uint32_t* rootRef, uint32_t* nodes, float* tris) {
    if (!ctx || mesh < 0 || mesh >= (int)ctx->meshes.size()) return PRT_ERR_ARG;
    const MeshBuild& b = *ctx->meshes[mesh]->build;
    const HostMesh& hm = *ctx->meshes[mesh];
    if (!hm.fastNodeCount && !(hm.fastRootRef & kLeafFlag)) return PRT_ERR_STATE;               // built without PRT_OPTION_BUILD_FAST_TRAVERSAL
    if (nodeCount) *nodeCount = hm.fastNodeCount;
    if (rootRef) *rootRef = b.wide.rootRef;
    if ((nodes && b.wide.nodes.size() != hm.fastNodeCount) || (tris && b.fastTris.size() != 3 * (size_t)hm.primCount)) return PRT_ERR_STATE;   // staging released after the upload
    if (nodes) memcpy(nodes, b.wide.nodes.data(), b.wide.nodes.size() * sizeof(Wide8));
    if (tris) memcpy(tris, b.fastTris.data(), b.fastTris.size() * sizeof(float4));
    return PRT_OK;
}
int prt_cuda_light_get_tables(const prt_ctx* ctx, float* vertP, float* horiP) {
    if (!ctx || ctx->light.width == 0) return PRT_ERR_STATE;
    if (vertP) memcpy(vertP, ctx->light.vertP.data(), ctx->light.vertP.size() * 4);
    if (horiP) memcpy(horiP, ctx->light.horiP.data(), ctx->light.horiP.size() * 4);
    return PRT_OK;
}
int prt_cuda_get_camera(const prt_ctx* ctx, float out[14]) {
    if (!ctx || !ctx->hasCamera) return PRT_ERR_STATE;
    const DCamera& c = ctx->scene.cam;
    memcpy(out, c.pos, 12); memcpy(out + 3, c.dir, 12); memcpy(out + 6, c.up, 12); memcpy(out + 9, c.right, 12);
    out[12] = c.invW; out[13] = c.invH;
    return PRT_OK;
}

int prt_cuda_synchronize(prt_ctx* ctx) {
    if (!ctx || ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "no CUDA device");
    return forEachDevice(ctx, [](prt_ctx* c, int) -> int {
        DeviceGuard g(c->device);
        PRT_CUDA_CHECK(c, cudaStreamSynchronize(c->stream));
        return PRT_OK;
    });
}
// Counters of a multi-GPU context are the sums over its GPUs (times: the leader's).
int prt_cuda_get_counters(prt_ctx* ctx, prt_counters* out) {
    if (!ctx || !out) return PRT_ERR_ARG;
    if (ctx->device >= 0) {
        int r = forEachDevice(ctx, [](prt_ctx* c, int) -> int {
            DeviceGuard g(c->device);
            int rr = foldWavefrontCounters(c);
            return rr ? rr : foldStatCounters(c);
        });
        if (r) return r;
        if (ctx->gatherRecorded) {
            DeviceGuard g(ctx->device);
            float ms = 0.0f;
            PRT_CUDA_CHECK(ctx, cudaEventSynchronize(ctx->evG1));
            PRT_CUDA_CHECK(ctx, cudaEventElapsedTime(&ms, ctx->evG0, ctx->evG1));
            ctx->counters.lastGatherMs = ms;
            ctx->gatherRecorded = false;
        }
    }
    *out = ctx->counters;
    for (auto p : ctx->peers) {
        const prt_counters& q = p->counters;
        out->extendRays += q.extendRays; out->shadowRays += q.shadowRays; out->primaryRays += q.primaryRays;
        out->kernelLaunches += q.kernelLaunches; out->shadowRaysSkipped += q.shadowRaysSkipped;
        for (int c = 0; c < 3; c++) {
            out->statRays[c] += q.statRays[c]; out->statNodeSteps[c] += q.statNodeSteps[c]; out->statLeafVisits[c] += q.statLeafVisits[c];
            out->statTriTests[c] += q.statTriTests[c]; out->statUvLoads[c] += q.statUvLoads[c];
        }
    }
    return PRT_OK;
}
int prt_cuda_reset_counters(prt_ctx* ctx) {
    if (!ctx) return PRT_ERR_ARG;
    return forEachDevice(ctx, [](prt_ctx* c, int) -> int {
        DeviceGuard g(c->device);
        if (c->device >= 0) { foldWavefrontCounters(c); foldStatCounters(c); }
        c->counters = prt_counters{};
        return PRT_OK;
    });
}
uint64_t prt_cuda_stream(prt_ctx* ctx) { return ctx ? (uint64_t)(uintptr_t)ctx->stream : 0; }

uint32_t prt_packet_seed(uint32_t seed, uint32_t pixel, uint32_t ppp, uint32_t packet) {
    uint32_t h = seed * 0x9E3779B9u + pixel * ppp + packet;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h | 1u;
}

}  // extern "C"
