// Multi-GPU half of the C ABI: communicators, the framebuffer all-gather and the de-tile kernel.
//
// The reference's only parallel driver is the tile loop of raytrace_scene (main.cpp:121-158): one PathTracer::TraceBlock per
// 16x16 tile on a thread pool, all tiles writing one shared Image.  Here the unit dealt out is a band of 8 image rows, the
// workers are GPUs with the scene replicated, and the shared Image becomes ONE collective: every GPU renders its bands into a
// rank-local tile-major buffer that IS the all-gather send buffer (wavefront.cu writes it in that layout, nothing is
// repacked), ncclAllGather moves the bands over NVLink / NVSwitch, and deTileBands turns the gathered [world][bands][rows]
// block back into the row-major Image layout.  Pixels do not depend on the partition (per-packet seeding), so there is no
// reduction and no halo; 1 / 2 / 4 / 8-GPU images are bit-identical.
//
// Two ways in, same data path: prt_cuda_create_multi (one process, one worker thread + stream per GPU, ncclCommInitAll) and
// prt_cuda_comm_init_rank (one process per GPU, ncclCommInitRank with an id the caller broadcast).
#include "prt_context.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <nccl.h>

using namespace prt_b200;

static_assert(PRT_COMM_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "PRT_COMM_ID_BYTES must be the size of ncclUniqueId");

namespace {

// NCCL is bound at the first communicator call instead of at load time.  Two copies with the SAME soname exist on the target
// image — the system libnccl.so.2 (2.27) and the one PyTorch bundles (2.28) — and a process can hold only one: had this
// library a DT_NEEDED entry for libnccl.so.2, loading it before `import torch` would hand torch the older copy
// ("undefined symbol: ncclDevCommCreate").  So: reuse whichever libnccl.so.2 the process already has (torch's, when the
// caller is a torch program), otherwise open the system one (a plain C++ host).  Single-GPU use never touches NCCL.
struct NcclApi {
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    std::string error;
    bool ok = false;
};

const NcclApi& nccl() {
    static const NcclApi api = [] {
        NcclApi a;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!h) { a.error = std::string("cannot load libnccl.so.2: ") + dlerror(); return a; }
        auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p && a.error.empty()) a.error = std::string("libnccl.so.2 lacks ") + name; return p; };
        a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
        a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
        a.CommInitAll = (decltype(a.CommInitAll))sym("ncclCommInitAll");
        a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
        a.AllGather = (decltype(a.AllGather))sym("ncclAllGather");
        a.GroupStart = (decltype(a.GroupStart))sym("ncclGroupStart");
        a.GroupEnd = (decltype(a.GroupEnd))sym("ncclGroupEnd");
        a.ok = a.error.empty();
        return a;
    }();
    return api;
}

#define PRT_NCCL_READY(ctx) do { if (!nccl().ok) return prt_b200::setError((ctx), PRT_ERR_CUDA, "NCCL unavailable: %s", nccl().error.c_str()); } while (0)
#define PRT_NCCL_CHECK(ctx, expr)                                                                  \
    do {                                                                                           \
        ncclResult_t _e = (expr);                                                                  \
        if (_e != ncclSuccess) return prt_b200::setError((ctx), PRT_ERR_CUDA, "%s: %s (%s:%d)", #expr, nccl().GetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// gathered: [world][bandsPerRank * bandRows][bw][3] floats (rank-major, as ncclAllGather leaves it).  Region row r belongs to
// band b = r / bandRows, which rank b % world rendered as its (b / world)-th band.
__global__ void deTileBands(const float* __restrict__ gathered, float* __restrict__ rgb, uint32_t x0, uint32_t y0, uint32_t bw, uint32_t regionH,
                            uint32_t bandRows, uint32_t world, size_t localFloats, uint32_t imageWidth) {
    const size_t rowFloats = 3 * (size_t)bw, total = rowFloats * regionH;
    for (size_t f = blockIdx.x * (size_t)blockDim.x + threadIdx.x; f < total; f += (size_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(f / rowFloats), col = (uint32_t)(f - (size_t)r * rowFloats);
        const uint32_t b = r / bandRows, rib = r - b * bandRows;
        const uint32_t owner = b % world, j = b / world;
        const float v = __ldcs(gathered + owner * localFloats + ((size_t)j * bandRows + rib) * rowFloats + col);
        rgb[3 * ((size_t)(y0 + r) * imageWidth + x0) + col] = v;
    }
}

}  // namespace

namespace prt_b200 {

uint32_t bandRowsDefault() {
    // 8-row bands dealt round-robin: every rank sees every part of the image (a sky-only band costs almost nothing, an
    // interior band runs 15-deep paths), and a 4K frame still gives each of 8 ranks 34 bands.  PRT_BAND_ROWS overrides (profiling).
    static const uint32_t v = getenv("PRT_BAND_ROWS") && atoi(getenv("PRT_BAND_ROWS")) > 0 ? (uint32_t)atoi(getenv("PRT_BAND_ROWS")) : 8u;
    return v;
}

int initCommAll(prt_ctx* ctx, const int* devices, int ndev) {
    PRT_NCCL_READY(ctx);
    std::vector<ncclComm_t> comms(ndev);
    PRT_NCCL_CHECK(ctx, nccl().CommInitAll(comms.data(), ndev, devices));
    ctx->comm = comms[0]; ctx->rank = 0; ctx->world = ndev;
    for (int k = 1; k < ndev; k++) { ctx->peers[k - 1]->comm = comms[k]; ctx->peers[k - 1]->rank = k; ctx->peers[k - 1]->world = ndev; }
    return PRT_OK;
}

void destroyMulti(prt_ctx* ctx) {
    auto drop = [](prt_ctx* c) {
        if (!c->comm) return;
        DeviceGuard g(c->device);
        cudaStreamSynchronize(c->stream);
        nccl().CommDestroy((ncclComm_t)c->comm);
        c->comm = nullptr;
    };
    for (auto p : ctx->peers) drop(p);
    drop(ctx);
    for (auto w : ctx->workers) delete w;
    ctx->workers.clear();
}

int ensureBandBuffers(prt_ctx* ctx, size_t localFloats) {
    const size_t allFloats = localFloats * (size_t)ctx->world;
    if (ctx->bandLocalFloats < localFloats) {
        if (ctx->d_bandLocal) cudaFree(ctx->d_bandLocal);
        ctx->d_bandLocal = nullptr; ctx->bandLocalFloats = 0;
        PRT_CUDA_CHECK(ctx, cudaMalloc((void**)&ctx->d_bandLocal, localFloats * sizeof(float)));
        ctx->bandLocalFloats = localFloats;
    }
    if (ctx->bandAllFloats < allFloats) {
        if (ctx->d_bandAll) cudaFree(ctx->d_bandAll);
        ctx->d_bandAll = nullptr; ctx->bandAllFloats = 0;
        PRT_CUDA_CHECK(ctx, cudaMalloc((void**)&ctx->d_bandAll, allFloats * sizeof(float)));
        ctx->bandAllFloats = allFloats;
    }
    if (!ctx->evG0) { PRT_CUDA_CHECK(ctx, cudaEventCreate(&ctx->evG0)); PRT_CUDA_CHECK(ctx, cudaEventCreate(&ctx->evG1)); }
    return PRT_OK;
}

// The collective: every GPU's band buffer to every GPU, then the de-tile on the GPU(s) that own an output image.
// One process per GPU: this rank's own all-gather on its stream.  One process, several GPUs: the calling thread issues the
// all-gathers of all its GPUs as one NCCL group, each on that GPU's stream, i.e. ordered after the render the GPU's worker
// thread enqueued there; only the leader (devices[0], where d_rgb lives) de-tiles.
int gatherBands(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t bandRows, uint32_t imageWidth, float* d_rgb) {
    const uint32_t bw = x1 - x0, regionH = y1 - y0;
    const uint32_t nb = (regionH + bandRows - 1) / bandRows, nbPerRank = (nb + ctx->world - 1) / ctx->world;
    const size_t localFloats = (size_t)nbPerRank * bandRows * bw * 3;
    DeviceGuard g(ctx->device);
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->evG0, ctx->stream));
    if (ctx->peers.empty()) {
        PRT_NCCL_CHECK(ctx, nccl().AllGather(ctx->d_bandLocal, ctx->d_bandAll, localFloats, ncclFloat, (ncclComm_t)ctx->comm, ctx->stream));
    } else {
        PRT_NCCL_CHECK(ctx, nccl().GroupStart());
        ncclResult_t e = nccl().AllGather(ctx->d_bandLocal, ctx->d_bandAll, localFloats, ncclFloat, (ncclComm_t)ctx->comm, ctx->stream);
        for (auto p : ctx->peers)
            if (e == ncclSuccess) e = nccl().AllGather(p->d_bandLocal, p->d_bandAll, localFloats, ncclFloat, (ncclComm_t)p->comm, p->stream);
        ncclResult_t e2 = nccl().GroupEnd();
        if (e != ncclSuccess || e2 != ncclSuccess) return setError(ctx, PRT_ERR_CUDA, "ncclAllGather (framebuffer bands): %s", nccl().GetErrorString(e != ncclSuccess ? e : e2));
    }
    const size_t total = 3 * (size_t)bw * regionH;
    const int grid = (int)std::min<size_t>((total + 255) / 256, (size_t)ctx->smCount * 16);
    deTileBands<<<grid, 256, 0, ctx->stream>>>(ctx->d_bandAll, d_rgb, x0, y0, bw, regionH, bandRows, (uint32_t)ctx->world, localFloats, imageWidth);
    PRT_CUDA_CHECK(ctx, cudaGetLastError());
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->evG1, ctx->stream));
    ctx->gatherRecorded = true;
    ctx->counters.kernelLaunches++;
    return PRT_OK;
}

}  // namespace prt_b200

extern "C" {

int prt_cuda_comm_unique_id(void* id) {
    if (!id) return PRT_ERR_ARG;
    PRT_NCCL_READY(nullptr);
    ncclUniqueId u;
    ncclResult_t e = nccl().GetUniqueId(&u);
    if (e != ncclSuccess) { setError(nullptr, PRT_ERR_CUDA, "ncclGetUniqueId: %s", nccl().GetErrorString(e)); return PRT_ERR_CUDA; }
    memcpy(id, &u, sizeof(u));
    return PRT_OK;
}

int prt_cuda_comm_init_rank(prt_ctx* ctx, const void* id, int rank, int world) {
    if (!ctx || !id || world < 1 || rank < 0 || rank >= world) return setError(ctx, PRT_ERR_ARG, "comm_init_rank: bad argument");
    if (ctx->device < 0) return setError(ctx, PRT_ERR_CUDA, "comm_init_rank: host-only context");
    if (!ctx->peers.empty() || ctx->comm) return setError(ctx, PRT_ERR_STATE, "comm_init_rank: the context already has a communicator");
    PRT_NCCL_READY(ctx);
    DeviceGuard g(ctx->device);
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    ncclComm_t comm = nullptr;
    PRT_NCCL_CHECK(ctx, nccl().CommInitRank(&comm, world, u, rank));
    ctx->comm = comm; ctx->rank = rank; ctx->world = world;
    return PRT_OK;
}

int prt_cuda_gather_bands_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, float* d_rgb) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->world <= 1 || !ctx->comm) return setError(ctx, PRT_ERR_STATE, "gather_bands: the context has no communicator");
    if (!d_rgb || x1 <= x0 || y1 <= y0 || x1 > ctx->scene.cam.width || y1 > ctx->scene.cam.height) return setError(ctx, PRT_ERR_ARG, "gather_bands: bad region");
    const uint32_t bandRows = bandRowsDefault();
    const uint32_t nb = (y1 - y0 + bandRows - 1) / bandRows, nbPerRank = (nb + ctx->world - 1) / ctx->world;
    const size_t localFloats = (size_t)nbPerRank * bandRows * (x1 - x0) * 3;
    if (ctx->bandLocalFloats < localFloats || ctx->bandAllFloats < localFloats * ctx->world) return setError(ctx, PRT_ERR_STATE, "gather_bands: no sharded render of this region has run");
    for (auto p : ctx->peers) if (p->bandLocalFloats < localFloats) return setError(ctx, PRT_ERR_STATE, "gather_bands: no sharded render of this region has run");
    return gatherBands(ctx, x0, y0, x1, y1, bandRows, ctx->scene.cam.width, d_rgb);
}

int prt_cuda_comm_info(const prt_ctx* ctx, int* rank, int* world, int* localDevices) {
    if (!ctx) return PRT_ERR_ARG;
    if (rank) *rank = ctx->rank;
    if (world) *world = ctx->world;
    if (localDevices) *localDevices = 1 + (int)ctx->peers.size();
    return PRT_OK;
}

}  // extern "C"
