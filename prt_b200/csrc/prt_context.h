// Host-side context shared by the C-ABI translation units (prt_cuda.cu, trace_kernels.cu, wavefront.cu, multi_gpu.cu ...).
#pragma once
#include <cstdint>
#include <cstdarg>
#include <cstdio>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include <unordered_map>
#include <cuda_runtime.h>

#include "../../include/prt_cuda.h"
#include "bvh_build.h"
#include "prt_device.cuh"

namespace prt_b200 {

// Everything add_mesh computes on the host, built ONCE per mesh and uploaded to every GPU of the context.
struct MeshBuild {
    BvhHost bvh;
    uint32_t primCount = 0, vertexCount = 0;
    float lower[3], upper[3];                 // Mesh::calculateBounds (mesh.cpp:300-309)
    std::vector<DMaterial> materials;         // texture ids are scene-wide already (every GPU adds textures in the same order)
    bool anyAlpha = false;
    // staging copies of the device arrays, released once every GPU has them
    std::vector<float4> tris, shade;
    std::vector<float2> triUv;
    std::vector<uint32_t> primToTri;          // inverse of bvh.remap: primitive id -> leaf-ordered triangle slot (mesh-local)
    // fast traversal mode (PRT_OPTION_BUILD_FAST_TRAVERSAL): its own tree (leaves <= 4), collapsed to Wide8 records, and the
    // triangle records in that tree's leaf order (vertex 0's w = the triangle's slot in `tris`)
    WideHost wide;
    std::vector<float4> fastTris;
    std::vector<float2> fastTriUv;
    void releaseStaging() {
        std::vector<float4>().swap(tris); std::vector<float4>().swap(shade); std::vector<float2>().swap(triUv);
        std::vector<uint32_t>().swap(primToTri); std::vector<Node64>().swap(bvh.nodes64);
        std::vector<Wide8>().swap(wide.nodes); std::vector<float4>().swap(fastTris); std::vector<float2>().swap(fastTriUv);
    }
};

struct HostMesh {
    std::shared_ptr<MeshBuild> build;
    uint32_t primCount = 0, vertexCount = 0, nodeCount = 0;
    // traversal geometry, uploaded with mesh-local references by add_mesh and moved into the scene-wide arrays
    // (references rebased) by the next commitScene, which frees these staging copies
    void *d_nodes = nullptr, *d_tris = nullptr, *d_triUv = nullptr, *d_triShade = nullptr, *d_primToTri = nullptr;
    void *d_fastNodes = nullptr, *d_fastTris = nullptr, *d_fastTriUv = nullptr;
    uint32_t fastNodeCount = 0, fastRootRef = 0, fastDepth = 0;
    DMesh d{};
};

struct HostLight {
    int width = 0, height = 0;
    std::vector<float> vertP, horiP;
};

struct WavefrontBuffers;                      // wavefront.cu

// One worker thread per additional GPU of a multi-device context ("one host thread + stream per GPU"): the
// calling thread serves devices[0], worker k serves devices[k] with that device current for its whole life.
class DeviceWorker {
public:
    explicit DeviceWorker(int device);
    ~DeviceWorker();
    void post(std::function<int()> fn);       // runs fn on the worker thread
    int wait();                               // result of the last posted task
private:
    void loop(int device);
    std::thread th_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::function<int()> task_;
    bool hasTask_ = false, busy_ = false, quit_ = false;
    int result_ = 0;
};

// Saves the caller's current device, makes `device` current, restores on scope exit.  Every extern "C" entry that touches
// CUDA opens one: a context keeps working when another context (another GPU) was used in between or when a thread other
// than the creating one calls in — the reference's ThreadPool pattern (main.cpp:132-158).
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int device) {
        if (device < 0) return;
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        if (prev != device) { cudaSetDevice(device); switched = true; }
    }
    ~DeviceGuard() { if (switched && prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

}  // namespace prt_b200

struct prt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // host-buffer entry points: H2D / kernels / D2H of consecutive chunks overlap on three streams
    cudaStream_t copyIn = nullptr, copyOut = nullptr;
    cudaStream_t laneStreams[3] = {nullptr, nullptr, nullptr};   // streams of the renderer's extra wavefront lanes (wavefront.cu)
    cudaEvent_t evLane = nullptr;
    std::vector<cudaEvent_t> pipeEv;                    // [2*chunk] = chunk uploaded, [2*chunk+1] = chunk traced
    bool inPipe = false;                                // chunk launches of a pipelined call: counters pre-zeroed, no per-launch events
    uint32_t counterSlot = 0;                           // which of the 64 work counters the next launch uses
    std::unordered_map<const void*, int> gridCache;     // persistent-grid size per kernel (occupancy query done once)
    std::string lastError;
    std::vector<prt_b200::HostMesh*> meshes;
    std::vector<prt_b200::DTexture> textures;           // device pointers inside
    std::vector<uint8_t> textureNeedsAlpha;             // Texture::isAlphaTestRequired (texture.cpp:338-350)
    std::vector<prt_b200::DMaterial> materials;
    std::vector<void*> deviceAllocs;
    std::vector<void*> lightAllocs;                     // infinite-area light tables (freed when the light is set again)
    prt_b200::HostLight light;
    prt_b200::DScene scene{};                           // host copy of the kernel parameter block
    void *d_meshes = nullptr, *d_materials = nullptr, *d_textures = nullptr;
    // DScene::nodes / tris / triUv: every committed mesh back to back (meshes are only ever appended)
    void *d_allNodes = nullptr, *d_allTris = nullptr, *d_allTriUv = nullptr, *d_allTriShade = nullptr, *d_allPrimToTri = nullptr;
    size_t mergedNodes = 0, mergedTris = 0, mergedMeshes = 0;
    void *d_allFastNodes = nullptr, *d_allFastTris = nullptr, *d_allFastTriUv = nullptr;
    size_t mergedFastNodes = 0;
    bool buildFast = false;                             // PRT_OPTION_BUILD_FAST_TRAVERSAL
    int traversal = 0;                                  // PRT_OPTION_TRAVERSAL
    bool dirty = true;
    float bboxLo[3], bboxHi[3];                         // Scene::m_bbox (scene.cpp:23)
    bool hasCamera = false;
    // scratch for host-pointer entry points
    void* d_scratch[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t scratchBytes[4] = {0, 0, 0, 0};
    uint32_t* d_counter = nullptr;                      // persistent-thread work counters
    unsigned long long* d_stats = nullptr;              // counting build only: [5][3] traversal counters (prt_counters.stat*)
    prt_counters counters{};
    int smCount = 148;
    bool traceFlagsForced = false;                      // PRT_TRACE_FLAGS set: do not pick traceFlags from the scene size
    int traceImpl = 1;                                  // 1: dynamic-fetch while-while loop (prt_trace.cuh); 0: one ray per lane per fetch (PRT_TRACE_IMPL=0)
    prt_b200::WavefrontBuffers* wf = nullptr;
    // per-batch queue counters copied back asynchronously (pinned), folded into `counters` on demand
    uint32_t* h_wfCounters = nullptr;
    size_t wfCountersCap = 0, wfCountersPending = 0;

    // ---- multi-GPU (multi_gpu.cu) ----
    // One process, several GPUs: the context the caller holds is the leader (rank 0); peers[k] is a full context on
    // devices[k+1] with rank k+1, served by workers[k].  One process per GPU: no peers, rank / world set by comm_init_rank.
    std::vector<prt_ctx*> peers;
    std::vector<prt_b200::DeviceWorker*> workers;
    prt_ctx* leader = nullptr;                          // peers only
    void* comm = nullptr;                               // ncclComm_t of this context's GPU
    int rank = 0, world = 1;
    float *d_bandLocal = nullptr, *d_bandAll = nullptr; // render bands of this rank (all-gather send buffer) / of every rank
    size_t bandLocalFloats = 0, bandAllFloats = 0;
    cudaEvent_t evG0 = nullptr, evG1 = nullptr;         // around the all-gather + de-tile
    bool gatherRecorded = false;                        // evG0 / evG1 hold a gather whose time has not been read yet
};

namespace prt_b200 {

int setError(prt_ctx* ctx, int code, const char* fmt, ...);
int commitScene(prt_ctx* ctx);                          // uploads tables if dirty
int ensureScratch(prt_ctx* ctx, int slot, size_t bytes);
void destroyWavefront(prt_ctx* ctx);
int foldWavefrontCounters(prt_ctx* ctx);               // syncs the stream and adds pending per-batch ray counts
int foldStatCounters(prt_ctx* ctx);                    // counting build: reads and clears the device traversal counters

// fn(c, k) on the leader (k = 0, calling thread) and on every peer (k = 1.., its worker thread), each with its device
// current; returns the first non-zero result (the leader's lastError carries the message).
int forEachDevice(prt_ctx* ctx, const std::function<int(prt_ctx*, int)>& fn);
void destroyMulti(prt_ctx* ctx);                        // communicators and worker threads
int initCommAll(prt_ctx* ctx, const int* devices, int ndev);   // ncclCommInitAll over the leader and its peers

// multi_gpu.cu: the collective half of a sharded render — all-gather of the band buffers + de-tile into d_rgb
// (full-image layout, row stride imageWidth).  Called on the leader / the rank's own context.
int gatherBands(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t bandRows, uint32_t imageWidth, float* d_rgb);
int ensureBandBuffers(prt_ctx* ctx, size_t localFloats);

// wavefront.cu: the local half — this rank's bands of region [y0,y1) x [x0,x1) into ctx->d_bandLocal (tile-major, row stride x1 - x0)
int renderBandsLocal(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t bandRows, uint32_t spp, uint32_t seed, float exposure);
uint32_t bandRowsDefault();

#define PRT_CUDA_CHECK(ctx, expr)                                                                  \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) return prt_b200::setError((ctx), PRT_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

}  // namespace prt_b200
