// Host-side context shared by the C-ABI translation units (prt_cuda.cu, wavefront.cu).
#pragma once
#include <cstdint>
#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>
#include <unordered_map>
#include <cuda_runtime.h>

#include "../../include/prt_cuda.h"
#include "bvh_build.h"
#include "prt_device.cuh"

namespace prt_b200 {

struct HostMesh {
    BvhHost bvh;
    uint32_t primCount = 0, vertexCount = 0;
    float lower[3], upper[3];                 // Mesh::calculateBounds (mesh.cpp:300-309)
    std::vector<void*> deviceAllocs;
    // traversal geometry, uploaded with mesh-local references by add_mesh and moved into the scene-wide arrays
    // (references rebased) by the next commitScene, which frees these staging copies
    void *d_nodes = nullptr, *d_tris = nullptr, *d_triUv = nullptr, *d_triShade = nullptr;
    DMesh d{};
};

struct HostLight {
    int width = 0, height = 0;
    std::vector<float> vertP, horiP;
};

struct WavefrontBuffers;                      // wavefront.cu

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
    prt_b200::HostLight light;
    prt_b200::DScene scene{};                           // host copy of the kernel parameter block
    void *d_meshes = nullptr, *d_materials = nullptr, *d_textures = nullptr;
    // DScene::nodes / tris / triUv: every committed mesh back to back (meshes are only ever appended)
    void *d_allNodes = nullptr, *d_allTris = nullptr, *d_allTriUv = nullptr, *d_allTriShade = nullptr;
    size_t mergedNodes = 0, mergedTris = 0, mergedMeshes = 0;
    bool dirty = true;
    float bboxLo[3], bboxHi[3];                         // Scene::m_bbox (scene.cpp:23)
    bool hasCamera = false;
    // scratch for host-pointer entry points
    void* d_scratch[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t scratchBytes[4] = {0, 0, 0, 0};
    uint32_t* d_counter = nullptr;                      // persistent-thread work counters
    prt_counters counters{};
    int smCount = 148;
    bool traceFlagsForced = false;                      // PRT_TRACE_FLAGS set: do not pick traceFlags from the scene size
    int traceImpl = 1;                                  // 1: dynamic-fetch while-while loop (prt_trace.cuh); 0: one ray per lane per fetch (PRT_TRACE_IMPL=0)
    prt_b200::WavefrontBuffers* wf = nullptr;
    // per-batch queue counters copied back asynchronously (pinned), folded into `counters` on demand
    uint32_t* h_wfCounters = nullptr;
    size_t wfCountersCap = 0, wfCountersPending = 0;
};

namespace prt_b200 {

int setError(prt_ctx* ctx, int code, const char* fmt, ...);
int commitScene(prt_ctx* ctx);                          // uploads tables if dirty
int ensureScratch(prt_ctx* ctx, int slot, size_t bytes);
void destroyWavefront(prt_ctx* ctx);
int foldWavefrontCounters(prt_ctx* ctx);               // syncs the stream and adds pending per-batch ray counts

#define PRT_CUDA_CHECK(ctx, expr)                                                                  \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) return prt_b200::setError((ctx), PRT_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

}  // namespace prt_b200
