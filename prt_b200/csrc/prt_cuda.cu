// C ABI (include/prt_cuda.h): context lifetime, scene upload, batch closest-hit / any-hit / primary kernels.
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
    if (ctx) ctx->lastError = buf; else g_createError = buf;
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

// Appends the meshes added since the last commit to the scene-wide node / triangle arrays.
static int mergeGeometry(prt_ctx* ctx) {
    if (ctx->mergedMeshes == ctx->meshes.size()) return PRT_OK;
    size_t nodes = ctx->mergedNodes, tris = ctx->mergedTris;
    bool uv = ctx->d_allTriUv != nullptr, shade = ctx->d_allTriShade != nullptr;
    for (size_t k = ctx->mergedMeshes; k < ctx->meshes.size(); k++) {
        nodes += ctx->meshes[k]->bvh.nodes64.size(); tris += ctx->meshes[k]->primCount; uv |= ctx->meshes[k]->d_triUv != nullptr; shade |= ctx->meshes[k]->d_triShade != nullptr;
    }
    if (tris >= (1u << 28) || nodes >= (1u << 31)) return setError(ctx, PRT_ERR_STATE, "scene too large for 28-bit triangle / 31-bit node references (%zu triangles, %zu nodes)", tris, nodes);
    void *nn = nullptr, *nt = nullptr, *nu = nullptr, *ns = nullptr;
    if (nodes) PRT_CUDA_CHECK(ctx, cudaMalloc(&nn, nodes * 64));
    if (tris) PRT_CUDA_CHECK(ctx, cudaMalloc(&nt, tris * 48));
    if (uv && tris) { PRT_CUDA_CHECK(ctx, cudaMalloc(&nu, tris * 24)); PRT_CUDA_CHECK(ctx, cudaMemsetAsync(nu, 0, tris * 24, ctx->stream)); }
    if (shade && tris) { PRT_CUDA_CHECK(ctx, cudaMalloc(&ns, tris * 64)); PRT_CUDA_CHECK(ctx, cudaMemsetAsync(ns, 0, tris * 64, ctx->stream)); }   // zero flags: geometric normal, default uv
    if (ctx->mergedTris && ctx->d_allTriShade) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(ns, ctx->d_allTriShade, ctx->mergedTris * 64, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->mergedNodes) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nn, ctx->d_allNodes, ctx->mergedNodes * 64, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->mergedTris) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nt, ctx->d_allTris, ctx->mergedTris * 48, cudaMemcpyDeviceToDevice, ctx->stream));
    if (ctx->mergedTris && ctx->d_allTriUv) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(nu, ctx->d_allTriUv, ctx->mergedTris * 24, cudaMemcpyDeviceToDevice, ctx->stream));
    size_t nb = ctx->mergedNodes, tb = ctx->mergedTris;
    for (size_t k = ctx->mergedMeshes; k < ctx->meshes.size(); k++) {
        HostMesh* hm = ctx->meshes[k];
        const size_t nc = hm->bvh.nodes64.size(), tc = hm->primCount;
        if (nc) {
            PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nn + nb * 64, hm->d_nodes, nc * 64, cudaMemcpyDeviceToDevice, ctx->stream));
            rebaseRefs<<<(unsigned)((nc + 255) / 256), 256, 0, ctx->stream>>>((float4*)nn + 4 * nb, (uint32_t)nc, (uint32_t)nb, (uint32_t)tb);
            PRT_CUDA_CHECK(ctx, cudaGetLastError());
        }
        if (tc) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nt + tb * 48, hm->d_tris, tc * 48, cudaMemcpyDeviceToDevice, ctx->stream));
        if (tc && hm->d_triUv) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)nu + tb * 24, hm->d_triUv, tc * 24, cudaMemcpyDeviceToDevice, ctx->stream));
        if (tc && hm->d_triShade) PRT_CUDA_CHECK(ctx, cudaMemcpyAsync((char*)ns + tb * 64, hm->d_triShade, tc * 64, cudaMemcpyDeviceToDevice, ctx->stream));
        hm->d.rootRef = hm->bvh.rootRef + (uint32_t)((hm->bvh.rootRef & kLeafBit) ? tb : nb);
        nb += nc; tb += tc;
    }
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t k = ctx->mergedMeshes; k < ctx->meshes.size(); k++) {
        HostMesh* hm = ctx->meshes[k];
        cudaFree(hm->d_nodes); cudaFree(hm->d_tris); cudaFree(hm->d_triUv); cudaFree(hm->d_triShade);
        hm->d_nodes = hm->d_tris = hm->d_triUv = hm->d_triShade = nullptr;
    }
    cudaFree(ctx->d_allNodes); cudaFree(ctx->d_allTris); cudaFree(ctx->d_allTriUv); cudaFree(ctx->d_allTriShade);
    ctx->d_allNodes = nn; ctx->d_allTris = nt; ctx->d_allTriUv = nu; ctx->d_allTriShade = ns;
    ctx->mergedNodes = nodes; ctx->mergedTris = tris; ctx->mergedMeshes = ctx->meshes.size();
    ctx->scene.nodes = (const float4*)nn; ctx->scene.tris = (const float4*)nt; ctx->scene.triUv = (const float2*)nu; ctx->scene.triShade = (const float4*)ns;
    return PRT_OK;
}

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
    if (!(ctx->scene.traceFlags >> 8)) ctx->scene.traceFlags |= 12u << 8;      // bits 8-15: idle lanes before a warp refills from the queue
    ctx->dirty = false;
    return PRT_OK;
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
    auto ctx = new prt_ctx;
    ctx->device = device;
    for (int k = 0; k < 3; k++) { ctx->bboxLo[k] = FLT_MAX; ctx->bboxHi[k] = -FLT_MAX; }
    ctx->scene.radius = FLT_MAX;                                                              // scene.cpp:16
    if (const char* e = getenv("PRT_TRACE_IMPL")) ctx->traceImpl = atoi(e);                   // A/B switch for profiling only
    ctx->scene.traceFlags = 0u;
    if (const char* e = getenv("PRT_TRACE_FLAGS")) { ctx->scene.traceFlags = (uint32_t)atoi(e); ctx->traceFlagsForced = true; }  // A/B switch for profiling only
    if (device >= 0) {
        cudaError_t e = cudaSetDevice(device);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev0);
        if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1);
        if (e == cudaSuccess) e = cudaMalloc((void**)&ctx->d_counter, 64 * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&ctx->smCount, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) {
            setError(nullptr, PRT_ERR_CUDA, "prt_cuda_create(device=%d): %s — no CPU fallback exists", device, cudaGetErrorString(e));
            cudaGetLastError();
            delete ctx;
            return PRT_ERR_CUDA;
        }
    }
    *out = ctx;
    return PRT_OK;
}

void prt_cuda_destroy(prt_ctx* ctx) {
    if (!ctx) return;
    if (ctx->device >= 0) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        destroyWavefront(ctx);
        for (auto m : ctx->meshes) { for (auto p : m->deviceAllocs) cudaFree(p); cudaFree(m->d_nodes); cudaFree(m->d_tris); cudaFree(m->d_triUv); cudaFree(m->d_triShade); }
        cudaFree(ctx->d_allNodes); cudaFree(ctx->d_allTris); cudaFree(ctx->d_allTriUv); cudaFree(ctx->d_allTriShade);
        for (auto p : ctx->deviceAllocs) cudaFree(p);
        for (auto p : ctx->d_scratch) if (p) cudaFree(p);
        if (ctx->d_meshes) cudaFree(ctx->d_meshes);
        if (ctx->d_materials) cudaFree(ctx->d_materials);
        if (ctx->d_textures) cudaFree(ctx->d_textures);
        if (ctx->d_counter) cudaFree(ctx->d_counter);
        if (ctx->ev0) cudaEventDestroy(ctx->ev0);
        if (ctx->ev1) cudaEventDestroy(ctx->ev1);
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

const char* prt_cuda_last_error(const prt_ctx* ctx) { return ctx ? ctx->lastError.c_str() : g_createError.c_str(); }

int prt_cuda_scene_add_texture(prt_ctx* ctx, int width, int height, int components, int channel, const void* texels) {
    if (!ctx || !texels || width <= 0 || height <= 0 || components <= 0) return setError(ctx, PRT_ERR_ARG, "add_texture: bad argument");
    DTexture t{};
    t.width = width; t.height = height; t.comp = components; t.channel = channel;
    size_t bytes = (size_t)width * height * components * (channel ? 4 : 1);
    if (ctx->device >= 0) {
        const uint8_t* d = nullptr;
        int r = uploadVec(ctx, ctx->deviceAllocs, (const uint8_t*)texels, bytes, &d);
        if (r) return r;
        PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        t.data = d;
    }
    // alpha-test requirement, as Texture::isAlphaTestRequired (texture.cpp:338-350): any alpha < 255 in an RGBA8 map
    bool alpha = false;
    if (channel == 0 && components == 4) {
        const uint8_t* p = (const uint8_t*)texels;
        for (size_t i = 0; i < (size_t)width * height; i++) if (p[4 * i + 3] < 255) { alpha = true; break; }
    }
    ctx->textures.push_back(t);
    ctx->textureNeedsAlpha.push_back(alpha ? 1 : 0);
    ctx->dirty = true;
    return (int)ctx->textures.size() - 1;
}

int prt_cuda_scene_add_mesh(prt_ctx* ctx, prt_mesh_desc* md) {
    if (!ctx || !md || !md->indices || !md->positions || !md->primMaterial || !md->materials || md->primCount == 0)
        return setError(ctx, PRT_ERR_ARG, "add_mesh: bad argument");
    auto hm = new HostMesh;
    hm->primCount = md->primCount; hm->vertexCount = md->vertexCount;
    const uint32_t meshId = (uint32_t)ctx->meshes.size();
    const uint32_t materialBase = (uint32_t)ctx->materials.size();

    bool anyAlpha = false;
    for (uint32_t i = 0; i < md->materialCount; i++) {
        prt_material& s = md->materials[i];
        DMaterial m{};
        memcpy(m.diffuse, s.diffuse, 12); memcpy(m.emissive, s.emissive, 12);
        m.refl = s.reflType; m.diffuseTex = s.diffuseTex; m.bumpTex = s.bumpTex;
        if (s.diffuseTex >= (int)ctx->textures.size() || s.bumpTex >= (int)ctx->textures.size()) { delete hm; return setError(ctx, PRT_ERR_ARG, "add_mesh: texture id out of range"); }
        m.alphaTest = (s.diffuseTex >= 0 && ctx->textureNeedsAlpha[s.diffuseTex]) ? 1u : 0u;   // material.cpp:77
        s.alphaTest = m.alphaTest;
        anyAlpha |= m.alphaTest != 0;
        ctx->materials.push_back(m);
    }
    if (anyAlpha && !md->texcoords) { delete hm; return setError(ctx, PRT_ERR_ARG, "add_mesh: alpha-tested material needs texcoords"); }

    // Mesh::calculateBounds (mesh.cpp:300-309), Scene::add (scene.cpp:19-27)
    for (int k = 0; k < 3; k++) { hm->lower[k] = FLT_MAX; hm->upper[k] = -FLT_MAX; }
    for (uint32_t v = 0; v < md->vertexCount; v++)
        for (int k = 0; k < 3; k++) { hm->lower[k] = std::fmin(hm->lower[k], md->positions[3 * v + k]); hm->upper[k] = std::fmax(hm->upper[k], md->positions[3 * v + k]); }
    for (int k = 0; k < 3; k++) { ctx->bboxLo[k] = std::fmin(ctx->bboxLo[k], hm->lower[k]); ctx->bboxHi[k] = std::fmax(ctx->bboxHi[k], hm->upper[k]); }
    {
        float c[3], d[3];
        for (int k = 0; k < 3; k++) { c[k] = 0.5f * (ctx->bboxHi[k] + ctx->bboxLo[k]); d[k] = ctx->bboxHi[k] - c[k]; }   // vecmath.h:1108
        float dd[3] = {d[0] * d[0], d[1] * d[1], d[2] * d[2]};
        ctx->scene.radius = sqrtf(dd[0] + dd[1] + dd[2]);
    }

    for (uint32_t i = 0; i < md->primCount; i++)      // the reference indexes its material table unchecked (mesh.cpp:362); e.g. OBJ faces before any usemtl carry id -1
        if (md->primMaterial[i] >= md->materialCount) { delete hm; return setError(ctx, PRT_ERR_ARG, "add_mesh: primMaterial[%u] = %u is outside the %u materials", i, md->primMaterial[i], md->materialCount); }
    buildBvh(hm->bvh, md->indices, md->primCount, md->positions, 0);
    if (hm->bvh.maxDepth >= (uint32_t)kStackSize) { delete hm; return setError(ctx, PRT_ERR_STATE, "add_mesh: BVH depth %u exceeds the traversal stack (%d), as the reference asserts (bvh.cpp:552)", hm->bvh.maxDepth, kStackSize); }

    // Leaf-ordered triangle records: tri[i] = primitive remap[i] (the reference copies the same positions into
    // its TriangleVector SoA, bvh.cpp:256-270).
    const uint32_t P = md->primCount;
    std::vector<float4> tris(3 * (size_t)P);   // 48-byte records; 64-byte records read with two 256-bit loads measured 1-2 % slower (bigger footprint)
    std::vector<float2> triUv;
    if (anyAlpha) triUv.resize(3 * (size_t)P);
    // De-indexed shading records (DScene::triShade): the vertex normals / texcoords Mesh::getSurfaceProperties gathers through
    // indices[3*prim + k] (mesh.cpp:318-356), copied per leaf-ordered triangle.
    std::vector<float4> shade;
    if (md->normals || md->texcoords) shade.resize(4 * (size_t)P);
    for (uint32_t i = 0; i < P; i++) {
        uint32_t prim = hm->bvh.remap[i];
        uint32_t mat = md->primMaterial[prim];
        const DMaterial& dmat = ctx->materials[materialBase + mat];
        for (int j = 0; j < 3; j++) {
            uint32_t v = md->indices[3 * (size_t)prim + j];
            float4 f;
            f.x = md->positions[3 * (size_t)v]; f.y = md->positions[3 * (size_t)v + 1]; f.z = md->positions[3 * (size_t)v + 2];
            uint32_t w = j == 0 ? prim : (j == 1 ? (dmat.alphaTest ? (uint32_t)dmat.diffuseTex + 1u : 0u) : materialBase + mat);
            memcpy(&f.w, &w, 4);
            tris[3 * (size_t)i + j] = f;
            if (anyAlpha) triUv[3 * (size_t)i + j] = make_float2(md->texcoords[2 * (size_t)v], md->texcoords[2 * (size_t)v + 1]);
        }
        if (!shade.empty()) {
            float n[3][3] = {}, t[3][2] = {};
            for (int j = 0; j < 3; j++) {
                uint32_t v = md->indices[3 * (size_t)prim + j];
                if (md->normals) memcpy(n[j], md->normals + 3 * (size_t)v, 12);
                if (md->texcoords) memcpy(t[j], md->texcoords + 2 * (size_t)v, 8);
            }
            const uint32_t flags = (md->normals ? 1u : 0u) | (md->texcoords ? 2u : 0u);
            float ff; memcpy(&ff, &flags, 4);
            float4* o = &shade[4 * (size_t)i];
            o[0] = make_float4(n[0][0], n[0][1], n[0][2], t[0][0]);
            o[1] = make_float4(n[1][0], n[1][1], n[1][2], t[0][1]);
            o[2] = make_float4(n[2][0], n[2][1], n[2][2], t[1][0]);
            o[3] = make_float4(t[1][1], t[2][0], t[2][1], ff);
        }
    }

    DMesh& d = hm->d;
    d.rootRef = hm->bvh.rootRef;
    memcpy(d.rootLo, hm->bvh.rootLower, 12); memcpy(d.rootHi, hm->bvh.rootUpper, 12);
    if (ctx->device >= 0) {
        int r = 0;
        auto& own = hm->deviceAllocs;
        std::vector<void*> staging;                  // moved into the scene-wide arrays by the next commitScene
        const Node64* dn = nullptr; const float4* dt = nullptr; const float2* du = nullptr;
        r |= uploadVec(ctx, staging, hm->bvh.nodes64.data(), hm->bvh.nodes64.size(), &dn); hm->d_nodes = (void*)dn;
        r |= uploadVec(ctx, staging, tris.data(), tris.size(), &dt); hm->d_tris = (void*)dt;
        r |= uploadVec(ctx, staging, triUv.data(), triUv.size(), &du); hm->d_triUv = (void*)du;
        const float4* ds = nullptr;
        r |= uploadVec(ctx, staging, shade.data(), shade.size(), &ds); hm->d_triShade = (void*)ds;
        if (r) { for (auto p : own) cudaFree(p); for (auto p : staging) cudaFree(p); delete hm; return PRT_ERR_CUDA; }
        cudaError_t e = cudaStreamSynchronize(ctx->stream);   // host staging vectors die at scope exit
        if (e != cudaSuccess) { delete hm; return setError(ctx, PRT_ERR_CUDA, "add_mesh upload: %s", cudaGetErrorString(e)); }
    }
    ctx->meshes.push_back(hm);
    ctx->dirty = true;
    return (int)meshId;
}

int prt_cuda_scene_set_directional_light(prt_ctx* ctx, const float dir[3], const float intensity[3]) {   // scene.h:30-35
    if (!ctx || !dir || !intensity) return setError(ctx, PRT_ERR_ARG, "set_directional_light: bad argument");
    memcpy(ctx->scene.lightDir, dir, 12); memcpy(ctx->scene.lightIntensity, intensity, 12);
    ctx->scene.lightType = 1;
    return PRT_OK;
}

int prt_cuda_scene_set_infinite_light(prt_ctx* ctx, int w, int h, const float* rgba) {                    // light.cpp:30-82
    if (!ctx || !rgba || w <= 1 || h <= 1 || w > 65535 || h > 65535) return setError(ctx, PRT_ERR_ARG, "set_infinite_light: bad argument");
    HostLight& l = ctx->light;
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
    ctx->scene.infW = w; ctx->scene.infH = h;
    ctx->scene.lightType = 2;
    if (ctx->device >= 0) {
        int r = 0;
        r |= uploadVec(ctx, ctx->deviceAllocs, rgba, 4 * (size_t)w * h, &ctx->scene.infImage);
        r |= uploadVec(ctx, ctx->deviceAllocs, l.vertP.data(), l.vertP.size(), &ctx->scene.infVertP);
        r |= uploadVec(ctx, ctx->deviceAllocs, l.horiP.data(), l.horiP.size(), &ctx->scene.infHoriP);
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
        r |= uploadVec(ctx, ctx->deviceAllocs, gv.data(), gv.size(), &ctx->scene.infGuideV);
        r |= uploadVec(ctx, ctx->deviceAllocs, gh.data(), gh.size(), &ctx->scene.infGuideH);
        if (r) return PRT_ERR_CUDA;
        PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return PRT_OK;
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
    DCamera& c = ctx->scene.cam;
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
    ctx->hasCamera = true;
    return PRT_OK;
}

int prt_cuda_bvh_node_count(const prt_ctx* ctx, int mesh, uint32_t* nodeCount) {
    if (!ctx || !nodeCount || mesh < 0 || mesh >= (int)ctx->meshes.size()) return PRT_ERR_ARG;
    *nodeCount = (uint32_t)ctx->meshes[mesh]->bvh.nodes32.size();
    return PRT_OK;
}
int prt_cuda_bvh_get(const prt_ctx* ctx, int mesh, uint32_t* nodes32, uint32_t* remap) {
    if (!ctx || mesh < 0 || mesh >= (int)ctx->meshes.size()) return PRT_ERR_ARG;
    const BvhHost& b = ctx->meshes[mesh]->bvh;
    if (nodes32) memcpy(nodes32, b.nodes32.data(), b.nodes32.size() * sizeof(Node32));
    if (remap) memcpy(remap, b.remap.data(), b.remap.size() * 4);
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
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return PRT_OK;
}
int prt_cuda_get_counters(prt_ctx* ctx, prt_counters* out) {
    if (!ctx || !out) return PRT_ERR_ARG;
    if (ctx->device >= 0) { int r = foldWavefrontCounters(ctx); if (r) return r; }
    *out = ctx->counters;
    return PRT_OK;
}
int prt_cuda_reset_counters(prt_ctx* ctx) {
    if (!ctx) return PRT_ERR_ARG;
    if (ctx->device >= 0) foldWavefrontCounters(ctx);
    ctx->counters = prt_counters{};
    return PRT_OK;
}
uint64_t prt_cuda_stream(prt_ctx* ctx) { return ctx ? (uint64_t)(uintptr_t)ctx->stream : 0; }

uint32_t prt_packet_seed(uint32_t seed, uint32_t pixel, uint32_t ppp, uint32_t packet) {
    uint32_t h = seed * 0x9E3779B9u + pixel * ppp + packet;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h | 1u;
}

}  // extern "C"
