// oracle/ref_harness.cpp — TEST INFRASTRUCTURE, not product code.
//
// C-ABI shim around the UNMODIFIED reference implementation (amada/PRT, sources compiled where
// they lie under /root/reference/src by oracle/Makefile into oracle/_ref/libprt_ref.so).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load the resulting library.  Nothing in prt_b200/ links or calls it.
//
// This file contains no reference code: it *calls* the reference's classes
// (Scene/Mesh/Bvh/Camera/PathTracer/Image, scene.h:11-72, bvh.h:90-120, path_tracer.h:16-38)
// and dumps their inputs/outputs as flat arrays.  It is compiled with -fno-access-control so
// that it can (a) reseed PathTracer::m_rand per packet (SURVEY §8c patch P2 — done from the
// outside, the reference source is untouched), (b) read Bvh::m_nodes / m_primRemapping for the
// builder-parity test, (c) fill Mesh texcoord buffers without an OBJ round trip (P5).
//
// Build flags: -std=c++17 -O3 -mavx2 (8-lane AVX path, platform.h:16-17; no -mfma, no
// fast-math, matching CMakeLists.txt:8,17), -include ref_p1_shim.h (P1: g++ needs a bit-cast for the
// implicit __m256i→__m256 conversions clang accepts in vecmath.h:332,384 — bit-neutral).
#include <cstdio>
#include <cstring>
#include <cstdint>
#include <chrono>
#include <vector>
#include <thread>
#include <atomic>
#include <string>
#include <unistd.h>

#include "vecmath.h"
#include "ray.h"
#include "triangle.h"
#include "mesh.h"
#include "material.h"
#include "texture.h"
#include "bvh.h"
#include "scene.h"
#include "camera.h"
#include "random.h"
#include "image.h"
#include "path_tracer.h"
#include "gbuffer_visualizer.h"
#include "sample_models.h"

// tinyexr is compiled inside the reference's texture.cpp; we only need the declaration.
extern "C" int LoadEXR(float** out_rgba, int* width, int* height, const char* filename, const char** err);
extern "C" int SaveEXR(const float* data, int width, int height, int components,
                       const int save_as_fp16, const char* filename, const char** err);

namespace prt { void convertNormalToBump(uint8_t* bump, const uint8_t* normal, uint32_t width, uint32_t height); }

using namespace prt;

namespace {

struct RefCtx {
    Scene scene;
    Camera camera;
    std::vector<Bvh*> bvhs;
    std::vector<Texture> textures;      // texel storage is owned here (malloc'd copies)
    uint32_t width = 0, height = 0;
};

// Per-packet RNG seed shared by oracle, reference harness and the CUDA path
// (fmix32 of (seed, pixel, packet) | 1 — the xorshift32 state must be non-zero).
inline uint32_t fmix32(uint32_t h) {
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}
inline uint32_t packetSeed(uint32_t seed, uint32_t pixel, uint32_t packetsPerPixel, uint32_t packet) {
    return fmix32(seed * 0x9E3779B9u + pixel * packetsPerPixel + packet) | 1u;
}

// Traversal counters of the reference itself (prt.h:5 PRT_ENABLE_STATS; incremented at bvh.cpp:464-466,555-557).
// Only compiled into libprt_ref_stats.so; they feed the bytes-per-ray constants of the roofline.
std::atomic<uint64_t> g_statRays{0}, g_statNodes{0}, g_statPrims{0};
template <typename HP>
inline void addStats(const HP& hp, uint64_t rays) {
#ifdef PRT_ENABLE_STATS
    g_statRays += rays; g_statNodes += hp.stats.nodesTraversed; g_statPrims += hp.stats.primsTraversed;
#else
    (void)hp; (void)rays;
#endif
}

template <typename F>
void parallelFor(uint32_t n, int nthreads, F&& f) {
    if (nthreads <= 1) { for (uint32_t i = 0; i < n; i++) f(i); return; }
    std::atomic<uint32_t> next{0};
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++)
        th.emplace_back([&]() { for (;;) { uint32_t i = next.fetch_add(1); if (i >= n) break; f(i); } });
    for (auto& t : th) t.join();
}

}  // namespace

extern "C" {

struct RefMaterial {
    float diffuse[3];
    float emissive[3];
    uint32_t reflType;     // ReflectionType (material.h:24-28)
    int32_t diffuseTex;    // texture id or -1
    int32_t bumpTex;       // texture id or -1
    uint32_t alphaTest;    // out: what Material::load would have computed (material.cpp:77)
};

// out: rays, nodesTraversed, primsTraversed (closest-hit calls since the last reset); zeros without PRT_ENABLE_STATS
void ref_stats_get(uint64_t* out) { out[0] = g_statRays; out[1] = g_statNodes; out[2] = g_statPrims; }
void ref_stats_reset() { g_statRays = 0; g_statNodes = 0; g_statPrims = 0; }

void* ref_create() {
    auto c = new RefCtx;
    c->scene.init();
    return c;
}

int ref_lane_count() { return (int)SoaConstants::kLaneCount; }

// channel: 0 = 8-bit unorm, 1 = fp32 (texture.h:8-11)
int ref_add_texture(void* h, int w, int hgt, int comp, int channel, const void* texels) {
    auto c = (RefCtx*)h;
    Texture t;
    t.init();
    t.width = (uint16_t)w; t.height = (uint16_t)hgt; t.format = 0; t.component = (uint8_t)comp;
    t.channel = channel ? TextureChannel::k32Float : TextureChannel::k8Unorm;
    size_t bytes = (size_t)w * hgt * comp * (channel ? 4 : 1);
    t.texels = malloc(bytes);
    memcpy(t.texels, texels, bytes);
    c->textures.push_back(t);
    return (int)c->textures.size() - 1;
}

int ref_add_mesh(void* h, uint32_t nprim, uint32_t nvert, const uint32_t* idx, const float* pos,
                 const float* normals, const float* uv, const uint32_t* primMat, uint32_t nmat,
                 RefMaterial* mats) {
    auto c = (RefCtx*)h;
    Mesh m;
    m.create(nprim, nvert, nmat, normals != nullptr);
    memcpy(m.getIndexBuffer(), idx, sizeof(uint32_t) * 3 * nprim);
    memcpy(m.m_texcoordIndices, idx, sizeof(uint32_t) * 3 * nprim);  // loadObj: texcoord index == vertex index (mesh.cpp:261)
    memcpy((void*)m.getPositionBuffer(), pos, sizeof(float) * 3 * nvert);
    if (normals) memcpy((void*)m.getNormalBuffer(), normals, sizeof(float) * 3 * nvert);
    if (uv) memcpy((void*)m.getTexcoordBuffer(), uv, sizeof(float) * 2 * nvert);
    else memset((void*)m.getTexcoordBuffer(), 0, sizeof(float) * 2 * nvert);
    m.m_hasTexcoord = (uv != nullptr);
    memcpy(m.getPrimMateialBuffer(), primMat, sizeof(uint32_t) * nprim);
    for (uint32_t i = 0; i < nmat; i++) {
        Material mat;
        mat.init();
        mat.diffuse = Vector3f(mats[i].diffuse[0], mats[i].diffuse[1], mats[i].diffuse[2]);
        mat.emissive = Vector3f(mats[i].emissive[0], mats[i].emissive[1], mats[i].emissive[2]);
        mat.reflectionType = (ReflectionType)mats[i].reflType;
        if (mats[i].diffuseTex >= 0) memcpy((void*)&mat.diffuseMap, &c->textures[mats[i].diffuseTex], sizeof(Texture));
        if (mats[i].bumpTex >= 0) memcpy((void*)&mat.bumpMap, &c->textures[mats[i].bumpTex], sizeof(Texture));
        mat.alphaTest = mat.diffuseMap.isAlphaTestRequired();
        mats[i].alphaTest = mat.alphaTest ? 1u : 0u;
        memcpy((void*)&m.getMaterialBuffer()[i], &mat, sizeof(Material));
    }
    m.calculateBounds();
    auto bvh = new Bvh;
    bvh->build(std::move(m));
    c->scene.add(bvh);
    c->bvhs.push_back(bvh);
    return (int)c->bvhs.size() - 1;
}

void ref_set_directional_light(void* h, const float* dir, const float* intensity) {
    auto c = (RefCtx*)h;
    c->scene.setDirectionalLight(Vector3f(dir[0], dir[1], dir[2]), Vector3f(intensity[0], intensity[1], intensity[2]));
}

// The reference can only ingest an infinite-area light from an .exr file (light.cpp:30-32); write
// the caller's RGBA32F lat-long image losslessly (fp32) to a temp file and load it through
// Scene::setInfiniteAreaLight.
int ref_set_infinite_light(void* h, int w, int hgt, const float* rgba) {
    auto c = (RefCtx*)h;
    char path[256];
    snprintf(path, sizeof(path), "/tmp/prt_ref_sky_%d_%p.exr", (int)getpid(), h);
    const char* err = nullptr;
    int r = SaveEXR(rgba, w, hgt, 4, 0, path, &err);
    if (r != 0) { fprintf(stderr, "ref_set_infinite_light: SaveEXR failed %d %s\n", r, err ? err : ""); return -1; }
    c->scene.setInfiniteAreaLight(path);
    unlink(path);
    return 0;
}

// Copies out the light tables the reference built (light.cpp:46-81) and the image as loaded.
void ref_get_infinite_light(void* h, float* vertP, float* horiP, float* rgba) {
    auto c = (RefCtx*)h;
    auto& l = c->scene.m_infiniteAreaLight;
    memcpy(vertP, l.m_verticalP, sizeof(float) * l.m_height);
    memcpy(horiP, l.m_horizontalP, sizeof(float) * l.m_width * l.m_height);
    if (rgba) memcpy(rgba, l.m_image.texels, sizeof(float) * 4 * l.m_width * l.m_height);
}

void ref_light_sample(void* h, uint32_t n, const float* u, float* dirOut, float* colOut) {
    auto c = (RefCtx*)h;
    auto& l = c->scene.getInfiniteAreaLight();
    for (uint32_t i = 0; i < n; i++) {
        Vector3f d, col;
        l.sample(d, col, Vector2f(u[2 * i], u[2 * i + 1]));
        memcpy(dirOut + 3 * i, &d, 12);
        memcpy(colOut + 3 * i, &col, 12);
    }
}

float ref_get_radius(void* h) { return ((RefCtx*)h)->scene.getRadius(); }

void ref_set_camera(void* h, const float* pos, const float* dir, uint32_t w, uint32_t hgt) {
    auto c = (RefCtx*)h;
    c->camera.create(Vector3f(pos[0], pos[1], pos[2]), Vector3f(dir[0], dir[1], dir[2]), w, hgt);
    c->width = w; c->height = hgt;
}

// out: pos, dir, up, right (3 floats each), invW, invH
void ref_get_camera(void* h, float* out) {
    auto& cam = ((RefCtx*)h)->camera;
    memcpy(out + 0, &cam.m_pos, 12); memcpy(out + 3, &cam.m_dir, 12);
    memcpy(out + 6, &cam.m_up, 12);  memcpy(out + 9, &cam.m_right, 12);
    out[12] = cam.m_invWidth; out[13] = cam.m_invHeight;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Part 2: ray batches, renders, BVH dumps.  Ray record = 8 floats {org.xyz, dir.xyz, maxT, pad};
// hit record = {t,i,j,k (float), primId, meshId (u32)} = RayHit (ray.h:182-198).
// ---------------------------------------------------------------------------------------------
extern "C" {

void ref_convert_normal_to_bump(uint8_t* bump, const uint8_t* normalRgba, uint32_t w, uint32_t hgt) {
    prt::convertNormalToBump(bump, normalRgba, w, hgt);
}

static inline void storeHit(float* o, const RayHit& hit) {
    o[0] = hit.t; o[1] = hit.i; o[2] = hit.j; o[3] = hit.k;
    memcpy(o + 4, &hit.primId, 4); memcpy(o + 5, &hit.meshId, 4);
}

static inline void storePacketHits(float* o, const RayHitPacket& hp) {
    const auto& hh = hp.hits[0];
    for (int l = 0; l < 8; l++) {
        RayHit r;
        r.t = hh.t.getLane(l); r.i = hh.i.getLane(l); r.j = hh.j.getLane(l); r.k = hh.k.getLane(l);
        r.primId = (uint32_t)hh.primId.getLane(l); r.meshId = (uint32_t)hh.meshId.getLane(l);
        storeHit(o + 6 * l, r);
    }
}

// Scene::intersect<SingleRayHitPacket,SingleRayPacket> (scene.cpp:47-63) after Ray::prepare (ray.h:26-40)
void ref_intersect_single(void* h, uint32_t n, const float* rays, float* hits, int nthreads) {
    auto c = (RefCtx*)h;
    const uint32_t chunk = 4096;
    parallelFor((n + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t i = b * chunk; i < std::min(n, (b + 1) * chunk); i++) {
            Ray r;
            r.org = Vector3f(rays[8 * i + 0], rays[8 * i + 1], rays[8 * i + 2]);
            r.dir = Vector3f(rays[8 * i + 3], rays[8 * i + 4], rays[8 * i + 5]);
            r.maxT = rays[8 * i + 6];
            r.prepare();
            SingleRayHitPacket hp;
            c->scene.intersect(hp, SingleRayPacket(r));
            addStats(hp, 1);
            storeHit(hits + 6 * i, hp.hit);
        }
    });
}

static inline void makePacket(RayPacket& p, const float* rays, const float* avgDir) {
    Vector3f org[8], dir[8];
    float maxT[8];
    for (int l = 0; l < 8; l++) {
        org[l] = Vector3f(rays[8 * l + 0], rays[8 * l + 1], rays[8 * l + 2]);
        dir[l] = Vector3f(rays[8 * l + 3], rays[8 * l + 4], rays[8 * l + 5]);
        maxT[l] = rays[8 * l + 6];
    }
    p.rays[0].org = SoaVector3f(org);
    p.rays[0].dir = SoaVector3f(dir);
    SoaFloat mt(0.0f);
    for (int l = 0; l < 8; l++) mt = mt.replaceLane(l, maxT[l]);
    p.rays[0].maxT = mt;
    p.rays[0].prepare();
    p.avgDir = avgDir ? Vector3f(avgDir[0], avgDir[1], avgDir[2]) : Vector3f(0.0f);
}

// Scene::intersect<RayHitPacket,RayPacket>; 8 consecutive rays = one packet, avgDir per packet.
void ref_intersect_packet(void* h, uint32_t npackets, const float* rays, const float* avgDir, float* hits, int nthreads) {
    auto c = (RefCtx*)h;
    const uint32_t chunk = 512;
    parallelFor((npackets + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t i = b * chunk; i < std::min(npackets, (b + 1) * chunk); i++) {
            RayPacket p;
            makePacket(p, rays + 64 * (size_t)i, avgDir + 3 * (size_t)i);
            RayHitPacket hp;
            c->scene.intersect(hp, p);
            addStats(hp, 8);
            storePacketHits(hits + 48 * (size_t)i, hp);
        }
    });
}

// Scene::occluded<bool,SingleRayPacket> (scene.cpp:69-94)
void ref_occluded_single(void* h, uint32_t n, const float* rays, uint8_t* out, int nthreads) {
    auto c = (RefCtx*)h;
    const uint32_t chunk = 4096;
    parallelFor((n + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t i = b * chunk; i < std::min(n, (b + 1) * chunk); i++) {
            Ray r;
            r.org = Vector3f(rays[8 * i + 0], rays[8 * i + 1], rays[8 * i + 2]);
            r.dir = Vector3f(rays[8 * i + 3], rays[8 * i + 4], rays[8 * i + 5]);
            r.maxT = rays[8 * i + 6];
            r.prepare();
            out[i] = c->scene.occluded<bool, SingleRayPacket>(RayPacketMask(), r) ? 1 : 0;
        }
    });
}

// Scene::occluded<RayPacketMask,RayPacket>; maskBits[i] = active lanes; out[i] = returned mask bits
// (inactive lanes come back set, exactly as the reference returns them).
void ref_occluded_packet(void* h, uint32_t npackets, const float* rays, const uint8_t* maskBits, uint8_t* out, int nthreads) {
    auto c = (RefCtx*)h;
    const uint32_t chunk = 512;
    parallelFor((npackets + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t i = b * chunk; i < std::min(npackets, (b + 1) * chunk); i++) {
            RayPacket p;
            makePacket(p, rays + 64 * (size_t)i, nullptr);
            auto m = c->scene.occluded<RayPacketMask, RayPacket>(RayPacketMask(maskBits[i]), p);
            out[i] = (uint8_t)(m.masks[0].ballot() & 0xff);
        }
    });
}

// Primary rays exactly as TracePacket generates them (path_tracer.cpp:65-66, camera.cpp:35-73), one
// RNG reseed per packet (P2).  Pixels [y0,y1) x [x0,x1), ppp packets per pixel.  Any output may be null.
// mode 0: packet closest hit; mode 1: rays only.
void ref_primary_packets(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t ppp, uint32_t seed,
                         int mode, float* raysOut, float* avgDirOut, float* hitsOut, int nthreads) {
    auto c = (RefCtx*)h;
    uint32_t bw = x1 - x0, bh = y1 - y0;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        Random rnd;
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) {
            for (uint32_t p = 0; p < ppp; p++) {
                rnd.m_state.a = packetSeed(seed, y * c->width + x, ppp, p);
                auto packet = c->camera.GenerateJitteredRayPacket(rnd, x, y);
                size_t pk = ((size_t)row * bw + (x - x0)) * ppp + p;
                if (raysOut) {
                    for (int l = 0; l < 8; l++) {
                        float* o = raysOut + (pk * 8 + l) * 8;
                        auto org = packet.rays[0].org.getLane(l);
                        auto dir = packet.rays[0].dir.getLane(l);
                        o[0] = org.x; o[1] = org.y; o[2] = org.z; o[3] = dir.x; o[4] = dir.y; o[5] = dir.z;
                        o[6] = packet.rays[0].maxT.getLane(l); o[7] = 0.0f;
                    }
                }
                if (avgDirOut) memcpy(avgDirOut + pk * 3, &packet.avgDir, 12);
                if (mode == 0) {
                    RayHitPacket hp;
                    c->scene.intersect(hp, packet);
                    if (hitsOut) storePacketHits(hitsOut + pk * 48, hp);
                }
            }
        }
    });
}

// Single jittered primary rays (camera.cpp:12-33): one ray per pixel per sample, RNG reseeded per
// (pixel, sample) with the same hash; single-ray closest hit.
void ref_primary_single(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed,
                        float* raysOut, float* hitsOut, int nthreads) {
    auto c = (RefCtx*)h;
    uint32_t bw = x1 - x0, bh = y1 - y0;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        Random rnd;
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) {
            for (uint32_t s = 0; s < spp; s++) {
                rnd.m_state.a = packetSeed(seed, y * c->width + x, spp, s);
                auto ray = c->camera.GenerateJitteredRay(rnd, x, y);
                size_t ri = ((size_t)row * bw + (x - x0)) * spp + s;
                if (raysOut) {
                    float* o = raysOut + ri * 8;
                    o[0] = ray.org.x; o[1] = ray.org.y; o[2] = ray.org.z; o[3] = ray.dir.x; o[4] = ray.dir.y; o[5] = ray.dir.z;
                    o[6] = ray.maxT; o[7] = 0.0f;
                }
                SingleRayHitPacket hp;
                c->scene.intersect(hp, SingleRayPacket(ray));
                if (hitsOut) storeHit(hitsOut + ri * 6, hp.hit);
            }
        }
    });
}

// Full render of pixels [y0,y1) x [x0,x1): the reference's TraceBlock body (path_tracer.cpp:23-31) with
// per-packet reseeding; each packet goes through the reference's own TracePacket(…, samples = 8) →
// GenerateJitteredRayPacket + Scene::intersect + ComputeRadiance.  rgbOut is the full-image RGB32F
// buffer (Image layout, image.cpp:44-50); only the requested region is written.
void ref_render(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed,
                float exposure, float* rgbOut, int nthreads) {
    auto c = (RefCtx*)h;
    uint32_t ppp = spp / RayPacket::kSize;   // integer division as path_tracer.cpp:65
    uint32_t W = c->width;
    uint32_t bh = y1 - y0;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        PathTracer tracer;
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) {
            Vector3f color(0);
            for (uint32_t p = 0; p < ppp; p++) {
                tracer.m_rand.m_state.a = packetSeed(seed, y * W + x, ppp, p);
                color = color + tracer.TracePacket(c->camera, c->scene, x, y, RayPacket::kSize, nullptr);
            }
            color = color / spp;                 // path_tracer.cpp:28
            auto px = exposure * color;          // image.cpp:45
            float* o = rgbOut + 3 * ((size_t)y * W + x);
            o[0] = px.x; o[1] = px.y; o[2] = px.z;
        }
    });
}

// The same per-pixel work over a list of tiles (x0,y0,x1,y1 exclusive, 4 words each), one tile per task — the reference's own
// parallel unit (raytrace_scene queues one TraceBlock per 16x16 tile on its pool, main.cpp:121-158).  Lets a bounded sample
// be spread over the whole frame while keeping every host thread busy.
void ref_render_tiles(void* h, uint32_t ntiles, const uint32_t* tiles, uint32_t spp, uint32_t seed, float exposure, float* rgbOut, int nthreads) {
    auto c = (RefCtx*)h;
    uint32_t ppp = spp / RayPacket::kSize;
    uint32_t W = c->width;
    parallelFor(ntiles, nthreads, [&](uint32_t t) {
        PathTracer tracer;
        const uint32_t* T = tiles + 4 * (size_t)t;
        for (uint32_t y = T[1]; y < T[3]; y++)
            for (uint32_t x = T[0]; x < T[2]; x++) {
                Vector3f color(0);
                for (uint32_t p = 0; p < ppp; p++) {
                    tracer.m_rand.m_state.a = packetSeed(seed, y * W + x, ppp, p);
                    color = color + tracer.TracePacket(c->camera, c->scene, x, y, RayPacket::kSize, nullptr);
                }
                color = color / spp;
                auto px = exposure * color;
                float* o = rgbOut + 3 * ((size_t)y * W + x);
                o[0] = px.x; o[1] = px.y; o[2] = px.z;
            }
    });
}

// G-buffer pass: GbufferVisualizer::TraceBlock body (gbuffer_visualizer.cpp:17-26) with the visualiser's RNG reseeded
// per pixel (same hash as a 1-spp single primary ray, so the camera ray is ref_primary_single's).  type: 0 diffuse,
// 1 mesh normal, 2 normal (gbuffer_visualizer.h:17-21).
void ref_gbuffer(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, int type, uint32_t seed, float exposure,
                 float* rgbOut, int nthreads) {
    auto c = (RefCtx*)h;
    uint32_t W = c->width, bh = y1 - y0;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        GbufferVisualizer vis((GbufferVisualizer::Type)type);
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) {
            vis.m_rand.m_state.a = packetSeed(seed, y * W + x, 1, 0);
            Vector3f color = vis.Trace(c->camera, c->scene, x, y);
            auto px = exposure * color;          // image.cpp:45
            float* o = rgbOut + 3 * ((size_t)y * W + x);
            o[0] = px.x; o[1] = px.y; o[2] = px.z;
        }
    });
}

// Scene::getSurfaceProperties<SurfaceProperties, RayHit> (scene.cpp:30-45) over hit records; out: 16 words per hit in
// prt_surface order (normal, material index within the mesh, uv, duv01, duv02, dp01, dp02); misses give zeros.
void ref_surface_properties(void* h, uint32_t n, const float* hits, float* out) {
    auto c = (RefCtx*)h;
    for (uint32_t i = 0; i < n; i++) {
        const float* hp = hits + 6 * (size_t)i;
        float* o = out + 16 * (size_t)i;
        memset(o, 0, 64);
        if (hp[0] == -1.0f) continue;
        RayHit ht; ht.t = hp[0]; ht.i = hp[1]; ht.j = hp[2]; ht.k = hp[3];
        memcpy(&ht.primId, hp + 4, 4); memcpy(&ht.meshId, hp + 5, 4);
        SurfaceProperties p;
        c->scene.getSurfaceProperties(p, ht);
        const Mesh& mesh = c->bvhs[ht.meshId]->m_mesh;
        uint32_t mat = (uint32_t)(p.material - &mesh.getMaterial(0));
        o[0] = p.normal.x; o[1] = p.normal.y; o[2] = p.normal.z; memcpy(o + 3, &mat, 4);
        o[4] = p.uv.x; o[5] = p.uv.y; o[6] = p.duv01.x; o[7] = p.duv01.y; o[8] = p.duv02.x; o[9] = p.duv02.y;
        o[10] = p.dp01.x; o[11] = p.dp01.y; o[12] = p.dp01.z; o[13] = p.dp02.x; o[14] = p.dp02.y; o[15] = p.dp02.z;
    }
}

// Output stage through the reference's own Image class: fills Image::m_pixels with rgb (exposure already applied),
// calls savePpm / saveExr (image.cpp:52-139) on a scratch file and returns what was written: the PPM payload bytes,
// or the EXR's pixels read back with tinyexr's LoadEXR as RGBA floats (exactly representable halfs).
int ref_image_ppm(uint32_t w, uint32_t hgt, int tonemap, const float* rgb, uint8_t* out, const char* scratchPath) {
    Image img(w, hgt, tonemap != 0, 1.0f);
    memcpy(img.m_pixels, rgb, sizeof(float) * 3 * (size_t)w * hgt);
    img.savePpm(scratchPath);
    FILE* fp = fopen(scratchPath, "rb");
    if (!fp) return -1;
    char line[64];
    for (int i = 0; i < 3; i++) if (!fgets(line, sizeof line, fp)) { fclose(fp); return -2; }     // "P6", "W H", "255"
    size_t n = fread(out, 1, 3 * (size_t)w * hgt, fp);
    fclose(fp);
    unlink(scratchPath);
    return n == 3 * (size_t)w * hgt ? 0 : -3;
}
int ref_image_exr(uint32_t w, uint32_t hgt, const float* rgb, float* rgbaOut, const char* scratchPath) {
    Image img(w, hgt, true, 1.0f);
    memcpy(img.m_pixels, rgb, sizeof(float) * 3 * (size_t)w * hgt);
    img.saveExr(scratchPath);
    float* rgba = nullptr; int rw = 0, rh = 0; const char* err = nullptr;
    if (LoadEXR(&rgba, &rw, &rh, scratchPath, &err) != 0 || rw != (int)w || rh != (int)hgt) return -1;
    memcpy(rgbaOut, rgba, sizeof(float) * 4 * (size_t)w * hgt);
    free(rgba);
    unlink(scratchPath);
    return 0;
}

// BVH dump for builder parity: walks the DFS array from the root to count nodes.
static uint32_t countNodes(const LinearBvhNode* nodes) {
    // DFS layout: subtree of node i occupies [i, end); walk iteratively.
    uint32_t maxIndex = 0;
    std::vector<uint32_t> stack{0};
    while (!stack.empty()) {
        uint32_t i = stack.back(); stack.pop_back();
        if (i > maxIndex) maxIndex = i;
        if (nodes[i].primCount == LinearBvhNode::kInternalNode) {
            stack.push_back(i + 1);
            stack.push_back(nodes[i].primOrSecondNodeIndex);
        }
    }
    return maxIndex + 1;
}

uint32_t ref_bvh_node_count(void* h, int mesh) { return countNodes(((RefCtx*)h)->bvhs[mesh]->m_nodes); }

// nodesOut: nodeCount x 8 words {lower.xyz, upper.xyz, primOrSecondNodeIndex, packed bitfield}
// decoded[4*i..]: {triVectorIndex, primCount, splitAxis, 0}; primRemapOut: primCount entries.
void ref_bvh_dump(void* h, int mesh, uint32_t* nodesOut, uint32_t* decoded, uint32_t* primRemapOut) {
    auto b = ((RefCtx*)h)->bvhs[mesh];
    uint32_t n = countNodes(b->m_nodes);
    static_assert(sizeof(LinearBvhNode) == 32, "");
    memcpy(nodesOut, b->m_nodes, 32 * (size_t)n);
    for (uint32_t i = 0; i < n; i++) {
        decoded[4 * i + 0] = b->m_nodes[i].triVectorIndex;
        decoded[4 * i + 1] = b->m_nodes[i].primCount;
        decoded[4 * i + 2] = b->m_nodes[i].splitAxis;
        decoded[4 * i + 3] = 0;
    }
    memcpy(primRemapOut, b->m_primRemapping, sizeof(uint32_t) * b->m_mesh.getPrimCount());
}

// SoA ray-triangle test on one triangle, 8 rays (triangle.cpp:90-166) — used for the reference's
// single known-answer vector (tests/tests.cpp:113-126) and for kernel unit parity.
void ref_intersect_triangle_soa(const float* rays /*8x8*/, const float* tri /*9*/, float* tijk /*8x4*/) {
    Vector3f org[8], dir[8];
    for (int l = 0; l < 8; l++) {
        org[l] = Vector3f(rays[8 * l], rays[8 * l + 1], rays[8 * l + 2]);
        dir[l] = Vector3f(rays[8 * l + 3], rays[8 * l + 4], rays[8 * l + 5]);
    }
    SoaRay r;
    r.org = SoaVector3f(org); r.dir = SoaVector3f(dir); r.maxT = SoaFloat(1e30f);
    r.prepare();
    SoaMask mask; mask.setAll(true);
    auto res = intersectTriangle(mask, r.org, r.dir, r.swapXZ, r.swapYZ,
        SoaVector3f(Vector3f(tri[0], tri[1], tri[2])), SoaVector3f(Vector3f(tri[3], tri[4], tri[5])), SoaVector3f(Vector3f(tri[6], tri[7], tri[8])));
    for (int l = 0; l < 8; l++) {
        tijk[4 * l] = res.t.getLane(l); tijk[4 * l + 1] = res.i.getLane(l);
        tijk[4 * l + 2] = res.j.getLane(l); tijk[4 * l + 3] = res.k.getLane(l);
    }
}

// ---- fixture generation helpers (used only by tests/golden/make_golden.py in the build container) ----

// Mesh::loadObj(path, mat) (mesh.cpp:151-209) + the setupCornellBox transform (main.cpp:40-46).
// Call with null outputs to query counts.
int ref_load_obj(const char* path, float scale, const float* translate, uint32_t* nprimOut, uint32_t* nvertOut,
                 uint32_t* idxOut, float* posOut, float* normalsOut) {
    Material mat; mat.init();
    Mesh m;
    m.loadObj(path, mat);
    if (m.getVertexCount() == 0) return -1;
    *nprimOut = m.getPrimCount(); *nvertOut = m.getVertexCount();
    if (!idxOut) return 0;
    auto pos = m.getPositionBuffer();
    for (uint32_t i = 0; i < m.getVertexCount(); i++)
        pos[i] = scale * pos[i] + Vector3f(translate[0], translate[1], translate[2]);
    m.calculateVertexNormals();
    memcpy(idxOut, m.getIndexBuffer(), sizeof(uint32_t) * 3 * m.getPrimCount());
    memcpy(posOut, (void*)m.getPositionBuffer(), 12 * (size_t)m.getVertexCount());
    memcpy(normalsOut, (void*)m.getNormalBuffer(), 12 * (size_t)m.getVertexCount());
    return 0;
}

// Mesh::loadObj(path) (mesh.cpp:212-298: tinyobjloader + Material::load + Texture::load) dumped to a flat binary file
// for the loader-parity test.  Layout (little endian): u32 nprim, nvert, nmat; indices[3*nprim] u32; positions[3*nvert]
// f32; texcoords[2*nvert] f32; primMaterial[nprim] u32; per material: diffuse[3], emissive[3] f32, u32 reflType, u32
// alphaTest, then two textures (diffuseMap, bumpMap): u32 width, height, component, then width*height*component bytes
// (0,0,0 when the map is absent).
int ref_load_obj_dump(const char* objPath, const char* outPath) {
    Mesh m;
    m.loadObj(objPath);
    if (m.getVertexCount() == 0) return -1;
    FILE* f = fopen(outPath, "wb");
    if (!f) return -2;
    uint32_t hdr[3] = {m.getPrimCount(), m.getVertexCount(), m.getMaterialCount()};
    fwrite(hdr, 4, 3, f);
    fwrite(m.getIndexBuffer(), 4, 3 * (size_t)hdr[0], f);
    fwrite((void*)m.getPositionBuffer(), 4, 3 * (size_t)hdr[1], f);
    fwrite((void*)m.m_texcoords, 4, 2 * (size_t)hdr[1], f);
    fwrite(m.getPrimMateialBuffer(), 4, hdr[0], f);
    for (uint32_t i = 0; i < hdr[2]; i++) {
        const Material& mt = m.getMaterialBuffer()[i];
        float v[6] = {mt.diffuse.x, mt.diffuse.y, mt.diffuse.z, mt.emissive.x, mt.emissive.y, mt.emissive.z};
        uint32_t u[2] = {(uint32_t)mt.reflectionType, mt.alphaTest ? 1u : 0u};
        fwrite(v, 4, 6, f); fwrite(u, 4, 2, f);
        for (const Texture* t : {&mt.diffuseMap, &mt.bumpMap}) {
            uint32_t d[3] = {0, 0, 0};
            if (t->isValid()) { d[0] = t->width; d[1] = t->height; d[2] = t->component; }
            fwrite(d, 4, 3, f);
            if (t->isValid()) fwrite(t->texels, 1, (size_t)d[0] * d[1] * d[2], f);
        }
    }
    fclose(f);
    return 0;
}

// Mesh::calculateVertexNormals (mesh.cpp:108-149) on caller data.
void ref_calc_vertex_normals(uint32_t nprim, uint32_t nvert, const uint32_t* idx, const float* pos, float* normalsOut) {
    Mesh m;
    m.create(nprim, nvert, 1, false);
    memcpy(m.getIndexBuffer(), idx, sizeof(uint32_t) * 3 * nprim);
    memcpy((void*)m.getPositionBuffer(), pos, 12 * (size_t)nvert);
    m.calculateVertexNormals();
    memcpy(normalsOut, (void*)m.getNormalBuffer(), 12 * (size_t)nvert);
}

// SampleModels::getCornellBox(true) (sample_models.cpp:13-209): 36 prims / 72 verts / 18 materials.
void ref_cornell_box(uint32_t* counts /*nprim,nvert,nmat*/, uint32_t* idx, float* pos, uint32_t* primMat, RefMaterial* mats) {
    Mesh m = SampleModels::getCornellBox(true);
    counts[0] = m.getPrimCount(); counts[1] = m.getVertexCount(); counts[2] = m.getMaterialCount();
    if (!idx) return;
    memcpy(idx, m.getIndexBuffer(), sizeof(uint32_t) * 3 * m.getPrimCount());
    memcpy(pos, (void*)m.getPositionBuffer(), 12 * (size_t)m.getVertexCount());
    memcpy(primMat, m.getPrimMateialBuffer(), sizeof(uint32_t) * m.getPrimCount());
    for (uint32_t i = 0; i < m.getMaterialCount(); i++) {
        auto& s = m.getMaterial(i);
        memcpy(mats[i].diffuse, &s.diffuse, 12); memcpy(mats[i].emissive, &s.emissive, 12);
        mats[i].reflType = (uint32_t)s.reflectionType; mats[i].diffuseTex = -1; mats[i].bumpTex = -1; mats[i].alphaTest = 0;
    }
}

}  // extern "C"
