// Device-side data layout and single-ray BVH traversal for sm_100a.
//
// Arithmetic contract: this translation unit is compiled with -fmad=false (no FMA
// contraction), IEEE divides/square roots (nvcc defaults, never --use_fast_math), in the reference's operation
// order, so hit t / barycentrics / box decisions are bit-identical to the reference's AVX path.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace prt_b200 {

constexpr uint32_t kLeafBit = 0x80000000u;
constexpr int kStackSize = 64;            // same capacity as the reference (bvh.cpp:432); depth is checked at build time

struct DMesh {
    uint32_t rootRef;         // scene-wide reference (see DScene::nodes)
    float rootLo[3], rootHi[3];
    uint32_t primBase;        // this mesh's first entry in DScene::primToTri (= its first scene-wide triangle index)
    uint32_t materialBase;    // scene-wide index of the mesh's material 0
    uint32_t fastRootRef;     // fast traversal mode: root of the mesh's Wide8 tree (scene-wide reference into DScene::fastNodes / fastTris)
};

struct DTexture { const void* data; int width, height, comp, channel; };   // texture.h:13-22

struct DMaterial {            // material.h:30-46, the fields the hot path reads
    float diffuse[3];
    float emissive[3];
    uint32_t refl;
    int diffuseTex, bumpTex;
    uint32_t alphaTest;
};

struct DCamera { float pos[3], dir[3], up[3], right[3]; float invW, invH; uint32_t width, height; };   // camera.h:43-50

struct DScene {
    // Geometry of ALL meshes in three scene-wide arrays; child / leaf references inside the node records are scene-wide
    // indices (rebased when the scene is committed).  The base pointers arrive in the kernel's constant bank, so the
    // traversal loop carries no per-mesh pointer in registers and the leaf loop has no dependent pointer load in front
    // of its triangle fetch (the per-mesh pointers cost an LDG per triangle, profiles/README.md).
    const float4* nodes;      // Node64 = 4 x float4: [c0.lo.xyz c0.hi.x][c0.hi.yz c1.lo.xy][c1.lo.z c1.hi.xyz][ref0 ref1 axis pad]
    const float4* tris;       // 3 x float4 per leaf-ordered triangle: (p0, primId) (p1, alphaTex+1 or 0) (p2, scene-wide material index)
    const float2* triUv;      // 3 x float2 per leaf-ordered triangle, or null when no mesh has an alpha-tested material
    // Shading record of the same triangle, 4 x float4: (n0, uv0.x) (n1, uv0.y) (n2, uv1.x) (uv1.y, uv2.x, uv2.y, flags); flags bit 0: the
    // mesh has vertex normals, bit 1: it has texcoords (Mesh buffers, mesh.h:83-98, de-indexed).  A hit carries the triangle's
    // index, so shading reaches everything Mesh::getSurfaceProperties reads with ONE level of dependent loads instead of
    // mesh table -> index buffer -> vertex buffers.  Null when no mesh has normals or texcoords.
    const float4* triShade;
    // primToTri[mesh.primBase + primId] = scene-wide leaf-ordered triangle index of the mesh's primitive primId (inverse of
    // m_primRemapping): turns a public hit record (primId, meshId) back into the internal one (prt_cuda_surface_properties).
    const uint32_t* primToTri;
    unsigned long long* stats;   // counting build (-DPRT_STATS=1): [5][3] traversal counters, see prt_counters; else null
    // Fast traversal mode (opt-in, prt_wide.cuh): Wide8 records (6 x float4 each), the fast tree's own leaf-ordered triangle
    // records (same 3 x float4 layout, but vertex 0's w holds the scene-wide index of the SAME triangle in `tris`, which is what
    // a hit carries) and their uv side records.  Null unless the scene was built with the option on.
    const float4* fastNodes;
    const float4* fastTris;
    const float2* fastTriUv;
    const DMesh* meshes;
    const DMaterial* materials;
    const DTexture* textures;
    uint32_t meshCount;
    uint32_t lightType;       // 0 none, 1 directional, 2 infinite-area
    float lightDir[3], lightIntensity[3];
    const float* infImage;    // RGBA32F
    const float* infVertP;
    const float* infHoriP;
    const uint16_t* infGuideV;   // [257] search bounds for infVertP, see prt_cuda_scene_set_infinite_light
    const uint16_t* infGuideH;   // [infH][257] per row of infHoriP
    int infW, infH;
    float radius;
    uint32_t traceFlags;      // bit 0: node loop yields to the leaf phase when leaf lanes outnumber node lanes; bits 1-2: any-hit child order;
                              // bits 8-15: idle lanes before a warp refills from the queue (prt_trace.cuh)
    DCamera cam;
};

struct Hit { float t, i, j, k; uint32_t tri, meshId; };

// Per-thread traversal counters of the counting build (prt_trace.cuh creates one per lane under PRT_STATS; everywhere
// else the pointer is a compile-time null and the increments vanish).
struct TravStats { uint32_t rays, nodeSteps, leafVisits, triTests, uvLoads; };   // tri: scene-wide leaf-ordered triangle index; the reference's primId is tris[3*tri].w

// RayHit::primId (ray.h:190-194) of a finished hit record; 0 for a miss, as the record is initialised.
__device__ __forceinline__ uint32_t hitPrimId(const DScene& sc, const Hit& h) { return h.t != -1.0f ? __float_as_uint(__ldg(reinterpret_cast<const float*>(sc.tris + 3 * (size_t)h.tri) + 3)) : 0u; }

// Ray with the per-ray precomputation hoisted out of the triangle test: the reference recomputes
// inv_dz = 1/d.z and invDzD = inv_dz*d per triangle from the same inputs (triangle.cpp:115-119).
struct RayD {
    float ox, oy, oz;
    float dx, dy, dz;
    float ix, iy, iz;          // invDir = 1/dir (ray.h:27, :59)
    float maxT;
    int axis;                  // 0: swapXZ, 1: swapYZ, 2: none
    float inv_dz, sx, sy;      // 1/d'.z and (inv_dz*d').xy for the swizzled direction d'
    bool nanSafe;              // any invDir component non-finite: NaNs can reach the slab min/max
};

// SSE min_ps/max_ps semantics (second operand when either is NaN).
__device__ __forceinline__ float sseMin(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float sseMax(float a, float b) { return a > b ? a : b; }

__device__ __forceinline__ int axisSigned(float x, float y, float z) {        // vecmath.h:216-233 via ray.h:34
    if (x > y) return x > z ? 0 : 2;
    return y > z ? 1 : 2;
}
__device__ __forceinline__ int axisAbs(float x, float y, float z) {           // ray.h:63-70, vecmath.h:1032-1038
    float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
    float m = sseMax(ax, sseMax(ay, az));
    if (m == ax) return 0;
    if (m == ay) return 1;
    return 2;
}

__device__ __forceinline__ RayD makeRay(float ox, float oy, float oz, float dx, float dy, float dz, float maxT, bool absRule) {
    RayD r;
    r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz; r.maxT = maxT;
    r.ix = 1.0f / dx; r.iy = 1.0f / dy; r.iz = 1.0f / dz;
    r.axis = absRule ? axisAbs(dx, dy, dz) : axisSigned(dx, dy, dz);
    float sdx = r.axis == 0 ? dz : dx;
    float sdy = r.axis == 1 ? dz : dy;
    float sdz = r.axis == 0 ? dx : (r.axis == 1 ? dy : dz);
    r.inv_dz = 1.0f / sdz;
    r.sx = r.inv_dz * sdx; r.sy = r.inv_dz * sdy;
    const float inf = __int_as_float(0x7f800000);
    r.nanSafe = !(fabsf(r.ix) < inf && fabsf(r.iy) < inf && fabsf(r.iz) < inf);
    return r;
}

// ---- slab tests.  NS = NaN-safe variant (exact SSE select semantics); the fast variant uses
// FMNMX and is only used for rays whose invDir is finite, where no NaN can arise and both agree. ----
template <bool NS> __device__ __forceinline__ float mn(float a, float b) { return NS ? sseMin(a, b) : fminf(a, b); }
template <bool NS> __device__ __forceinline__ float mx(float a, float b) { return NS ? sseMax(a, b) : fmaxf(a, b); }

struct Slab { float t0x, t0y, t0z, t1x, t1y, t1z; };
template <bool NS>
__device__ __forceinline__ Slab slab(const RayD& r, float lx, float ly, float lz, float hx, float hy, float hz) {
    float a0 = (lx - r.ox) * r.ix, a1 = (hx - r.ox) * r.ix;
    float b0 = (ly - r.oy) * r.iy, b1 = (hy - r.oy) * r.iy;
    float c0 = (lz - r.oz) * r.iz, c1 = (hz - r.oz) * r.iz;
    Slab s;
    s.t0x = mn<NS>(a0, a1); s.t0y = mn<NS>(b0, b1); s.t0z = mn<NS>(c0, c1);
    s.t1x = mx<NS>(a0, a1); s.t1y = mx<NS>(b0, b1); s.t1z = mx<NS>(c0, c1);
    return s;
}
// float BBox::intersect(org,dir,invDir) (vecmath.h:1402-1424): 4th lane (-inf,+inf); reduce = OP(OP(x,z),OP(y,w)).
template <bool NS>
__device__ __forceinline__ float boxEntry(const Slab& s) {
    const float inf = __int_as_float(0x7f800000);
    float maxT0 = mx<NS>(mx<NS>(s.t0x, s.t0z), NS ? sseMax(s.t0y, -inf) : s.t0y);
    float minT1 = mn<NS>(mn<NS>(s.t1x, s.t1z), NS ? sseMin(s.t1y, inf) : s.t1y);
    return (minT1 < maxT0) ? inf : maxT0;
}
// bool BBox::intersect(org,dir,invDir,maxT) (vecmath.h:1449-1466): strict; 4th lane t0 = min(-FLT_MAX,maxT), t1 = max(..).
template <bool NS>
__device__ __forceinline__ bool boxStrict(const Slab& s, float maxT) {
    const float lowest = -3.402823466e+38f;
    float t0w = sseMin(lowest, maxT), t1w = sseMax(lowest, maxT);
    float maxT0 = mx<NS>(mx<NS>(s.t0x, s.t0z), mx<NS>(s.t0y, t0w));
    float minT1 = mn<NS>(mn<NS>(s.t1x, s.t1z), mn<NS>(s.t1y, t1w));
    return minT1 > maxT0;
}
// SoaMask BBox::intersect(org,dir,invDir,maxT) (vecmath.h:1504-1518): OP(x,OP(y,z)); returns entry via *entry,
// and "min_t1 >= max_t0"; the caller applies "max_t0 < maxT" (at push and again at pop).
template <bool NS>
__device__ __forceinline__ bool boxSoa(const Slab& s, float* entry) {
    float maxT0 = mx<NS>(s.t0x, mx<NS>(s.t0y, s.t0z));
    float minT1 = mn<NS>(s.t1x, mn<NS>(s.t1y, s.t1z));
    *entry = maxT0;
    return minT1 >= maxT0;
}

// ---- bilinear footprint (texture.cpp:31-57) and alpha test (texture.cpp:142-156) ----
__device__ __forceinline__ void bilinear(float* k, int* ind, int comp, float u, float v, int width, int height) {
    float s = u - floorf(u);
    float t = v - floorf(v);
    float xt = fmaxf(s * (float)width - 0.5f, 0.0f);     // std::max(a, 0.0f): a is never NaN for finite uv
    float yt = fmaxf(t * (float)height - 0.5f, 0.0f);
    int x0 = (int)floorf(xt), x1 = min(x0 + 1, width - 1);
    int y0 = (int)floorf(yt), y1 = min(y0 + 1, height - 1);
    ind[0] = comp * (x0 + y0 * width); ind[1] = comp * (x1 + y0 * width);
    ind[2] = comp * (x0 + y1 * width); ind[3] = comp * (x1 + y1 * width);
    s = xt - (float)x0; t = yt - (float)y0;
    k[0] = (1.0f - s) * (1.0f - t); k[1] = s * (1.0f - t); k[2] = (1.0f - s) * t; k[3] = s * t;
}
__device__ __forceinline__ bool testAlpha(const DTexture& tex, float u, float v) {
    float k[4]; int ind[4];
    bilinear(k, ind, tex.comp, u, v, tex.width, tex.height);
    const uint8_t* p = (const uint8_t*)tex.data;
    float alpha = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; i++) alpha = alpha + k[i] * (float)__ldg(p + ind[i] + 3);
    return alpha > 127.0f;
}

// ---- one triangle: SoA intersectTriangle, one lane (triangle.cpp:90-166) + the acceptance rules of the leaf loops.
// Returns true when the triangle is accepted: t > kEpsilon (triangle.cpp:156), t >= epsilon && t < limit
// (bvh.cpp:323 / :392-395), alpha test (bvh.cpp:328-340 / 400-410).  Closest hit: the record is written. ----
template <bool OCCLUDE, bool FAST = false>
__device__ __forceinline__ bool triTest(const DScene& sc, const DMesh& m, uint32_t meshId, uint32_t tri, const RayD& r, Hit& hit, float tLimit, TravStats* ts = nullptr,
                                        bool skipHidden = false) {
    if (ts) ts->triTests++;
    const float4* tp = (FAST ? sc.fastTris : sc.tris) + 3 * (size_t)tri;
    const float4 A = __ldg(tp), B = __ldg(tp + 1), C = __ldg(tp + 2);
    // fast mode: bit 31 of vertex 1's w marks a triangle that lies under a flat box of the reference's tree — a box its strict
    // single-ray any-hit test (vecmath.h:1449-1466, "minT1 > maxT0") never enters, whatever the ray (prt_cuda.cu, add_mesh)
    if (FAST && OCCLUDE && skipHidden && (__float_as_uint(B.w) >> 31)) return false;
    // translate, then move the dominant axis to z (triangle.cpp:96-113)
    float ax = A.x - r.ox, ay = A.y - r.oy, az = A.z - r.oz;
    float bx = B.x - r.ox, by = B.y - r.oy, bz = B.z - r.oz;
    float cx = C.x - r.ox, cy = C.y - r.oy, cz = C.z - r.oz;
    const bool s0 = r.axis == 0, s1 = r.axis == 1;
    float v0x = s0 ? az : ax, v0y = s1 ? az : ay, v0z = s0 ? ax : (s1 ? ay : az);
    float v1x = s0 ? bz : bx, v1y = s1 ? bz : by, v1z = s0 ? bx : (s1 ? by : bz);
    float v2x = s0 ? cz : cx, v2y = s1 ? cz : cy, v2z = s0 ? cx : (s1 ? cy : cz);
    // shear: v = v - v.z*(inv_dz*d)   (triangle.cpp:119-124); only x,y are used afterwards
    v0x = v0x - v0z * r.sx; v0y = v0y - v0z * r.sy;
    v1x = v1x - v1z * r.sx; v1y = v1y - v1z * r.sy;
    v2x = v2x - v2z * r.sx; v2y = v2y - v2z * r.sy;
    float e0 = v1x * v2y - v1y * v2x;
    float e1 = v2x * v0y - v2y * v0x;
    float e2 = v0x * v1y - v0y * v1x;
    if ((e0 < 0.0f || e1 < 0.0f || e2 < 0.0f) && (e0 > 0.0f || e1 > 0.0f || e2 > 0.0f)) return false;   // :132-134
    float det = e0 + e1 + e2;
    if (!(det != 0.0f)) return false;
    float inv_det = 1.0f / det;
    float t = ((e0 * v0z + e1 * v1z + e2 * v2z) * r.inv_dz) * inv_det;                                  // :150-152
    if (!(t > 0.0001f && t >= 0.0001f && t < (OCCLUDE ? tLimit : hit.t))) return false;
    float bi = e0 * inv_det, bj = e1 * inv_det, bk = e2 * inv_det;
    uint32_t alphaTex = __float_as_uint(B.w) & (FAST ? 0x7fffffffu : 0xffffffffu);
    if (alphaTex) {
        if (ts) ts->uvLoads++;
        const float2* uvp = (FAST ? sc.fastTriUv : sc.triUv) + 3 * (size_t)tri;
        float2 uv0 = __ldg(uvp), uv1 = __ldg(uvp + 1), uv2 = __ldg(uvp + 2);
        float u = bi * uv0.x + bj * uv1.x + bk * uv2.x;
        float v = bi * uv0.y + bj * uv1.y + bk * uv2.y;
        if (!testAlpha(sc.textures[alphaTex - 1], u, v)) return false;
    }
    if (!OCCLUDE) { hit.t = t; hit.i = bi; hit.j = bj; hit.k = bk; hit.tri = FAST ? __float_as_uint(A.w) : tri; hit.meshId = meshId; }
    return true;
}

// ---- leaf: the leaf's triangles in ascending order.  intersectSingleRay<kNearest> (bvh.cpp:302-368) and one lane of
// intersectRayPacket<kNearest> (bvh.cpp:370-427) accept exactly the same triangles: "t >= 1e-4 && t < (entry hit.t)"
// + strict "nearestT > t" with lowest lane winning is the sequential "t < current hit.t" scan.  OCCLUDE returns on
// the first accepted triangle. ----
template <bool OCCLUDE, bool FAST = false>
__device__ __forceinline__ bool leafTest(const DScene& sc, const DMesh& m, uint32_t meshId, uint32_t ref, const RayD& r, Hit& hit, float tLimit, TravStats* ts = nullptr,
                                         bool skipHidden = false) {
    const uint32_t first = ref & 0x0fffffffu;
    const uint32_t count = ((ref >> 28) & 7u) + 1u;
    if (ts) ts->leafVisits++;
    for (uint32_t l = 0; l < count; l++)
        if (triTest<OCCLUDE, FAST>(sc, m, meshId, first + l, r, hit, tLimit, ts, skipHidden) && OCCLUDE) return true;
    return false;
}

struct NodeD { float4 a, b, c; uint32_t ref0, ref1, axis; };
// One 64-byte node record = two 256-bit read-only loads (LDG.E.256, new on sm_100): half the L1 requests of four
// 128-bit loads.  On C5 (tree larger than L2) the L1/TEX pipe was the busiest unit at 83 % with 64 % of its tag
// requests coming from node fetches (profiles/r1/r1_v7_c5_*).
__device__ __forceinline__ NodeD loadNode64(const float4* p) {
    NodeD n;
    float w0, w1, w2, w3;
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(n.a.x), "=f"(n.a.y), "=f"(n.a.z), "=f"(n.a.w), "=f"(n.b.x), "=f"(n.b.y), "=f"(n.b.z), "=f"(n.b.w) : "l"(p));
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(n.c.x), "=f"(n.c.y), "=f"(n.c.z), "=f"(n.c.w), "=f"(w0), "=f"(w1), "=f"(w2), "=f"(w3) : "l"(p + 2));
    n.ref0 = __float_as_uint(w0); n.ref1 = __float_as_uint(w1); n.axis = __float_as_uint(w2);
    return n;
}
__device__ __forceinline__ NodeD loadNode(const DScene& sc, uint32_t idx) { return loadNode64(sc.nodes + 4 * (size_t)idx); }
template <bool NS> __device__ __forceinline__ Slab slab0(const RayD& r, const NodeD& n) { return slab<NS>(r, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y); }
template <bool NS> __device__ __forceinline__ Slab slab1(const RayD& r, const NodeD& n) { return slab<NS>(r, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w); }

// ---- closest hit, single-ray rules: Bvh::intersect<SingleRayHitPacket,SingleRayPacket> (bvh.cpp:429-570) ----
// Root culled with the strict bool test against the current hit.t (:458); children tested with the float test
// and culled against hit.t at PUSH time only (:499-521); nearer child first, child1 first on ties.
template <bool NS>
__device__ __forceinline__ void intersectSingleMesh(const DScene& sc, const DMesh& m, uint32_t meshId, const RayD& r, Hit& hit, uint32_t* stack) {
    {
        Slab s = slab<NS>(r, m.rootLo[0], m.rootLo[1], m.rootLo[2], m.rootHi[0], m.rootHi[1], m.rootHi[2]);
        if (!boxStrict<NS>(s, hit.t)) return;
    }
    int sp = 0;
    uint32_t cur = m.rootRef;
    for (;;) {
        if (cur & kLeafBit) {
            leafTest<false>(sc, m, meshId, cur, r, hit, 0.0f);
            if (sp == 0) return;
            cur = stack[--sp];
            continue;
        }
        NodeD n = loadNode(sc, cur);
        float t0 = boxEntry<NS>(slab0<NS>(r, n));
        float t1 = boxEntry<NS>(slab1<NS>(r, n));
        bool h0 = t0 < hit.t, h1 = t1 < hit.t;
        if (h0 && h1) {
            if (t0 < t1) { stack[sp++] = n.ref1; cur = n.ref0; }
            else { stack[sp++] = n.ref0; cur = n.ref1; }
        } else if (h0) cur = n.ref0;
        else if (h1) cur = n.ref1;
        else { if (sp == 0) return; cur = stack[--sp]; }
    }
}

// ---- closest hit, packet rules, one lane: Bvh::intersect<RayHitPacket,RayPacket> (bvh.cpp:429-570) ----
// The reference tests a node's own box when it is popped ("min_t1 >= max_t0 && max_t0 < hit.t", :480-485).  We
// evaluate the same test when the child is pushed (hit.t can only shrink, so a child failing now also fails at
// pop) and repeat the hit.t half at pop with the stored entry distance — the visited set is identical.
template <bool NS>
__device__ __forceinline__ void intersectPacketMesh(const DScene& sc, const DMesh& m, uint32_t meshId, const RayD& r, uint32_t revMask, Hit& hit,
                                                    uint32_t* stack, float* stackT) {
    {
        Slab s = slab<NS>(r, m.rootLo[0], m.rootLo[1], m.rootLo[2], m.rootHi[0], m.rootHi[1], m.rootHi[2]);
        float e;
        if (!(boxSoa<NS>(s, &e) && e < hit.t)) return;
    }
    int sp = 0;
    uint32_t cur = m.rootRef;
    for (;;) {
        if (cur & kLeafBit) {
            leafTest<false>(sc, m, meshId, cur, r, hit, 0.0f);
        } else {
            NodeD n = loadNode(sc, cur);
            float e0, e1;
            bool h0 = boxSoa<NS>(slab0<NS>(r, n), &e0) && e0 < hit.t;
            bool h1 = boxSoa<NS>(slab1<NS>(r, n), &e1) && e1 < hit.t;
            bool rev = (revMask >> n.axis) & 1u;          // reversed: child1 is visited first (bvh.cpp:523-529)
            uint32_t firstRef = rev ? n.ref1 : n.ref0, secondRef = rev ? n.ref0 : n.ref1;
            bool hf = rev ? h1 : h0, hs = rev ? h0 : h1;
            float es = rev ? e0 : e1;
            if (hs) { stack[sp] = secondRef; stackT[sp] = es; sp++; }
            if (hf) { cur = firstRef; continue; }
        }
        // pop, re-applying "max_t0 < hit.t" with the current hit.t
        for (;;) {
            if (sp == 0) return;
            --sp;
            if (stackT[sp] < hit.t) { cur = stack[sp]; break; }
        }
    }
}

// ---- any hit: Bvh::occluded (bvh.cpp:576-657).  Box test against ray.maxT (constant), so testing at push is
// the same as testing at pop; order is irrelevant for a boolean.  PACKET: non-strict SoA test; else strict bool. ----
template <bool NS, bool PACKET>
__device__ __forceinline__ bool occludedMesh(const DScene& sc, const DMesh& m, uint32_t meshId, const RayD& r, uint32_t* stack) {
    Hit dummy;
    {
        Slab s = slab<NS>(r, m.rootLo[0], m.rootLo[1], m.rootLo[2], m.rootHi[0], m.rootHi[1], m.rootHi[2]);
        float e;
        bool h = PACKET ? (boxSoa<NS>(s, &e) && e < r.maxT) : boxStrict<NS>(s, r.maxT);
        if (!h) return false;
    }
    int sp = 0;
    uint32_t cur = m.rootRef;
    for (;;) {
        if (cur & kLeafBit) {
            if (leafTest<true>(sc, m, meshId, cur, r, dummy, r.maxT)) return true;
            if (sp == 0) return false;
            cur = stack[--sp];
            continue;
        }
        NodeD n = loadNode(sc, cur);
        Slab s0 = slab0<NS>(r, n), s1 = slab1<NS>(r, n);
        float e;
        bool h0 = PACKET ? (boxSoa<NS>(s0, &e) && e < r.maxT) : boxStrict<NS>(s0, r.maxT);
        bool h1 = PACKET ? (boxSoa<NS>(s1, &e) && e < r.maxT) : boxStrict<NS>(s1, r.maxT);
        if (h0 && h1) { stack[sp++] = n.ref1; cur = n.ref0; }
        else if (h0) cur = n.ref0;
        else if (h1) cur = n.ref1;
        else { if (sp == 0) return false; cur = stack[--sp]; }
    }
}

// ---- Scene::intersect (scene.cpp:47-63) / Scene::occluded (scene.cpp:69-94): BVHs in insertion order ----
__device__ __forceinline__ Hit sceneIntersectSingle(const DScene& sc, const RayD& r, uint32_t* stack) {
    Hit hit; hit.t = r.maxT; hit.i = hit.j = hit.k = 0.0f; hit.tri = 0; hit.meshId = 0;
    for (uint32_t mi = 0; mi < sc.meshCount; mi++) {
        if (r.nanSafe) intersectSingleMesh<true>(sc, sc.meshes[mi], mi, r, hit, stack);
        else intersectSingleMesh<false>(sc, sc.meshes[mi], mi, r, hit, stack);
    }
    if (hit.t == r.maxT) hit.t = -1.0f;
    return hit;
}
__device__ __forceinline__ Hit sceneIntersectPacket(const DScene& sc, const RayD& r, uint32_t revMask, uint32_t* stack, float* stackT) {
    Hit hit; hit.t = r.maxT; hit.i = hit.j = hit.k = 0.0f; hit.tri = 0; hit.meshId = 0;
    for (uint32_t mi = 0; mi < sc.meshCount; mi++) {
        if (r.nanSafe) intersectPacketMesh<true>(sc, sc.meshes[mi], mi, r, revMask, hit, stack, stackT);
        else intersectPacketMesh<false>(sc, sc.meshes[mi], mi, r, revMask, hit, stack, stackT);
    }
    if (hit.t == r.maxT) hit.t = -1.0f;
    return hit;
}
template <bool PACKET>
__device__ __forceinline__ bool sceneOccluded(const DScene& sc, const RayD& r, uint32_t* stack) {
    for (uint32_t mi = 0; mi < sc.meshCount; mi++) {
        bool o = r.nanSafe ? occludedMesh<true, PACKET>(sc, sc.meshes[mi], mi, r, stack) : occludedMesh<false, PACKET>(sc, sc.meshes[mi], mi, r, stack);
        if (o) return true;
    }
    return false;
}

}  // namespace prt_b200
