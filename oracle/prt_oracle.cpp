// oracle/prt_oracle.cpp — CPU restatement of the reference's ray-tracing hot path.  TEST INFRASTRUCTURE.
//
// Plain scalar C++ (no SIMD, no reference headers), each function citing the reference file:line it
// restates.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load the library
// built from this file (oracle/_build/libprt_oracle.so); prt_b200/ never links or calls it.
//
// PARITY PINNING: this restatement is pinned against (a) the reference's one hot-path known-answer vector
// (tests/tests.cpp:113-126, t == 1.0f exactly) and (b) outputs of the reference itself, executed in the
// build container through oracle/_ref/libprt_ref.so and committed as fixtures under tests/golden/
// (tests/golden/make_golden.py) — see tests/test_oracle_vs_ref.py / test_oracle_golden.py.
//
// Arithmetic contract: IEEE fp32, every multiply/add individually rounded (compiled with
// -ffp-contract=off, the reference is built -mavx2 without -mfma), true divides and square roots, SSE
// min/max NaN semantics (second operand returned when either is NaN).  The reference's 8-lane AVX packet
// code is restated per lane: every SoA operation in the reference is lane-wise, and the only cross-lane
// coupling in packet traversal is "visit the node if ANY lane hits it" — a lane whose own box test fails is
// masked off for the whole subtree (bvh.cpp:480-485, 532-533), so walking the same fixed child order with
// the lane's own box test visits, for that lane, exactly the leaves the packet visits with the lane enabled.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <cfloat>
#include <limits>
#include <vector>
#include <thread>
#include <atomic>
#include <algorithm>

namespace orc {

struct V2 { float x, y; };
struct V3 {
    float x, y, z;
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
static inline V3 v3(float a, float b, float c) { return V3{a, b, c}; }
static inline V3 v3(float a) { return V3{a, a, a}; }
static inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
static inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline V3 operator*(float f, V3 a) { return v3(f * a.x, f * a.y, f * a.z); }   // vecmath.h:280-282
static inline V3 operator*(V3 a, float f) { return v3(f * a.x, f * a.y, f * a.z); }   // vecmath.h:276-278
static inline V2 operator+(V2 a, V2 b) { return V2{a.x + b.x, a.y + b.y}; }
static inline V2 operator-(V2 a, V2 b) { return V2{a.x - b.x, a.y - b.y}; }
static inline V2 operator*(float f, V2 a) { return V2{f * a.x, f * a.y}; }
static inline float dot(V3 a, V3 b) { V3 v = a * b; return v.x + v.y + v.z; }          // vecmath.h:1183-1189
static inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }  // :1191-1194
static inline float length(V3 v) { return sqrtf(dot(v, v)); }                            // :1197-1200
static inline V3 normalize(V3 v) { float inv = 1.0f / length(v); return inv * v; }       // :1202-1206
static inline float dot(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }                    // :1128-1132
static inline float length(V2 v) { return sqrtf(dot(v, v)); }
static inline V2 safeNormalize(V2 v) {                                                    // :1145-1153
    float len = length(v);
    if (len < 0.00001f) return V2{0.0f, 0.0f};
    float inv = 1.0f / len;
    return inv * v;
}
// SSE min_ps / max_ps: "a OP b ? a : b" — the second operand comes back when either is NaN.
static inline float sseMin(float a, float b) { return a < b ? a : b; }
static inline float sseMax(float a, float b) { return a > b ? a : b; }

struct BBox {
    V3 lower, upper;
    static BBox init() { return BBox{v3(FLT_MAX), v3(-FLT_MAX)}; }                       // vecmath.cpp:46-52
    void merge(V3 p) {                                                                    // vecmath.cpp:69-73 (std::fmin/fmax)
        lower = v3(std::fmin(lower.x, p.x), std::fmin(lower.y, p.y), std::fmin(lower.z, p.z));
        upper = v3(std::fmax(upper.x, p.x), std::fmax(upper.y, p.y), std::fmax(upper.z, p.z));
    }
    void merge(const BBox& b) { merge2(b); }
    void merge2(const BBox& b) {                                                          // vecmath.cpp:63-67
        lower = v3(std::fmin(lower.x, b.lower.x), std::fmin(lower.y, b.lower.y), std::fmin(lower.z, b.lower.z));
        upper = v3(std::fmax(upper.x, b.upper.x), std::fmax(upper.y, b.upper.y), std::fmax(upper.z, b.upper.z));
    }
    float surfaceArea() const { V3 e = upper - lower; return 2.0f * (e.x * e.y + e.y * e.z + e.z * e.x); }  // vecmath.cpp:75-79
};

// ------------------------------------------------------------------------------------------------
// Flattened BVH (bvh.h:49-62).  32-byte node: bbox, primOrSecondNodeIndex, {triVectorIndex:26, primCount:4, splitAxis:2}
// ------------------------------------------------------------------------------------------------
struct Node {
    BBox bbox;
    uint32_t primOrSecond;
    uint32_t triVectorIndex;
    uint32_t primCount;   // 0xf = interior
    uint32_t splitAxis;
};
static const uint32_t kInternal = 0xf;
static const int kMaxLeaf = 8;   // kLaneCount on the AVX build (bvh.h:21, platform.h:16-17)

struct Texture {                 // texture.h:13-22
    int width = 0, height = 0, comp = 0, channel = 0;   // channel 0 = u8, 1 = f32
    std::vector<uint8_t> data;
    bool valid() const { return !data.empty(); }
    float onePixel() const { return 0.5f / width + 0.5f / height; }   // texture.h:28 (int -> float division)
};

struct Material {                // material.h:30-46
    V3 diffuse{0, 0, 0}, emissive{0, 0, 0};
    uint32_t refl = 0;
    int diffuseTex = -1, bumpTex = -1;
    bool alphaTest = false;
};

struct Mesh {
    std::vector<uint32_t> idx;   // 3 per prim
    std::vector<V3> pos, nrm;
    std::vector<V2> uv;
    std::vector<uint32_t> primMat;
    std::vector<Material> mats;
    bool hasNormals = false, hasUv = false;
    BBox bbox;
    uint32_t id = 0;
    // BVH
    std::vector<Node> nodes;
    std::vector<uint32_t> remap;
    uint32_t primCount() const { return (uint32_t)(idx.size() / 3); }
};

// ---- builder: BvhBuildNode::build (bvh.cpp:31-171) + buildLinearBvhNodes (bvh.cpp:230-299), sequential ----
struct BuildNode {
    BBox bbox = BBox::init();
    BuildNode* child[2] = {nullptr, nullptr};
    int32_t primIndex = -1;
    int32_t primCount = 0;
    int32_t splitAxis = 0;
};

static V3 centroid(const Mesh& m, uint32_t prim) {   // bvh.cpp:78-87 — ((0+v0)+v1)+v2, then (1.0f/3)*sum
    V3 t = v3(0.0f);
    for (int j = 0; j < 3; j++) t = t + m.pos[m.idx[3 * prim + j]];
    return (1.0f / 3) * t;
}

static void buildRec(BuildNode* n, const Mesh& m, uint32_t* remap, int32_t start, int32_t end) {
    for (int32_t i = start; i <= end; i++)                                               // calculateBounds bvh.cpp:21-29
        for (int j = 0; j < 3; j++) n->bbox.merge(m.pos[m.idx[3 * remap[i] + j]]);
    int32_t primCount = end - start + 1;
    if (primCount <= kMaxLeaf) { n->primIndex = start; n->primCount = primCount; return; }   // :39-48
    const V3 extent = n->bbox.upper - n->bbox.lower;
    const uint32_t kBucketCount = 32;
    float lowestCost = std::numeric_limits<float>::max();
    uint32_t lowestDim = 0;
    int32_t lowestCostSplit = -1;
    for (uint32_t dim = 0; dim < 3; dim++) {                                             // :66-118
        const float splitExtent = extent[dim] == 0.0f ? 0.0001f : extent[dim];
        const float lowerPos = n->bbox.lower[dim];
        uint32_t count[32] = {0};
        BBox bounds[32];
        for (auto& b : bounds) b = BBox::init();
        for (int32_t i = start; i <= end; i++) {
            V3 temp = v3(0.0f);
            BBox pb = BBox::init();
            for (int j = 0; j < 3; j++) { V3 v = m.pos[m.idx[3 * remap[i] + j]]; temp = temp + v; pb.merge(v); }
            temp = (1.0f / 3) * temp;
            int32_t b = static_cast<int32_t>(kBucketCount * (temp[dim] - lowerPos) / splitExtent);
            if (b >= (int32_t)kBucketCount) b = kBucketCount - 1;
            if (b < 0) b = 0;
            count[b]++;
            bounds[b].merge2(pb);
        }
        for (uint32_t i = 0; i < kBucketCount - 1; i++) {
            uint32_t cl = 0, cr = 0;
            BBox bl = BBox::init(), br = BBox::init();
            for (uint32_t j = 0; j <= i; j++) { cl += count[j]; bl.merge2(bounds[j]); }
            for (uint32_t j = i + 1; j < kBucketCount; j++) { cr += count[j]; br.merge2(bounds[j]); }
            float cost = 0.125f + (cl * bl.surfaceArea() + cr * br.surfaceArea());      // 0*inf = NaN never wins (:111-116)
            if (lowestCost > cost) { lowestDim = dim; lowestCost = cost; lowestCostSplit = (int32_t)i; }
        }
    }
    const uint32_t dim = lowestDim;
    const float splitExtent = extent[dim] == 0.0f ? 0.0001f : extent[dim];
    const float lowerPos = n->bbox.lower[dim];
    float splitPos = lowerPos + (lowestCostSplit + 1) * splitExtent / kBucketCount;     // :124
    auto left = [&](int32_t i) { return centroid(m, remap[i])[dim] < splitPos; };
    int32_t cursor;
    for (cursor = start; cursor <= end; cursor++) if (!left(cursor)) break;              // :137-149
    for (int32_t i = cursor + 1; i <= end; i++) if (left(i)) { std::swap(remap[i], remap[cursor]); cursor++; }
    int32_t mid = cursor - 1;
    if (mid <= start || end <= mid) mid = (start + end) / 2;                             // :151-154
    n->splitAxis = (int32_t)dim;
    n->child[0] = new BuildNode; n->child[1] = new BuildNode;
    buildRec(n->child[0], m, remap, start, mid);
    buildRec(n->child[1], m, remap, mid + 1, end);
}

static void flatten(std::vector<Node>& out, BuildNode* n, uint32_t& triVec) {          // bvh.cpp:230-299
    uint32_t me = (uint32_t)out.size();
    out.push_back(Node{});
    out[me].bbox = n->bbox;
    out[me].splitAxis = (uint32_t)n->splitAxis;
    if (n->primIndex == -1) {
        out[me].primCount = kInternal;
        out[me].triVectorIndex = 0;
        flatten(out, n->child[0], triVec);
        out[me].primOrSecond = (uint32_t)out.size();
        flatten(out, n->child[1], triVec);
    } else {
        out[me].primOrSecond = (uint32_t)n->primIndex;
        out[me].primCount = (uint32_t)n->primCount;
        out[me].triVectorIndex = triVec++;
    }
    delete n;
}

static void buildBvh(Mesh& m) {                                                          // Bvh::build bvh.cpp:173-228
    uint32_t n = m.primCount();
    m.remap.resize(n);
    for (uint32_t i = 0; i < n; i++) m.remap[i] = i;
    auto root = new BuildNode;
    buildRec(root, m, m.remap.data(), 0, (int32_t)n - 1);
    m.nodes.clear();
    uint32_t triVec = 0;
    flatten(m.nodes, root, triVec);
}


// ------------------------------------------------------------------------------------------------
// Rays (ray.h:14-71).  axis: which component is moved to z before the shear (0 = swapXZ, 1 = swapYZ, 2 = none).
// ------------------------------------------------------------------------------------------------
struct Ray {
    V3 org, dir, invDir;
    float maxT;
    int axis;
};

// Ray::prepare (ray.h:26-40): longest element of the SIGNED direction (vecmath.h:216-233; abs is commented out).
static inline int axisSigned(V3 d) {
    if (d.x > d.y) return d.x > d.z ? 0 : 2;
    return d.y > d.z ? 1 : 2;
}
// SoaRay::prepare (ray.h:58-71): max_e = max_ps(|x|, max_ps(|y|,|z|)) (vecmath.h:1032-1038); swapXZ where
// max_e == |x|; swapYZ where max_e == |y| and not swapXZ.
static inline int axisAbs(V3 d) {
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    float m = sseMax(ax, sseMax(ay, az));
    if (m == ax) return 0;
    if (m == ay) return 1;
    return 2;
}
static inline Ray makeRay(V3 org, V3 dir, float maxT, bool absRule) {
    Ray r;
    r.org = org; r.dir = dir; r.maxT = maxT;
    r.invDir = v3(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
    r.axis = absRule ? axisAbs(dir) : axisSigned(dir);
    return r;
}

// ---- the three slab tests traversal uses ----
struct Slab { float maxT0, minT1; };
// 4-wide versions: lanes x,y,z + a 4th lane w; reduceMax4/Min4 = OP(OP(x,z), OP(y,w)) (vecmath.h:1325-1345).
static inline Slab slab4(const BBox& b, const Ray& r, float lowW, float upW, float invW) {
    float a0 = (b.lower.x - r.org.x) * r.invDir.x, a1 = (b.upper.x - r.org.x) * r.invDir.x;
    float b0 = (b.lower.y - r.org.y) * r.invDir.y, b1 = (b.upper.y - r.org.y) * r.invDir.y;
    float c0 = (b.lower.z - r.org.z) * r.invDir.z, c1 = (b.upper.z - r.org.z) * r.invDir.z;
    float w0 = (lowW - 0.0f) * invW, w1 = (upW - 0.0f) * invW;
    float t0x = sseMin(a0, a1), t0y = sseMin(b0, b1), t0z = sseMin(c0, c1), t0w = sseMin(w0, w1);
    float t1x = sseMax(a0, a1), t1y = sseMax(b0, b1), t1z = sseMax(c0, c1), t1w = sseMax(w0, w1);
    Slab s;
    s.maxT0 = sseMax(sseMax(t0x, t0z), sseMax(t0y, t0w));
    s.minT1 = sseMin(sseMin(t1x, t1z), sseMin(t1y, t1w));
    return s;
}
// float BBox::intersect(org,dir,invDir) (vecmath.h:1402-1424): entry distance (may be negative) or +inf.
static inline float boxEntry(const BBox& b, const Ray& r) {
    Slab s = slab4(b, r, -1.0f, 1.0f, std::numeric_limits<float>::infinity());
    if (s.minT1 < s.maxT0) return std::numeric_limits<float>::infinity();
    return s.maxT0;
}
// bool BBox::intersect(org,dir,invDir,maxT) (vecmath.h:1449-1466): strict, 4th lane (-FLT_MAX, maxT).
static inline bool boxStrict(const BBox& b, const Ray& r, float maxT) {
    Slab s = slab4(b, r, std::numeric_limits<float>::lowest(), maxT, 1.0f);
    return s.minT1 > s.maxT0;
}
// SoaMask BBox::intersect(org,dir,invDir,maxT) (vecmath.h:1504-1518): per lane, 3 axes, OP(x, OP(y,z)), non-strict.
static inline bool boxSoa(const BBox& b, const Ray& r, float maxT) {
    float a0 = (b.lower.x - r.org.x) * r.invDir.x, a1 = (b.upper.x - r.org.x) * r.invDir.x;
    float b0 = (b.lower.y - r.org.y) * r.invDir.y, b1 = (b.upper.y - r.org.y) * r.invDir.y;
    float c0 = (b.lower.z - r.org.z) * r.invDir.z, c1 = (b.upper.z - r.org.z) * r.invDir.z;
    float t0x = sseMin(a0, a1), t0y = sseMin(b0, b1), t0z = sseMin(c0, c1);
    float t1x = sseMax(a0, a1), t1y = sseMax(b0, b1), t1z = sseMax(c0, c1);
    float maxT0 = sseMax(t0x, sseMax(t0y, t0z));
    float minT1 = sseMin(t1x, sseMin(t1y, t1z));
    return (maxT0 < maxT) && (minT1 >= maxT0);
}

// ---- SoA intersectTriangle, one lane (triangle.cpp:90-166) ----
struct TriHit { float t, i, j, k; };   // t == -1 : no intersection (triangle.h:11)
static inline V3 swz(V3 v, int axis) { return axis == 0 ? v3(v.z, v.y, v.x) : (axis == 1 ? v3(v.x, v.z, v.y) : v); }
static inline TriHit intersectTriangle(const Ray& r, V3 p0, V3 p1, V3 p2) {
    V3 d = r.dir;
    V3 v0 = p0 - r.org, v1 = p1 - r.org, v2 = p2 - r.org;
    d = swz(d, r.axis); v0 = swz(v0, r.axis); v1 = swz(v1, r.axis); v2 = swz(v2, r.axis);   // :101-113
    float v0z = v0.z, v1z = v1.z, v2z = v2.z;
    float inv_dz = 1.0f / d.z;
    V3 invDzD = inv_dz * d;                                                                  // :119
    v0 = v0 - v0z * invDzD; v1 = v1 - v1z * invDzD; v2 = v2 - v2z * invDzD;                  // :122-124
    float e0 = v1.x * v2.y - v1.y * v2.x;
    float e1 = v2.x * v0.y - v2.y * v0.x;
    float e2 = v0.x * v1.y - v0.y * v1.x;
    TriHit h; h.t = -1.0f; h.i = h.j = h.k = 0.0f;
    bool eb = (e0 < 0.0f || e1 < 0.0f || e2 < 0.0f) && (e0 > 0.0f || e1 > 0.0f || e2 > 0.0f);   // :132
    if (eb) return h;
    float det = e0 + e1 + e2;
    bool ok = det != 0.0f;
    float inv_det = 1.0f / det;
    float t = ((e0 * v0z + e1 * v1z + e2 * v2z) * inv_dz) * inv_det;                         // :150-152
    ok = ok && t > 0.0001f;                                                                  // :156 (NaN fails)
    h.i = e0 * inv_det; h.j = e1 * inv_det; h.k = e2 * inv_det;
    if (ok) h.t = t;
    return h;
}

// ---- texture sampling used inside traversal (alpha) and shading ----
static inline void bilinear(float* k, int32_t* ind, int comp, V2 uv, int width, int height) {   // texture.cpp:31-57
    float s = uv.x - std::floor(uv.x);
    float t = uv.y - std::floor(uv.y);
    float xt = std::max(s * width - 0.5f, 0.0f);
    float yt = std::max(t * height - 0.5f, 0.0f);
    int32_t x0 = (int32_t)std::floor(xt), x1 = std::min(x0 + 1, width - 1);
    int32_t y0 = (int32_t)std::floor(yt), y1 = std::min(y0 + 1, height - 1);
    ind[0] = comp * (x0 + y0 * width); ind[1] = comp * (x1 + y0 * width);
    ind[2] = comp * (x0 + y1 * width); ind[3] = comp * (x1 + y1 * width);
    s = xt - x0; t = yt - y0;
    k[0] = (1.0f - s) * (1.0f - t); k[1] = s * (1.0f - t); k[2] = (1.0f - s) * t; k[3] = s * t;
}
static inline bool testAlpha(const Texture& tex, V2 uv) {                                       // texture.cpp:142-156
    float k[4]; int32_t ind[4];
    bilinear(k, ind, tex.comp, uv, tex.width, tex.height);
    float alpha = 0.0f;
    for (int i = 0; i < 4; i++) alpha = alpha + k[i] * tex.data[ind[i] + 3];
    return alpha > 127;
}

struct Hit { float t, i, j, k; uint32_t primId, meshId; };

// Optional traversal counters (per ray class: node records fetched = interior pops of the 2-child device layout are
// derived from these; "nodes" counts reference node pops, "prims" counts triangles of visited leaves) — they define the
// algorithmic bytes per ray of the roofline.  Enabled with orc_stats_enable(1); single-threaded use only.
static bool g_stats = false;
static uint64_t g_cnt[12] = {0};

struct Scene;
static bool alphaPass(const Scene& sc, const Mesh& m, uint32_t prim, const TriHit& th);

// ---- leaves ----
// intersectSingleRay<kNearest> (bvh.cpp:302-368): all lanes tested against the SAME maxT (the hit.t on entry),
// lanes scanned in ascending order, strict "nearestT > t".
static inline void leafNearestSingle(const Scene& sc, const Mesh& m, const Node& n, const Ray& r, Hit& hit) {
    float maxT = hit.t, nearestT = std::numeric_limits<float>::max();
    int nearest = -1; TriHit best{};
    for (uint32_t l = 0; l < n.primCount; l++) {
        uint32_t prim = m.remap[n.primOrSecond + l];
        TriHit th = intersectTriangle(r, m.pos[m.idx[3 * prim]], m.pos[m.idx[3 * prim + 1]], m.pos[m.idx[3 * prim + 2]]);
        if (!(th.t >= 0.0001f && th.t < maxT)) continue;
        if (!alphaPass(sc, m, prim, th)) continue;
        if (nearestT > th.t) { nearestT = th.t; nearest = (int)l; best = th; }
    }
    if (nearest >= 0) { hit.t = nearestT; hit.i = best.i; hit.j = best.j; hit.k = best.k; hit.primId = m.remap[n.primOrSecond + nearest]; hit.meshId = m.id; }
}
// intersectRayPacket<kNearest>, one lane (bvh.cpp:370-427): triangles in ascending order, each against the CURRENT hit.t.
static inline void leafNearestPacket(const Scene& sc, const Mesh& m, const Node& n, const Ray& r, Hit& hit) {
    for (uint32_t l = 0; l < n.primCount; l++) {
        uint32_t prim = m.remap[n.primOrSecond + l];
        TriHit th = intersectTriangle(r, m.pos[m.idx[3 * prim]], m.pos[m.idx[3 * prim + 1]], m.pos[m.idx[3 * prim + 2]]);
        if (!(th.t >= 0.0001f && th.t < hit.t)) continue;
        if (!alphaPass(sc, m, prim, th)) continue;
        hit.t = th.t; hit.i = th.i; hit.j = th.j; hit.k = th.k; hit.primId = prim; hit.meshId = m.id;
    }
}
static inline bool leafOcclude(const Scene& sc, const Mesh& m, const Node& n, const Ray& r) {    // kOcclude variants
    for (uint32_t l = 0; l < n.primCount; l++) {
        uint32_t prim = m.remap[n.primOrSecond + l];
        TriHit th = intersectTriangle(r, m.pos[m.idx[3 * prim]], m.pos[m.idx[3 * prim + 1]], m.pos[m.idx[3 * prim + 2]]);
        if (!(th.t >= 0.0001f && th.t < r.maxT)) continue;
        if (!alphaPass(sc, m, prim, th)) continue;
        return true;
    }
    return false;
}

// ---- Bvh::intersect<SingleRayHitPacket,SingleRayPacket> (bvh.cpp:429-570, SORT_CHILDREN path) ----
static void bvhIntersectSingle(const Scene& sc, const Mesh& m, const Ray& r, Hit& hit, uint64_t* stats) {
    const Node* nodes = m.nodes.data();
    if (!boxStrict(nodes[0].bbox, r, hit.t)) return;                                         // :458
    uint32_t stack[64]; int cur = 0; stack[0] = 0;
    do {
        if (stats) stats[0]++;
        if (g_stats) g_cnt[0]++;
        const Node& n = nodes[stack[cur]];
        if (n.primCount == kInternal) {
            uint32_t i0 = stack[cur] + 1, i1 = n.primOrSecond;
            float t0 = boxEntry(nodes[i0].bbox, r), t1 = boxEntry(nodes[i1].bbox, r);        // :499-500
            bool h0 = t0 < hit.t, h1 = t1 < hit.t;
            if (!h0) { if (h1) { stack[cur] = i1; cur++; } }
            else if (!h1) { stack[cur] = i0; cur++; }
            else if (t0 < t1) { stack[cur] = i1; stack[cur + 1] = i0; cur += 2; }
            else { stack[cur] = i0; stack[cur + 1] = i1; cur += 2; }
        } else {
            if (stats) stats[1] += n.primCount;
            if (g_stats) g_cnt[1] += n.primCount;
            leafNearestSingle(sc, m, n, r, hit);
        }
        cur--;
    } while (cur >= 0);
}

// ---- Pop-time culling study (orc_set_popcull): the reference culls a child against hit.t only when it is PUSHED
// (bvh.cpp:499-521); a pushed node is visited even when hit.t has shrunk below its entry distance since.  Mode 1 skips a
// popped node whose stored entry distance is >= hit.t (what most ray tracers do; NOT provably result-preserving in
// fp32: a triangle's computed t can be below its box's computed entry on a non-dominant axis).  Mode 2 is the
// provably result-preserving rule the CUDA kernels use: along the ray's swizzle axis k (Ray::prepare, ray.h:34) the
// computed t of any accepted triangle is a convex combination of fl(p_k - o_k) * invDir_k over its vertices, up to
// < 10 ulp of T = max(|t0k|, |t1k|) of the enclosing box's slab on that axis (boxes are exact min/max of vertex
// positions, bvh.cpp:19-29; subtraction and multiplication are monotone, 1/dir_k is the same bit pattern in the slab
// test and in intersectTriangle).  So t >= t0k - 2^-19 T for every triangle of the subtree, and a node with
// t0k - 2^-19 T >= hit.t holds nothing the strict "t < hit.t" test could accept.  Mode 0 = the reference.
static int g_popcull = 0;
static inline float axisLowerBound(const BBox& b, const Ray& r) {
    float lo = r.axis == 0 ? b.lower.x : (r.axis == 1 ? b.lower.y : b.lower.z);
    float hi = r.axis == 0 ? b.upper.x : (r.axis == 1 ? b.upper.y : b.upper.z);
    float o = r.axis == 0 ? r.org.x : (r.axis == 1 ? r.org.y : r.org.z);
    float inv = r.axis == 0 ? r.invDir.x : (r.axis == 1 ? r.invDir.y : r.invDir.z);
    float a0 = (lo - o) * inv, a1 = (hi - o) * inv;
    float t0 = fminf(a0, a1), T = fmaxf(fabsf(a0), fabsf(a1));
    if (!(a0 == a0) || !(a1 == a1)) return -std::numeric_limits<float>::infinity();
    return t0 - (T * 1.9073486328125e-06f + 1e-30f);
}
static void bvhIntersectSingleCull(const Scene& sc, const Mesh& m, const Ray& r, Hit& hit) {
    const Node* nodes = m.nodes.data();
    if (!boxStrict(nodes[0].bbox, r, hit.t)) return;
    uint32_t stack[64]; float bound[64]; int cur = 0; stack[0] = 0; bound[0] = -std::numeric_limits<float>::infinity();
    do {
        const uint32_t ni = stack[cur];
        if (bound[cur] >= hit.t) { if (g_stats) g_cnt[9]++; cur--; continue; }
        if (g_stats) g_cnt[0]++;
        const Node& n = nodes[ni];
        if (n.primCount == kInternal) {
            uint32_t i0 = ni + 1, i1 = n.primOrSecond;
            float t0 = boxEntry(nodes[i0].bbox, r), t1 = boxEntry(nodes[i1].bbox, r);
            float b0 = g_popcull == 1 ? t0 : axisLowerBound(nodes[i0].bbox, r), b1 = g_popcull == 1 ? t1 : axisLowerBound(nodes[i1].bbox, r);
            bool h0 = t0 < hit.t, h1 = t1 < hit.t;
            if (!h0) { if (h1) { stack[cur] = i1; bound[cur] = b1; cur++; } }
            else if (!h1) { stack[cur] = i0; bound[cur] = b0; cur++; }
            else if (t0 < t1) { stack[cur] = i1; bound[cur] = b1; stack[cur + 1] = i0; bound[cur + 1] = b0; cur += 2; }
            else { stack[cur] = i0; bound[cur] = b0; stack[cur + 1] = i1; bound[cur + 1] = b1; cur += 2; }
        } else {
            if (g_stats) g_cnt[1] += n.primCount;
            leafNearestSingle(sc, m, n, r, hit);
        }
        cur--;
    } while (cur >= 0);
}

// ---- Bvh::intersect<RayHitPacket,RayPacket>, one lane (bvh.cpp:429-570): fixed order from avgDir signs ----
static void bvhIntersectPacketLane(const Scene& sc, const Mesh& m, const Ray& r, const bool* reverse, Hit& hit) {
    const Node* nodes = m.nodes.data();
    uint32_t stack[64]; int cur = 0; stack[0] = 0;
    do {
        uint32_t ni = stack[cur];
        const Node& n = nodes[ni];
        bool h = boxSoa(n.bbox, r, hit.t);                                                    // :480-485 (lane's own mask)
        if (g_stats) { g_cnt[2]++; if (h && n.primCount != kInternal) g_cnt[3] += n.primCount; }
        if (!h) {
        } else if (n.primCount == kInternal) {
            if (reverse[n.splitAxis]) { stack[cur] = ni + 1; stack[cur + 1] = n.primOrSecond; }   // :523-529
            else { stack[cur] = n.primOrSecond; stack[cur + 1] = ni + 1; }
            cur += 2;
        } else {
            leafNearestPacket(sc, m, n, r, hit);
        }
        cur--;
    } while (cur >= 0);
}

// ---- Bvh::occluded (bvh.cpp:576-657): unordered DFS, box test at pop against ray.maxT ----
static bool bvhOccluded(const Scene& sc, const Mesh& m, const Ray& r, bool packetRules) {
    const Node* nodes = m.nodes.data();
    uint32_t stack[64]; int cur = 0; stack[0] = 0;
    do {
        uint32_t ni = stack[cur];
        const Node& n = nodes[ni];
        bool h = packetRules ? boxSoa(n.bbox, r, r.maxT) : boxStrict(n.bbox, r, r.maxT);   // :601 / :605
        if (g_stats) { g_cnt[4]++; if (h && n.primCount != kInternal) g_cnt[5] += n.primCount; }
        if (!h) {
        } else if (n.primCount == kInternal) {
            stack[cur] = n.primOrSecond; stack[cur + 1] = ni + 1; cur += 2;                   // :617-618
        } else {
            if (leafOcclude(sc, m, n, r)) return true;
        }
        cur--;
    } while (cur >= 0);
    return false;
}


// ------------------------------------------------------------------------------------------------
// Scene (scene.h:11-72, scene.cpp), lights (light.cpp), camera (camera.h/.cpp), RNG (random.h)
// ------------------------------------------------------------------------------------------------
struct InfiniteLight {                      // light.cpp:30-82
    int width = 0, height = 0;
    std::vector<float> image;               // RGBA32F
    std::vector<float> vertP, horiP;
    bool valid() const { return width > 0; }
};

struct Camera {                             // camera.h:17-36
    V3 pos, dir, up, right;
    uint32_t width = 0, height = 0;
    float invW = 0, invH = 0;
    void create(V3 p, V3 d, uint32_t w, uint32_t h) {
        pos = p; dir = normalize(d); width = w; height = h; invW = 1.0f / w; invH = 1.0f / h;
        V3 u = v3(0, 1.0f, 0);
        V3 r = cross(d, u);                                      // un-normalised dir, as the reference
        if (length(r) < 0.00001f) r = cross(d, v3(1, 0, 0));
        r = normalize(r);
        u = normalize(cross(r, d));
        up = u; right = r;
    }
};

struct Rng {                                // random.h:12-51
    uint32_t a;
    uint32_t next() { uint32_t x = a; x ^= x << 13; x ^= x >> 17; x ^= x << 5; return a = x; }
    float generate() { uint32_t u = (next() & 0x007fffffu) | 0x3f800000u; float f; memcpy(&f, &u, 4); return f - 1.0f; }
    float generateMinus1to1() { return 2.0f * generate() - 1.0f; }
};

static inline uint32_t fmix32(uint32_t h) { h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16; return h; }
// Per-packet seed — identical in ref_harness.cpp and the CUDA generate kernel.
static inline uint32_t packetSeed(uint32_t seed, uint32_t pixel, uint32_t ppp, uint32_t packet) {
    return fmix32(seed * 0x9E3779B9u + pixel * ppp + packet) | 1u;
}

struct Scene {
    std::vector<Mesh*> meshes;
    std::vector<Texture> textures;
    bool hasDirectional = false, hasInfinite = false;
    V3 lightDir{0, 0, 0}, lightIntensity{0, 0, 0};
    InfiniteLight inf;
    BBox bbox = BBox::init();
    float radius = std::numeric_limits<float>::max();          // scene.cpp:16
    Camera camera;
    void add(Mesh* m) {                                         // scene.cpp:19-27
        m->id = (uint32_t)meshes.size();
        meshes.push_back(m);
        bbox.merge2(m->bbox);
        V3 center = 0.5f * (bbox.upper + bbox.lower);          // vecmath.h:1108
        radius = length(bbox.upper - center);
    }
};

static bool alphaPass(const Scene& sc, const Mesh& m, uint32_t prim, const TriHit& th) {   // bvh.cpp:328-340 / 400-410
    const Material& mat = m.mats[m.primMat[prim]];
    if (!mat.alphaTest) return true;
    V2 uv0 = m.uv[m.idx[3 * prim]], uv1 = m.uv[m.idx[3 * prim + 1]], uv2 = m.uv[m.idx[3 * prim + 2]];
    V2 uv = th.i * uv0 + th.j * uv1 + th.k * uv2;
    return testAlpha(sc.textures[mat.diffuseTex], uv);
}

// Scene::intersect (scene.cpp:47-63): all BVHs in insertion order sharing hit.t; miss = t stayed at maxT.
static Hit sceneIntersectSingle(const Scene& sc, const Ray& r, uint64_t* stats = nullptr) {
    if (g_stats) g_cnt[6]++;
    Hit hit; memset(&hit, 0, sizeof(hit)); hit.t = r.maxT;
    for (auto m : sc.meshes) { if (g_popcull) bvhIntersectSingleCull(sc, *m, r, hit); else bvhIntersectSingle(sc, *m, r, hit, stats); }
    if (hit.t == r.maxT) hit.t = -1.0f;
    return hit;
}
static Hit sceneIntersectPacketLane(const Scene& sc, const Ray& r, V3 avgDir) {
    bool rev[3] = {avgDir.x < 0.0f, avgDir.y < 0.0f, avgDir.z < 0.0f};                         // bvh.cpp:444-445
    if (g_stats) g_cnt[7]++;
    Hit hit; memset(&hit, 0, sizeof(hit)); hit.t = r.maxT;
    for (auto m : sc.meshes) bvhIntersectPacketLane(sc, *m, r, rev, hit);
    if (hit.t == r.maxT) hit.t = -1.0f;
    return hit;
}
static bool sceneOccluded(const Scene& sc, const Ray& r, bool packetRules) {                   // scene.cpp:69-94
    if (g_stats) g_cnt[8]++;
    for (auto m : sc.meshes) if (bvhOccluded(sc, *m, r, packetRules)) return true;
    return false;
}

// ---- InfiniteAreaLight::create (light.cpp:30-82) ----
static void buildInfiniteLight(InfiniteLight& l, int w, int h, const float* rgba) {
    l.width = w; l.height = h;
    l.image.assign(rgba, rgba + 4 * (size_t)w * h);
    l.vertP.resize(h); l.horiP.resize((size_t)w * h);
    float vsum = 0.0f;
    for (int y = 0; y < h; y++) {
        float hsum = 0.0f;
        for (int x = 0; x < w; x++) {
            size_t b = x + (size_t)y * w;
            float len = length(v3(rgba[4 * b], rgba[4 * b + 1], rgba[4 * b + 2]));
            l.horiP[b] = len; hsum += len;
        }
        float sinPhi = std::sin(3.14159265358979323846f * (y + 0.5f) / h);
        l.vertP[y] = hsum * sinPhi;
        vsum += hsum * sinPhi;
        float invH = 1.0f / hsum, acc = 0.0f;
        for (int x = 0; x < w; x++) { size_t b = x + (size_t)y * w; float ph = acc + invH * l.horiP[b]; l.horiP[b] = ph; acc = ph; }
    }
    float invV = 1.0f / vsum, acc = 0.0f;
    for (int y = 0; y < h; y++) { float pv = acc + invV * l.vertP[y]; l.vertP[y] = pv; acc = pv; }
}

static V3 sampleTex3f32(const float* p, int comp, int w, int h, V2 uv) {                       // texture.cpp:105-139, C = float
    float k[4]; int32_t ind[4];
    bilinear(k, ind, comp, uv, w, h);
    V3 c = v3(0.0f);
    for (int i = 0; i < 4; i++) c = c + k[i] * v3(p[ind[i]], p[ind[i] + 1], p[ind[i] + 2]);
    return c;
}

// InfiniteAreaLight::sample (light.cpp:84-128).  If the vertical scan never breaks, y == height and the
// reference reads past its table (UB); we clamp the ROW to height-1 in that case and flag it — the sky images
// used by the configs are checked never to trigger it (vertP[H-1] >= largest possible u).
static void sampleInfinite(const InfiniteLight& l, V3& dir, V3& color, V2 u) {
    const float kPi = 3.14159265358979323846f;
    int y; float pdfV = 1.0f, yf = 0.0f;
    for (y = 1; y < l.height; y++) {
        if (l.vertP[y] > u.y) {
            float prev = l.vertP[y - 1], pdf = l.vertP[y] - prev;
            if (pdf == 0.0f) continue;
            pdfV = pdf; yf = y + (u.y - prev) / pdfV - 1.0f; break;
        }
    }
    if (y >= l.height) y = l.height - 1;
    float pdfH = 1.0f, xf = 0.0f;
    for (int x = 1; x < l.width; x++) {
        if (l.horiP[x + (size_t)y * l.width] > u.x) {
            float prev = l.horiP[x - 1 + (size_t)y * l.width], pdf = l.horiP[x + (size_t)y * l.width] - prev;
            if (pdf == 0.0f) continue;
            pdfH = pdf; xf = x + (u.x - prev) / pdfH - 1.0f; break;
        }
    }
    V2 uv{xf / l.width, yf / l.height};
    color = sampleTex3f32(l.image.data(), 4, l.width, l.height, uv);
    float d0 = pdfH * pdfV;
    color = v3(color.x / d0, color.y / d0, color.z / d0);                                      // Vector3f / Vector3f(float)
    float d1 = (float)(l.width * l.height);
    color = v3(color.x / d1, color.y / d1, color.z / d1);
    float theta = 2.0f * kPi * (uv.x + 0.5f);
    float phi = kPi * uv.y;
    float sinPhi = std::sin(phi);
    dir = normalize(v3(std::cos(theta) * sinPhi, std::cos(phi), std::sin(theta) * sinPhi));
}

// ---- Mesh::getSurfaceProperties (mesh.cpp:311-372) ----
struct SurfProps { V3 normal; int mesh; uint32_t material; V2 uv, duv01, duv02; V3 dp01, dp02; };
static void getSurfaceProperties(const Scene& sc, SurfProps& p, const Hit& hit) {
    const Mesh& m = *sc.meshes[hit.meshId];
    uint32_t b = hit.primId * 3;
    uint32_t v0 = m.idx[b], v1 = m.idx[b + 1], v2 = m.idx[b + 2];
    V3 p0 = m.pos[v0], p1 = m.pos[v1], p2 = m.pos[v2];
    if (m.hasNormals) p.normal = normalize(hit.i * m.nrm[v0] + hit.j * m.nrm[v1] + hit.k * m.nrm[v2]);
    else p.normal = normalize(cross(p1 - p0, p2 - p0));
    V2 t0, t1, t2;
    if (m.hasUv) { t0 = m.uv[v0]; t1 = m.uv[v1]; t2 = m.uv[v2]; }
    else { t0 = V2{0, 0}; t1 = V2{1.0f, 0}; t2 = V2{0, 1.0f}; }
    p.uv = hit.i * t0 + hit.j * t1 + hit.k * t2;
    p.mesh = (int)hit.meshId; p.material = m.primMat[hit.primId];
    p.dp01 = normalize(p1 - p0); p.dp02 = normalize(p2 - p0);
    p.duv01 = safeNormalize(t1 - t0); p.duv02 = safeNormalize(t2 - t0);
}

static float sampleTex1(const Texture& t, V2 uv) {                                             // sample<float> (texture.cpp:105-139)
    float k[4]; int32_t ind[4];
    bilinear(k, ind, t.comp, uv, t.width, t.height);
    float c = 0.0f;
    if (t.channel == 0) { for (int i = 0; i < 4; i++) c = c + k[i] * t.data[ind[i]]; return (1.0f / 255.0f) * c; }
    const float* p = (const float*)t.data.data();
    for (int i = 0; i < 4; i++) c = c + k[i] * p[ind[i]];
    return c;
}
static V3 sampleTex3(const Texture& t, V2 uv) {
    if (t.channel == 1) return sampleTex3f32((const float*)t.data.data(), t.comp, t.width, t.height, uv);
    float k[4]; int32_t ind[4];
    bilinear(k, ind, t.comp, uv, t.width, t.height);
    V3 c = v3(0.0f);
    for (int i = 0; i < 4; i++) c = c + k[i] * v3(t.data[ind[i]], t.data[ind[i] + 1], t.data[ind[i] + 2]);
    return (1.0f / 255.0f) * c;
}
struct MatRef { int mesh; uint32_t material; };
static const Material& mat(const Scene& sc, MatRef r) { return sc.meshes[r.mesh]->mats[r.material]; }

static V3 sampleDiffuse(const Scene& sc, const Material& m, V2 uv) {                           // material.cpp:87-96
    V3 color = m.diffuse;
    if (m.diffuseTex >= 0) {
        V3 c = sampleTex3(sc.textures[m.diffuseTex], uv);
        color = color * v3(powf(c.x, 2.2f), powf(c.y, 2.2f), powf(c.z, 2.2f));               // degamma, material.cpp:24-28
    }
    return color;
}
static V3 sampleBump(const Scene& sc, const Material& m, const SurfProps& p) {                 // material.cpp:98-114
    V3 normal = p.normal;
    if (m.bumpTex >= 0) {
        const Texture& t = sc.textures[m.bumpTex];
        float b = sampleTex1(t, p.uv);
        float b01 = sampleTex1(t, p.uv + t.onePixel() * p.duv01) - b;
        float b02 = sampleTex1(t, p.uv + t.onePixel() * p.duv02) - b;
        float nk = 4.0f;
        normal = normalize(normal + nk * b01 * p.dp01 + nk * b02 * p.dp02);
    }
    return normal;
}

// ---- Camera::GenerateJitteredRayPacket (camera.cpp:35-73) ----
static void generatePacket(const Camera& c, Rng& rnd, uint32_t x, uint32_t y, Ray* rays, V3& avgDir) {
    V3 avg = v3(0.0f);
    const float kScreenScale = 0.6f;
    const float kAspect = (float)c.width / c.height;
    const float kScaleX = 0.5f * c.invW, kScaleY = 0.5f * c.invH;
    for (int i = 0; i < 8; i++) {
        float dx = rnd.generateMinus1to1() * kScaleX;
        float dy = rnd.generateMinus1to1() * kScaleY;
        float nx = 2.0f * (x * c.invW - 0.5f + dx) * kScreenScale * kAspect;
        float ny = -2.0f * (y * c.invH - 0.5f + dy) * kScreenScale;
        V3 d = normalize(nx * c.right + ny * c.up + c.dir);
        avg = avg + d;
        rays[i] = makeRay(c.pos, d, 100000.0f, true);
    }
    avgDir = v3(avg.x / 8.0f, avg.y / 8.0f, avg.z / 8.0f);                                    // avgDir/RayPacket::kSize
}
static Ray generateSingle(const Camera& c, Rng& rnd, uint32_t x, uint32_t y) {                 // camera.cpp:12-33
    const float kScreenScale = 0.6f;
    const float kAspect = (float)c.width / c.height;
    const float kScaleX = 0.5f * c.invW, kScaleY = 0.5f * c.invH;
    float dx = rnd.generateMinus1to1() * kScaleX;
    float dy = rnd.generateMinus1to1() * kScaleY;
    float nx = 2.0f * (x * c.invW - 0.5f + dx) * kScreenScale * kAspect;
    float ny = -2.0f * (y * c.invH - 0.5f + dy) * kScreenScale;
    return makeRay(c.pos, normalize(nx * c.right + ny * c.up + c.dir), 100000.0f, false);
}


// ------------------------------------------------------------------------------------------------
// PathTracer::ComputeRadiance(packet) (path_tracer.cpp:77-308) — including its slot-compaction behaviour:
// beta[] is only moved for depth > 4 (:261), result[] is never compacted (:303-307 sums all 8 slots), and
// materials[ci] comes from props[i] (:286-287), which is the NEW hit when ci == i and the path's PREVIOUS
// vertex when ci < i.  With P3 (zero-initialised lightDir/lightIntensity/nextRayDir) a specular path's shadow
// ray contributes beta*0 == 0 whether or not it is occluded, so it is not traced here.
// ------------------------------------------------------------------------------------------------
struct PathStats { uint64_t extendRays = 0, shadowRays = 0; };

static V3 computeRadiancePacket(const Scene& sc, Rng& rnd, const Ray* prim, const Hit* primHit, PathStats* st) {
    const float kPi = 3.14159265358979323846f;
    V3 result[8], beta[8], pos[8], rayDir[8], normals[8];
    SurfProps props[8];
    MatRef materials[8];
    for (int i = 0; i < 8; i++) { beta[i] = v3(1.0f); result[i] = v3(0.0f); }
    uint32_t alive = 0;
    for (int lane = 0; lane < 8; lane++) {                                                    // :97-120
        const Hit& h = primHit[lane];
        if (h.t != -1.0f) {
            getSurfaceProperties(sc, props[alive], h);
            materials[alive] = MatRef{props[alive].mesh, props[alive].material};
            normals[alive] = sampleBump(sc, mat(sc, materials[alive]), props[alive]);
            pos[alive] = h.t * prim[lane].dir + prim[lane].org;
            rayDir[alive] = prim[lane].dir;
            alive++;
        }
    }
    uint32_t depth = 0;
    while (depth < 14) {
        V3 lightDir[8], lightIntensity[8], nextRayDir[8];
        bool shadowable[8];
        for (int i = 0; i < 8; i++) { lightDir[i] = lightIntensity[i] = nextRayDir[i] = v3(0.0f); shadowable[i] = false; }   // P3
        bool directLighting = false;
        for (uint32_t path = 0; path < alive; path++) {                                       // :132-190
            const Material& material = mat(sc, materials[path]);
            V3 normal = normals[path];
            if (material.emissive.x != 0) result[path] = result[path] + beta[path] * material.emissive;
            if (material.refl == 0 || material.refl == 1) {
                float r2 = rnd.generate();
                float r2sq = std::sqrt(r2);
                float r1 = rnd.generate();
                V3 u = (fabsf(normal.x) > 0.1f) ? v3(0, 1.0f, 0.0f) : v3(1.0f, 0, 0);
                V3 tangent = normalize(cross(normal, u));
                V3 binormal = normalize(cross(tangent, normal));
                float theta = 2.0f * kPi * r1;
                V3 diffuseDir = r2sq * std::cos(theta) * binormal + r2sq * std::sin(theta) * tangent + (1 - r2) * normal;
                if (material.refl == 0) {
                    nextRayDir[path] = diffuseDir;
                    beta[path] = beta[path] * sampleDiffuse(sc, material, props[path].uv);
                    if (sc.hasInfinite) {
                        float ux = rnd.generate(); float uy = rnd.generate();                  // P3: x first, then y
                        sampleInfinite(sc.inf, lightDir[path], lightIntensity[path], V2{ux, uy});
                        directLighting = true; shadowable[path] = true;
                    } else if (sc.hasDirectional) {
                        lightDir[path] = sc.lightDir; lightIntensity[path] = sc.lightIntensity;
                        directLighting = true; shadowable[path] = true;
                    }
                } else {
                    V3 rdir = rayDir[path];
                    V3 reflectDir = rdir - normal * 2 * dot(normal, rdir);                    // (normal*2)*dot
                    nextRayDir[path] = 0.9f * reflectDir + 0.1f * diffuseDir;
                }
            }
        }
        const float kFar = 2.0f * sc.radius;
        const float kEpsilon = 0.0008f;
        if (directLighting) {                                                                  // :196-252
            bool grouping = (alive & 0xf) > 2;
            for (uint32_t i = 0; i < alive; i++) {
                if (!shadowable[i]) continue;
                V3 org = grouping ? (pos[i] + v3(kFar) * lightDir[i]) : (pos[i] + kFar * lightDir[i]);
                V3 dir = grouping ? (lightDir[i] * -1.0f) : (-lightDir[i]);
                Ray sr = makeRay(org, dir, kFar - kEpsilon, grouping);
                if (st) st->shadowRays++;
                if (!sceneOccluded(sc, sr, grouping)) {
                    float c = std::max(dot(lightDir[i], normals[i]), 0.0f);
                    V3 li = lightIntensity[i] * c;                                             // Vector3f*float
                    V3 lightRadiance = v3(li.x / kPi, li.y / kPi, li.z / kPi);
                    result[i] = result[i] + beta[i] * lightRadiance;
                }
            }
        }
        uint32_t ci = 0;
        for (uint32_t i = 0; i < alive; i++) {                                                // :255-293
            if (depth > 4) {
                float q = std::max(0.05f, 1.0f - length(beta[i]));
                if (rnd.generate() < q) continue;
                float d = 1.0f - q;
                beta[ci] = v3(beta[i].x / d, beta[i].y / d, beta[i].z / d);
            }
            V3 dir = normalize(nextRayDir[i]);
            Ray ray = makeRay(pos[i], dir, kFar, false);
            if (st) st->extendRays++;
            Hit hit = sceneIntersectSingle(sc, ray);
            if (hit.t != -1.0f) {
                getSurfaceProperties(sc, props[ci], hit);
                pos[ci] = hit.t * ray.dir + ray.org;
                materials[ci] = MatRef{props[i].mesh, props[i].material};                     // :286-287 (index i!)
                normals[ci] = sampleBump(sc, mat(sc, materials[ci]), props[ci]);
                rayDir[ci] = dir;
                ci++;
            }
        }
        if (ci == 0) break;
        alive = ci;
        depth++;
    }
    V3 res = v3(0.0f);
    for (int i = 0; i < 8; i++) res = res + result[i];
    return res;
}

template <typename F>
static void parallelFor(uint32_t n, int nthreads, F&& f) {
    if (nthreads <= 1) { for (uint32_t i = 0; i < n; i++) f(i); return; }
    std::atomic<uint32_t> next{0};
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back([&]() { for (;;) { uint32_t i = next.fetch_add(1); if (i >= n) break; f(i); } });
    for (auto& t : th) t.join();
}

}  // namespace orc

// ------------------------------------------------------------------------------------------------
// C ABI — same shapes as ref_harness.cpp's ref_* so the Python wrapper and the tests are shared.
// Ray record = 8 floats {org, dir, maxT, pad}; hit record = {t,i,j,k, primId, meshId}.
// ------------------------------------------------------------------------------------------------
using namespace orc;

extern "C" {

struct OrcMaterial { float diffuse[3]; float emissive[3]; uint32_t reflType; int32_t diffuseTex; int32_t bumpTex; uint32_t alphaTest; };

void* orc_create() { return new Scene; }
// counters: [0] single-ray node pops [1] single-ray leaf triangles [2] packet-lane node pops [3] packet-lane leaf triangles
// [4] occlusion node pops [5] occlusion leaf triangles [6] single rays [7] packet-lane rays [8] occlusion rays
void orc_stats_enable(int on) { g_stats = on != 0; }
void orc_set_popcull(int mode) { g_popcull = mode; }   // study knob, see bvhIntersectSingleCull; 0 = the reference's traversal
void orc_stats_reset() { for (auto& c : g_cnt) c = 0; }
void orc_stats_get(uint64_t* out) { for (int i = 0; i < 12; i++) out[i] = g_cnt[i]; }

int orc_add_texture(void* h, int w, int hgt, int comp, int channel, const void* texels) {
    auto sc = (Scene*)h;
    Texture t; t.width = w; t.height = hgt; t.comp = comp; t.channel = channel;
    size_t bytes = (size_t)w * hgt * comp * (channel ? 4 : 1);
    t.data.assign((const uint8_t*)texels, (const uint8_t*)texels + bytes);
    sc->textures.push_back(std::move(t));
    return (int)sc->textures.size() - 1;
}

int orc_add_mesh(void* h, uint32_t nprim, uint32_t nvert, const uint32_t* idx, const float* pos, const float* normals,
                 const float* uv, const uint32_t* primMat, uint32_t nmat, OrcMaterial* mats) {
    auto sc = (Scene*)h;
    auto m = new Mesh;
    m->idx.assign(idx, idx + 3 * (size_t)nprim);
    m->pos.resize(nvert); memcpy(m->pos.data(), pos, 12 * (size_t)nvert);
    if (normals) { m->nrm.resize(nvert); memcpy(m->nrm.data(), normals, 12 * (size_t)nvert); m->hasNormals = true; }
    m->uv.resize(nvert);
    if (uv) { memcpy(m->uv.data(), uv, 8 * (size_t)nvert); m->hasUv = true; } else memset(m->uv.data(), 0, 8 * (size_t)nvert);
    m->primMat.assign(primMat, primMat + nprim);
    for (uint32_t i = 0; i < nmat; i++) {
        Material mt;
        mt.diffuse = v3(mats[i].diffuse[0], mats[i].diffuse[1], mats[i].diffuse[2]);
        mt.emissive = v3(mats[i].emissive[0], mats[i].emissive[1], mats[i].emissive[2]);
        mt.refl = mats[i].reflType; mt.diffuseTex = mats[i].diffuseTex; mt.bumpTex = mats[i].bumpTex;
        mt.alphaTest = false;
        if (mt.diffuseTex >= 0) {                                                             // texture.cpp:338-350
            const Texture& t = sc->textures[mt.diffuseTex];
            for (int p = 0; p < t.width * t.height; ++p) if (t.data[(size_t)p * 4 + 3] < 255) { mt.alphaTest = true; break; }
        }
        mats[i].alphaTest = mt.alphaTest ? 1u : 0u;
        m->mats.push_back(mt);
    }
    m->bbox = BBox::init();                                                                   // mesh.cpp:300-309
    for (auto& p : m->pos) m->bbox.merge(p);
    buildBvh(*m);
    sc->add(m);
    return (int)sc->meshes.size() - 1;
}

void orc_set_directional_light(void* h, const float* dir, const float* intensity) {           // scene.h:30-35
    auto sc = (Scene*)h;
    sc->lightDir = v3(dir[0], dir[1], dir[2]); sc->lightIntensity = v3(intensity[0], intensity[1], intensity[2]);
    sc->hasDirectional = true; sc->hasInfinite = false;
}
int orc_set_infinite_light(void* h, int w, int hgt, const float* rgba) {                      // scene.h:41-45
    auto sc = (Scene*)h;
    buildInfiniteLight(sc->inf, w, hgt, rgba);
    sc->hasInfinite = true; sc->hasDirectional = false;
    return 0;
}
void orc_get_infinite_light(void* h, float* vertP, float* horiP, float* rgba) {
    auto sc = (Scene*)h;
    memcpy(vertP, sc->inf.vertP.data(), 4 * sc->inf.vertP.size());
    memcpy(horiP, sc->inf.horiP.data(), 4 * sc->inf.horiP.size());
    if (rgba) memcpy(rgba, sc->inf.image.data(), 4 * sc->inf.image.size());
}
void orc_light_sample(void* h, uint32_t n, const float* u, float* dirOut, float* colOut) {
    auto sc = (Scene*)h;
    for (uint32_t i = 0; i < n; i++) {
        V3 d, c; sampleInfinite(sc->inf, d, c, V2{u[2 * i], u[2 * i + 1]});
        memcpy(dirOut + 3 * i, &d, 12); memcpy(colOut + 3 * i, &c, 12);
    }
}
float orc_get_radius(void* h) { return ((Scene*)h)->radius; }
void orc_set_camera(void* h, const float* pos, const float* dir, uint32_t w, uint32_t hgt) {
    ((Scene*)h)->camera.create(v3(pos[0], pos[1], pos[2]), v3(dir[0], dir[1], dir[2]), w, hgt);
}
void orc_get_camera(void* h, float* out) {
    auto& c = ((Scene*)h)->camera;
    memcpy(out, &c.pos, 12); memcpy(out + 3, &c.dir, 12); memcpy(out + 6, &c.up, 12); memcpy(out + 9, &c.right, 12);
    out[12] = c.invW; out[13] = c.invH;
}

static inline Ray rayFrom(const float* r, bool absRule) { return makeRay(v3(r[0], r[1], r[2]), v3(r[3], r[4], r[5]), r[6], absRule); }
static inline void storeHit(float* o, const Hit& h) { o[0] = h.t; o[1] = h.i; o[2] = h.j; o[3] = h.k; memcpy(o + 4, &h.primId, 4); memcpy(o + 5, &h.meshId, 4); }

void orc_intersect_single(void* h, uint32_t n, const float* rays, float* hits, int nthreads) {
    auto sc = (Scene*)h; const uint32_t chunk = 4096;
    parallelFor((n + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t i = b * chunk; i < std::min(n, (b + 1) * chunk); i++) storeHit(hits + 6 * (size_t)i, sceneIntersectSingle(*sc, rayFrom(rays + 8 * (size_t)i, false)));
    });
}
void orc_intersect_packet(void* h, uint32_t npackets, const float* rays, const float* avgDir, float* hits, int nthreads) {
    auto sc = (Scene*)h; const uint32_t chunk = 512;
    parallelFor((npackets + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t p = b * chunk; p < std::min(npackets, (b + 1) * chunk); p++)
            for (int l = 0; l < 8; l++) {
                size_t i = (size_t)p * 8 + l;
                storeHit(hits + 6 * i, sceneIntersectPacketLane(*sc, rayFrom(rays + 8 * i, true), v3(avgDir[3 * p], avgDir[3 * p + 1], avgDir[3 * p + 2])));
            }
    });
}
void orc_occluded_single(void* h, uint32_t n, const float* rays, uint8_t* out, int nthreads) {
    auto sc = (Scene*)h; const uint32_t chunk = 4096;
    parallelFor((n + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t i = b * chunk; i < std::min(n, (b + 1) * chunk); i++) out[i] = sceneOccluded(*sc, rayFrom(rays + 8 * (size_t)i, false), false) ? 1 : 0;
    });
}
void orc_occluded_packet(void* h, uint32_t npackets, const float* rays, const uint8_t* maskBits, uint8_t* out, int nthreads) {
    auto sc = (Scene*)h; const uint32_t chunk = 512;
    parallelFor((npackets + chunk - 1) / chunk, nthreads, [&](uint32_t b) {
        for (uint32_t p = b * chunk; p < std::min(npackets, (b + 1) * chunk); p++) {
            uint8_t m = (uint8_t)~maskBits[p];                                               // inactive lanes report occluded (bvh.cpp:588)
            for (int l = 0; l < 8; l++)
                if ((maskBits[p] >> l) & 1) if (sceneOccluded(*sc, rayFrom(rays + 8 * ((size_t)p * 8 + l), true), true)) m |= (uint8_t)(1u << l);
            out[p] = m;
        }
    });
}

void orc_primary_packets(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t ppp, uint32_t seed, int mode,
                         float* raysOut, float* avgDirOut, float* hitsOut, int nthreads) {
    auto sc = (Scene*)h; uint32_t bw = x1 - x0, bh = y1 - y0, W = sc->camera.width;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) for (uint32_t p = 0; p < ppp; p++) {
            Rng rnd{packetSeed(seed, y * W + x, ppp, p)};
            Ray rays[8]; V3 avg;
            generatePacket(sc->camera, rnd, x, y, rays, avg);
            size_t pk = ((size_t)row * bw + (x - x0)) * ppp + p;
            for (int l = 0; l < 8; l++) {
                if (raysOut) { float* o = raysOut + (pk * 8 + l) * 8; memcpy(o, &rays[l].org, 12); memcpy(o + 3, &rays[l].dir, 12); o[6] = rays[l].maxT; o[7] = 0; }
                if (mode == 0) { Hit ht = sceneIntersectPacketLane(*sc, rays[l], avg); if (hitsOut) storeHit(hitsOut + (pk * 8 + l) * 6, ht); }
            }
            if (avgDirOut) memcpy(avgDirOut + pk * 3, &avg, 12);
        }
    });
}
void orc_primary_single(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed,
                        float* raysOut, float* hitsOut, int nthreads) {
    auto sc = (Scene*)h; uint32_t bw = x1 - x0, bh = y1 - y0, W = sc->camera.width;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) for (uint32_t s = 0; s < spp; s++) {
            Rng rnd{packetSeed(seed, y * W + x, spp, s)};
            Ray r = generateSingle(sc->camera, rnd, x, y);
            size_t ri = ((size_t)row * bw + (x - x0)) * spp + s;
            if (raysOut) { float* o = raysOut + ri * 8; memcpy(o, &r.org, 12); memcpy(o + 3, &r.dir, 12); o[6] = r.maxT; o[7] = 0; }
            Hit ht = sceneIntersectSingle(*sc, r);
            if (hitsOut) storeHit(hitsOut + ri * 6, ht);
        }
    });
}

// TraceBlock body (path_tracer.cpp:23-31) with per-packet seeding; counters[0..1] += extend / shadow rays.
void orc_render(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed, float exposure,
                float* rgbOut, int nthreads) {
    auto sc = (Scene*)h; uint32_t ppp = spp / 8, W = sc->camera.width, bh = y1 - y0;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) {
            V3 color = v3(0.0f);
            for (uint32_t p = 0; p < ppp; p++) {
                Rng rnd{packetSeed(seed, y * W + x, ppp, p)};
                Ray rays[8]; V3 avg; Hit hits[8];
                generatePacket(sc->camera, rnd, x, y, rays, avg);
                for (int l = 0; l < 8; l++) hits[l] = sceneIntersectPacketLane(*sc, rays[l], avg);
                color = color + computeRadiancePacket(*sc, rnd, rays, hits, nullptr);
            }
            float s = (float)spp;
            color = v3(color.x / s, color.y / s, color.z / s);
            V3 px = exposure * color;
            float* o = rgbOut + 3 * ((size_t)y * W + x);
            o[0] = px.x; o[1] = px.y; o[2] = px.z;
        }
    });
}

// The same over a list of tiles (x0,y0,x1,y1 exclusive), one tile per task: the reference's parallel unit (main.cpp:121-158).
void orc_render_tiles(void* h, uint32_t ntiles, const uint32_t* tiles, uint32_t spp, uint32_t seed, float exposure, float* rgbOut, int nthreads) {
    auto sc = (Scene*)h; uint32_t ppp = spp / 8, W = sc->camera.width;
    parallelFor(ntiles, nthreads, [&](uint32_t t) {
        const uint32_t* T = tiles + 4 * (size_t)t;
        for (uint32_t y = T[1]; y < T[3]; y++)
            for (uint32_t x = T[0]; x < T[2]; x++) {
                V3 color = v3(0.0f);
                for (uint32_t p = 0; p < ppp; p++) {
                    Rng rnd{packetSeed(seed, y * W + x, ppp, p)};
                    Ray rays[8]; V3 avg; Hit hits[8];
                    generatePacket(sc->camera, rnd, x, y, rays, avg);
                    for (int l = 0; l < 8; l++) hits[l] = sceneIntersectPacketLane(*sc, rays[l], avg);
                    color = color + computeRadiancePacket(*sc, rnd, rays, hits, nullptr);
                }
                float s = (float)spp;
                color = v3(color.x / s, color.y / s, color.z / s);
                V3 px = exposure * color;
                float* o = rgbOut + 3 * ((size_t)y * W + x);
                o[0] = px.x; o[1] = px.y; o[2] = px.z;
            }
    });
}

// GbufferVisualizer::Trace (gbuffer_visualizer.cpp:28-51), RNG reseeded per pixel; type 0 diffuse, 1/2 normal.
void orc_gbuffer(void* h, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, int type, uint32_t seed, float exposure,
                 float* rgbOut, int nthreads) {
    auto sc = (Scene*)h; uint32_t W = sc->camera.width, bh = y1 - y0;
    parallelFor(bh, nthreads, [&](uint32_t row) {
        uint32_t y = y0 + row;
        for (uint32_t x = x0; x < x1; x++) {
            Rng rnd{packetSeed(seed, y * W + x, 1, 0)};
            Ray r = generateSingle(sc->camera, rnd, x, y);
            Hit ht = sceneIntersectSingle(*sc, r);
            V3 color = v3(0.0f);
            if (ht.t != -1.0f) {
                SurfProps p; getSurfaceProperties(*sc, p, ht);
                const Material& m = mat(*sc, MatRef{p.mesh, p.material});
                if (type == 0) color = sampleDiffuse(*sc, m, p.uv);
                else color = sampleBump(*sc, m, p) * 0.5f + v3(0.5f);
            }
            V3 px = exposure * color;
            float* o = rgbOut + 3 * ((size_t)y * W + x);
            o[0] = px.x; o[1] = px.y; o[2] = px.z;
        }
    });
}

// Image::savePpm payload (image.cpp:60-69): c/(c+1) if tonemap, clamp via fmin(fmax(c,0),1) (vecmath.h:1171-1174),
// powf(c, 1/2.2f), *255, truncation.
int orc_image_ppm(uint32_t w, uint32_t hgt, int tonemap, const float* rgb, uint8_t* out, const char*) {
    size_t n = 3 * (size_t)w * hgt;
    for (size_t i = 0; i < n; i++) {
        float c = rgb[i];
        if (tonemap) c = c / (c + 1.0f);
        c = std::fmin(std::fmax(c, 0.0f), 1.0f);
        c = powf(c, 1 / 2.2f) * 255.0f;
        out[i] = (uint8_t)c;
    }
    return 0;
}
// FP32 -> half exactly as tinyexr's float_to_half_full (tinyexr.h:899-934) does it for Image::saveExr
// (image.cpp:82-139): round half UP on the magnitude (not ties-to-even), denormals flushed on input, NaN -> 0x200.
static uint16_t halfBits(float f) {
    uint32_t u; memcpy(&u, &f, 4);
    uint32_t sign = u >> 31, e = (u >> 23) & 0xffu, m = u & 0x7fffffu;
    uint16_t o = 0;
    if (e == 0) o = 0;
    else if (e == 255) o = (uint16_t)((31u << 10) | (m ? 0x200u : 0u));
    else {
        int ne = (int)e - 127 + 15;
        if (ne >= 31) o = (uint16_t)(31u << 10);
        else if (ne <= 0) {
            if ((14 - ne) <= 24) {
                uint32_t mant = m | 0x800000u;
                o = (uint16_t)((mant >> (14 - ne)) & 0x3ffu);
                if ((mant >> (13 - ne)) & 1u) o++;
            }
        } else {
            o = (uint16_t)(((uint32_t)ne << 10) | (m >> 13));
            if (m & 0x1000u) o++;
        }
    }
    return (uint16_t)(o | (sign << 15));
}
// rgbaOut: what reading the reference's EXR back gives — R,G,B as the half value widened to float, A = 1.
int orc_image_exr(uint32_t w, uint32_t hgt, const float* rgb, float* rgbaOut, const char*) {
    auto widen = [](uint16_t hb) {
        uint32_t s = hb >> 15, e = (hb >> 10) & 31u, m = hb & 0x3ffu; float v;
        if (e == 0) v = std::ldexp((float)m, -24);
        else if (e == 31) { uint32_t u = 0x7f800000u | (m << 13); memcpy(&v, &u, 4); }
        else v = std::ldexp((float)(m | 0x400u), (int)e - 25);
        return s ? -v : v;
    };
    for (size_t i = 0; i < (size_t)w * hgt; i++) {
        for (int c = 0; c < 3; c++) rgbaOut[4 * i + c] = widen(halfBits(rgb[3 * i + c]));
        rgbaOut[4 * i + 3] = 1.0f;
    }
    return 0;
}
void orc_half_bits(uint32_t n, const float* in, uint16_t* out) { for (uint32_t i = 0; i < n; i++) out[i] = halfBits(in[i]); }

// Scene::getSurfaceProperties (scene.cpp:30-45) over hit records (t,i,j,k,primId,meshId as 6 words); out: 16 words per hit in
// prt_surface order (normal, material index, uv, duv01, duv02, dp01, dp02); misses give zeros.
void orc_surface_properties(void* h, uint32_t n, const float* hits, float* out) {
    auto sc = (Scene*)h;
    for (uint32_t i = 0; i < n; i++) {
        const float* hp = hits + 6 * (size_t)i;
        float* o = out + 16 * (size_t)i;
        memset(o, 0, 64);
        if (hp[0] == -1.0f) continue;
        Hit ht; ht.t = hp[0]; ht.i = hp[1]; ht.j = hp[2]; ht.k = hp[3];
        memcpy(&ht.primId, hp + 4, 4); memcpy(&ht.meshId, hp + 5, 4);
        SurfProps p; getSurfaceProperties(*sc, p, ht);
        o[0] = p.normal.x; o[1] = p.normal.y; o[2] = p.normal.z; memcpy(o + 3, &p.material, 4);
        o[4] = p.uv.x; o[5] = p.uv.y; o[6] = p.duv01.x; o[7] = p.duv01.y; o[8] = p.duv02.x; o[9] = p.duv02.y;
        o[10] = p.dp01.x; o[11] = p.dp01.y; o[12] = p.dp01.z; o[13] = p.dp02.x; o[14] = p.dp02.y; o[15] = p.dp02.z;
    }
}

uint32_t orc_bvh_node_count(void* h, int mesh) { return (uint32_t)((Scene*)h)->meshes[mesh]->nodes.size(); }
void orc_bvh_dump(void* h, int mesh, uint32_t* nodesOut, uint32_t* decoded, uint32_t* primRemapOut) {
    auto m = ((Scene*)h)->meshes[mesh];
    for (size_t i = 0; i < m->nodes.size(); i++) {
        const Node& n = m->nodes[i];
        memcpy(nodesOut + 8 * i, &n.bbox, 24);
        nodesOut[8 * i + 6] = n.primOrSecond;
        nodesOut[8 * i + 7] = (n.triVectorIndex & 0x3ffffffu) | (n.primCount << 26) | (n.splitAxis << 30);   // bvh.h:57-59 bitfield order
        decoded[4 * i] = n.triVectorIndex; decoded[4 * i + 1] = n.primCount; decoded[4 * i + 2] = n.splitAxis; decoded[4 * i + 3] = 0;
    }
    memcpy(primRemapOut, m->remap.data(), 4 * m->remap.size());
}

void orc_intersect_triangle_soa(const float* rays, const float* tri, float* tijk) {
    for (int l = 0; l < 8; l++) {
        Ray r = rayFrom(rays + 8 * l, true);
        TriHit t = intersectTriangle(r, v3(tri[0], tri[1], tri[2]), v3(tri[3], tri[4], tri[5]), v3(tri[6], tri[7], tri[8]));
        tijk[4 * l] = t.t; tijk[4 * l + 1] = t.i; tijk[4 * l + 2] = t.j; tijk[4 * l + 3] = t.k;
    }
}

// Mesh::calculateVertexNormals (mesh.cpp:108-149): faces visited from last to first; an accumulation that would
// make the running normal exactly zero-length is skipped.
void orc_calc_vertex_normals(uint32_t nprim, uint32_t nvert, const uint32_t* idx, const float* pos, float* out) {
    std::vector<V3> n(nvert, v3(0.0f));
    const V3* p = (const V3*)pos;
    for (int32_t i = (int32_t)nprim - 1; i >= 0; i--) {
        uint32_t v[3] = {idx[3 * i], idx[3 * i + 1], idx[3 * i + 2]};
        V3 fn = normalize(cross(p[v[1]] - p[v[0]], p[v[2]] - p[v[0]]));
        if (std::isnan(fn.x) || std::isnan(fn.y) || std::isnan(fn.z)) fn = v3(0.0f);
        for (int j = 0; j < 3; j++) { V3 s = n[v[j]] + fn; if (length(s) > 0.0f) n[v[j]] = s; }
    }
    for (uint32_t i = 0; i < nvert; i++) { V3 r = normalize(n[i]); memcpy(out + 3 * i, &r, 12); }
}

}  // extern "C"
