// Device-side shading inputs: Mesh::getSurfaceProperties, Material::sampleDiffuse/sampleBump, texture
// sampling, InfiniteAreaLight::sample, xorshift32 — in the reference's fp32 operation order (-fmad=false).
//
// libm calls: the reference calls glibc sinf/cosf (direction sampling — any last-bit difference sends the path
// somewhere else) and powf (texture de-gamma).  Both are restated bit-exactly: sinf/cosf below, powf in prt_powf.h.
#pragma once
#include "prt_device.cuh"
#include "prt_powf.h"

namespace prt_b200 {

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 f3(float x, float y, float z) { F3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ F3 operator*(F3 a, F3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ F3 operator*(float f, F3 a) { return f3(f * a.x, f * a.y, f * a.z); }
__device__ __forceinline__ F3 operator*(F3 a, float f) { return f3(f * a.x, f * a.y, f * a.z); }
__device__ __forceinline__ F3 divs(F3 a, float d) { return f3(a.x / d, a.y / d, a.z / d); }          // Vector3f / Vector3f(float)
__device__ __forceinline__ float dot3(F3 a, F3 b) { float x = a.x * b.x, y = a.y * b.y, z = a.z * b.z; return x + y + z; }   // vecmath.h:1183
__device__ __forceinline__ F3 cross3(F3 a, F3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ float length3(F3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ F3 normalize3(F3 a) { float inv = 1.0f / length3(a); return inv * a; }    // vecmath.h:1200-1204
__device__ __forceinline__ F3 ld3(const float* p) { return f3(__ldg(p), __ldg(p + 1), __ldg(p + 2)); }

// sinf / cosf exactly as glibc 2.39 evaluates them (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf.h — the
// ARM optimized-routines algorithm): fp64 range reduction by pi/2 with a 2^24-prescaled 2/pi, then a degree-7 sine
// or degree-8 cosine polynomial in fp64, rounded once to fp32.  Rounding an accurate fp64 sin/cos to fp32 is NOT the
// same function (1.3 % of arguments differ in the last bit); this restatement matched glibc on 6e7 random arguments
// in [0, 3*pi) with zero mismatches (with and without FMA contraction).  Valid for |x| < 120, which covers every
// call site on the path (theta in [0, 3*pi), phi in [0, pi]).
__device__ __forceinline__ float sincosPoly(double x, double x2, bool negc, int n) {
    if ((n & 1) == 0) {
        const double s1c = -0x1.555545995a603p-3, s2c = 0x1.1107605230bc4p-7, s3c = -0x1.994eb3774cf24p-13;
        double x3 = __dmul_rn(x, x2);
        double s1 = fma(x2, s3c, s2c);
        double x7 = __dmul_rn(x3, x2);
        double sv = fma(x3, s1c, x);
        return (float)fma(x7, s1, sv);
    }
    double c0 = 0x1p0, c1 = -0x1.ffffffd0c621cp-2, c2 = 0x1.55553e1068f19p-5, c3 = -0x1.6c087e89a359dp-10, c4 = 0x1.99343027bf8c3p-16;
    if (negc) { c0 = -c0; c1 = -c1; c2 = -c2; c3 = -c3; c4 = -c4; }
    double x4 = __dmul_rn(x2, x2);
    double cc2 = fma(x2, c4, c3);
    double cc1 = fma(x2, c1, c0);
    double x6 = __dmul_rn(x4, x2);
    double c = fma(x4, c2, cc1);
    return (float)fma(x6, cc2, c);
}
__device__ __forceinline__ uint32_t absTop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ffu; }
template <bool COS>
__device__ __forceinline__ float sincosRef(float y) {
    double x = (double)y;
    if (absTop12(y) < absTop12(0x1.921FB6p-1f)) {
        double x2 = __dmul_rn(x, x);
        if (absTop12(y) < absTop12(0x1p-12f)) return COS ? 1.0f : y;
        return sincosPoly(x, x2, false, COS ? 1 : 0);
    }
    double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
    int n = (__double2int_rz(r) + 0x800000) >> 24;
    x = fma(-(double)n, 0x1.921FB54442D18p0, x);
    double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    return sincosPoly(__dmul_rn(x, sgn), __dmul_rn(x, x), (n & 2) != 0, COS ? (n ^ 1) : n);
}
__device__ __forceinline__ float sinRef(float x) { return sincosRef<false>(x); }
__device__ __forceinline__ float cosRef(float x) { return sincosRef<true>(x); }
__device__ __forceinline__ float powRef(float x, float y) { return prt_powf::powfRef(x, y); }     // glibc powf, bit-exact (prt_powf.h)

__device__ __forceinline__ uint32_t xs32(uint32_t x) { x ^= x << 13; x ^= x >> 17; x ^= x << 5; return x; }   // random.h:41-48
__device__ __forceinline__ float rngFloat(uint32_t x) { return __uint_as_float((x & 0x007fffffu) | 0x3f800000u) - 1.0f; }   // random.h:23-33
__device__ __forceinline__ uint32_t xsAdvance(uint32_t s, uint32_t n) { for (uint32_t k = 0; k < n; k++) s = xs32(s); return s; }

struct Surf {                 // SurfaceProperties (mesh.h:19-28)
    F3 normal;
    uint32_t matRef;          // global material index (mesh materialBase + primMaterial)
    float u, v;
    float duv01x, duv01y, duv02x, duv02y;
    F3 dp01, dp02;
};

__device__ __forceinline__ void safeNormalize2(float x, float y, float& ox, float& oy) {             // vecmath.h:1145-1153
    float len = sqrtf(x * x + y * y);
    if (len < 0.00001f) { ox = 0.0f; oy = 0.0f; return; }
    float inv = 1.0f / len;
    ox = inv * x; oy = inv * y;
}

__device__ __forceinline__ F3 xyz(float4 v) { return f3(v.x, v.y, v.z); }

// props.material of a hit (mesh.cpp:362): scene-wide material index, kept in the triangle record.
__device__ __forceinline__ uint32_t surfaceMaterial(const DScene& sc, const Hit& hit) {
    return __float_as_uint(__ldg(reinterpret_cast<const float*>(sc.tris + 3 * (size_t)hit.tri + 2) + 3));
}

// Mesh::getSurfaceProperties (mesh.cpp:311-372).  dp01/dp02/duv01/duv02 are only ever read by Material::sampleBump when
// the material has a bump map (material.cpp:102-111), so they (and the vertex positions, when the mesh has vertex
// normals) are only fetched/computed when needBumpFrame is set — the fields that are read are computed identically.
// Vertex data comes from the de-indexed per-triangle records (DScene::tris / triShade): the same floats the reference
// reaches through indices[3*primId + k].
__device__ __forceinline__ Surf surfaceProps(const DScene& sc, const Hit& hit, bool needBumpFrame) {
    const float4* tp = sc.tris + 3 * (size_t)hit.tri;
    float4 S0 = make_float4(0, 0, 0, 0), S1 = S0, S2 = S0, S3 = S0;
    if (sc.triShade) { const float4* sp = sc.triShade + 4 * (size_t)hit.tri; S0 = __ldg(sp); S1 = __ldg(sp + 1); S2 = __ldg(sp + 2); S3 = __ldg(sp + 3); }
    const uint32_t flags = __float_as_uint(S3.w);
    const bool hasNormals = flags & 1u, hasUv = flags & 2u;
    Surf s;
    s.matRef = __float_as_uint(__ldg(reinterpret_cast<const float*>(tp + 2) + 3));
    F3 p0 = f3(0, 0, 0), p1 = p0, p2 = p0;
    if (!hasNormals || needBumpFrame) { const float4 A = __ldg(tp), B = __ldg(tp + 1), C = __ldg(tp + 2); p0 = xyz(A); p1 = xyz(B); p2 = xyz(C); }
    if (hasNormals) {
        F3 n0 = xyz(S0), n1 = xyz(S1), n2 = xyz(S2);
        s.normal = normalize3(hit.i * n0 + hit.j * n1 + hit.k * n2);
    } else {
        s.normal = normalize3(cross3(p1 - p0, p2 - p0));
    }
    float2 t0, t1, t2;
    if (hasUv) { t0 = make_float2(S0.w, S1.w); t1 = make_float2(S2.w, S3.x); t2 = make_float2(S3.y, S3.z); }
    else { t0 = make_float2(0.0f, 0.0f); t1 = make_float2(1.0f, 0.0f); t2 = make_float2(0.0f, 1.0f); }
    s.u = hit.i * t0.x + hit.j * t1.x + hit.k * t2.x;
    s.v = hit.i * t0.y + hit.j * t1.y + hit.k * t2.y;
    s.dp01 = s.dp02 = f3(0.0f, 0.0f, 0.0f);
    s.duv01x = s.duv01y = s.duv02x = s.duv02y = 0.0f;
    if (needBumpFrame) {
        s.dp01 = normalize3(p1 - p0);
        s.dp02 = normalize3(p2 - p0);
        safeNormalize2(t1.x - t0.x, t1.y - t0.y, s.duv01x, s.duv01y);
        safeNormalize2(t2.x - t0.x, t2.y - t0.y, s.duv02x, s.duv02y);
    }
    return s;
}

// Texture::sample<float> / sample<Vector3f> (texture.cpp:105-139): taps accumulated in index order, u8 scaled by 1/255 at the end
__device__ __forceinline__ float sampleTex1(const DTexture& t, float u, float v) {
    float k[4]; int ind[4];
    bilinear(k, ind, t.comp, u, v, t.width, t.height);
    float c = 0.0f;
    if (t.channel == 0) {
        const uint8_t* p = (const uint8_t*)t.data;
#pragma unroll
        for (int i = 0; i < 4; i++) c = c + k[i] * (float)__ldg(p + ind[i]);
        return (1.0f / 255.0f) * c;
    }
    const float* p = (const float*)t.data;
#pragma unroll
    for (int i = 0; i < 4; i++) c = c + k[i] * __ldg(p + ind[i]);
    return c;
}
__device__ __forceinline__ F3 sampleTex3(const void* data, int channel, int comp, int w, int h, float u, float v) {
    float k[4]; int ind[4];
    bilinear(k, ind, comp, u, v, w, h);
    F3 c = f3(0.0f, 0.0f, 0.0f);
    if (channel == 0) {
        const uint8_t* p = (const uint8_t*)data;
#pragma unroll
        for (int i = 0; i < 4; i++) c = c + k[i] * f3((float)__ldg(p + ind[i]), (float)__ldg(p + ind[i] + 1), (float)__ldg(p + ind[i] + 2));
        return (1.0f / 255.0f) * c;
    }
    const float* p = (const float*)data;
#pragma unroll
    for (int i = 0; i < 4; i++) c = c + k[i] * f3(__ldg(p + ind[i]), __ldg(p + ind[i] + 1), __ldg(p + ind[i] + 2));
    return c;
}

__device__ __forceinline__ F3 sampleDiffuse(const DScene& sc, const DMaterial& m, float u, float v) {   // material.cpp:87-96
    F3 color = f3(m.diffuse[0], m.diffuse[1], m.diffuse[2]);
    if (m.diffuseTex >= 0) {
        const DTexture& t = sc.textures[m.diffuseTex];
        F3 c = sampleTex3(t.data, t.channel, t.comp, t.width, t.height, u, v);
        color = color * f3(powRef(c.x, 2.2f), powRef(c.y, 2.2f), powRef(c.z, 2.2f));                    // degamma, material.cpp:24-28
    }
    return color;
}

__device__ __forceinline__ F3 sampleBump(const DScene& sc, const DMaterial& m, const Surf& p) {          // material.cpp:98-114
    F3 normal = p.normal;
    if (m.bumpTex >= 0) {
        const DTexture& t = sc.textures[m.bumpTex];
        float one = 0.5f / (float)t.width + 0.5f / (float)t.height;                                      // texture.h:28
        float b = sampleTex1(t, p.u, p.v);
        float b01 = sampleTex1(t, p.u + one * p.duv01x, p.v + one * p.duv01y) - b;
        float b02 = sampleTex1(t, p.u + one * p.duv02x, p.v + one * p.duv02y) - b;
        const float nk = 4.0f;
        normal = normalize3(normal + nk * b01 * p.dp01 + nk * b02 * p.dp02);
    }
    return normal;
}

// InfiniteAreaLight::sample (light.cpp:84-128).  The reference scans the CDFs linearly for the first entry
// (from index 1) that exceeds u and has a non-zero step; the tables are non-decreasing, so that entry is the
// upper bound (binary search) followed by a skip over an initial zero-step plateau — same index, O(log n).
__device__ __forceinline__ int cdfFind(const float* __restrict__ cdf, const uint16_t* __restrict__ guide, int n, float u, float& pdf, float& xf) {
    // first index in [1,n) with cdf[i] > u, or n.  u*256 is exact, so u lies in bucket g = floor(u*256) and the answer in
    // [guide[g], guide[g+1]] (host-built bounds); u >= 1 cannot occur (rngFloat < 1) but is clamped anyway.
    const int g = min((int)(u * 256.0f), 255);
    int lo = __ldg(guide + g), hi = __ldg(guide + g + 1);
    while (lo < hi) { int mid = (lo + hi) >> 1; if (__ldg(cdf + mid) > u) hi = mid; else lo = mid + 1; }
    int i = lo;
    pdf = 1.0f; xf = 0.0f;
    for (; i < n; i++) {
        float prev = __ldg(cdf + i - 1);
        float p = __ldg(cdf + i) - prev;
        if (p == 0.0f) continue;
        pdf = p;
        xf = (float)i + (u - prev) / pdf - 1.0f;
        break;
    }
    return i;
}
__device__ __forceinline__ void sampleInfinite(const DScene& sc, float ux, float uy, F3& dir, F3& color) {
    const float kPi = 3.14159265358979323846f;
    float pdfV, yf, pdfH, xf;
    int y = cdfFind(sc.infVertP, sc.infGuideV, sc.infH, uy, pdfV, yf);
    if (y >= sc.infH) y = sc.infH - 1;         // reference reads out of bounds here (UB); configs are checked never to reach it
    cdfFind(sc.infHoriP + (size_t)y * sc.infW, sc.infGuideH + (size_t)y * 257, sc.infW, ux, pdfH, xf);
    float u = xf / (float)sc.infW, v = yf / (float)sc.infH;
    color = sampleTex3(sc.infImage, 1, 4, sc.infW, sc.infH, u, v);
    color = divs(color, pdfH * pdfV);
    color = divs(color, (float)(sc.infW * sc.infH));
    float theta = 2.0f * kPi * (u + 0.5f);
    float phi = kPi * v;
    float sinPhi = sinRef(phi);
    dir = normalize3(f3(cosRef(theta) * sinPhi, cosRef(phi), sinRef(theta) * sinPhi));
}

}  // namespace prt_b200
