// Wavefront path tracer: PathTracer::TraceBlock / TracePacket / ComputeRadiance(packet) (path_tracer.cpp:17-308)
// as a sequence of kernels over a batch of packets:
//
//   wfPrimary            generate 8 jittered camera rays per packet (camera.cpp:35-73) + packet-rule closest hit
//   for depth 0..13:     wfShade(depth)   compact the packet's surviving paths (ballot/prefix inside each 8-lane
//                                         segment), surface properties, emissive, BSDF sampling, light sampling,
//                                         Russian roulette; marks the slot's extend ray / shadow ray as pending
//                        wfCompact(stage) stream compaction of the pending rays into dense slot-id lists (warp ballot/prefix,
//                                         one atomicAdd per block)
//                        wfTrace(stage)   persistent kernel with dynamic ray fetch: closest hit for extend rays
//                                         (single-ray rules), any hit for shadow rays (packet or single rules per ray)
//   wfShade(14)          resolves the last shadow rays, sums the packet's 8 result slots
//   wfAccumulate         per pixel: sum of its packets in order, /spp, *exposure (path_tracer.cpp:28, image.cpp:44-50)
//
// The unit of state is the reference's: the 8-sample packet of one pixel (4 packets per warp).  The reference's
// slot-compaction behaviour is reproduced exactly: beta[] moves only for depth > 4, result[] never
// moves, and materials[ci] is taken from props[i] (the new hit when ci == i, path i's previous vertex otherwise).
// The packet's single xorshift32 stream is consumed in the reference's order: a lane's offset into the stream is
// an 8-lane exclusive prefix sum of per-path draw counts.
//
// Rays are never copied into queues: an extend ray is (pos[s], extDir[s], kFar), a shadow ray is rebuilt from
// (pos[s], lightDir[s]) with the reference's expressions (path_tracer.cpp:209-212 / 236-238) by the trace kernel.
// ncu on the first version (atomic queues) showed 29 % of wfShade's stall samples waiting on the queue atomics and
// another 14 % on serialised dependent loads at kernel entry (profiles/r1_v2_*); both are gone in this layout.
#include <cstdlib>
#include "prt_context.h"
#include "prt_shade.cuh"
#define PRT_STACK_STRIDE 128   // wfTrace keeps its traversal stacks in shared memory, one column per thread
#include "prt_trace.cuh"
#include "prt_wide.cuh"

using namespace prt_b200;

namespace prt_b200 {

constexpr uint32_t kDone = 0xffu;
constexpr int kWfBlock = 128;
constexpr uint32_t kStages = 16;            // stage 0 = primary, stage d+1 = depth d
constexpr uint32_t kCounterWords = 13 * kStages;  // [stage][4] queue / ray counters, [stage][4] live-class counts, [stage] skipped shadow rays,
                                                  // [stage][4] parked rays (second trace launch of multi-mesh scenes): extFetch, shadowFetch, extRays, shadowRays

struct WfDev {
    float4 *pos;        // xyz: path vertex position = origin of the slot's next rays
    float4 *betaP;      // xyz: throughput beta[slot]; w: props[slot].material (global material index bits)
    float4 *result;     // xyz: result[slot]
    float4 *pend;       // xyz: beta * lightRadiance of the pending shadow ray
    float4 *betaRR;     // xyz: beta[i]/(1-q) of the path that survived Russian roulette
    float4 *extDir;     // xyz: normalised direction of the slot's extend ray
    float4 *lightD;     // xyz: light direction of the slot's shadow ray
    prt_hit* hits;      // extend result per slot
    uint8_t *extFlag;   // 1: extend ray pending for this slot
    uint8_t *shFlag;    // 0: no shadow ray, 1: single-ray rules, 2: packet rules (path_tracer.cpp:198), 3: not traced (its radiance is exactly zero)
    uint8_t *occl;      // shadow result per slot
    uint32_t *rng, *alive;
    float4* packetResult;
    uint32_t *extList, *shList;   // compacted slot ids of pending extend / shadow rays (shadow: bit 31 = packet rules)
    // Multi-mesh scenes (prt_trace.cuh, PHASE): rays that leave mesh 0 for a later mesh are parked by the first trace launch
    // (extend: partial hit in hits[], extPark[slot] = 1; shadow: occl[slot] = 2), compacted into these lists by wfCompactParked
    // and finished by the second launch.  Null for single-mesh scenes.
    uint32_t *extList2, *shList2;
    uint8_t *extPark;
    // Round 2 tried re-ordering both lists per stage by (Morton code of the origin's cell in a 32^3 grid over the scene, direction
    // octant) with a counting sort (count / scan / scatter kernels), so that the rays a warp holds walk the same part of the tree:
    // on c5 the trace launches got 7 % shorter (5.07 -> 4.72 ms average) and the three sort kernels cost 15 % of them; end to end
    // c5 446 -> 409, c4 625 -> 536, c3 1251 -> 1128 Msamples/s (profiles/r2/b1_ab.log, s0 / s1 launch shares).  Removed: diffuse
    // bounce rays leave neighbouring pixels in every direction of the hemisphere, and a 7 % gain does not pay for any sort.
    uint32_t* counters; // [stage][4]: extFetch, shadowFetch, extRays, shadowRays
    // Packets still alive before the shade of depth = stage (stage >= 1), in four classes by the number of paths they
    // still carry: class c holds packets with alive <= 8 >> c (8, 4, 2, 1) and is shaded by groups of 8 >> c lanes, so a
    // warp shades 4 / 8 / 16 / 32 packets at once instead of idling the lanes of paths that ended (wfShade ran at
    // 15.9 / 32 lanes with fixed 8-lane groups, profiles/r1/r1_v15_*).
    uint32_t* liveCnt;         // [stage][4]
    uint32_t* liveList[2][4];  // packet ids per class, ping-pong by stage parity (built by wfCompact)
};

// Thread index space over the four class lists: class c occupies n[c] * (8 >> c) threads, padded to a whole warp so that
// the group width is warp-uniform.
struct LiveMap {
    uint32_t n[4], b[5];
    __device__ __forceinline__ LiveMap(const uint32_t* cnt4) {
        b[0] = 0;
#pragma unroll
        for (uint32_t c = 0; c < 4; c++) { n[c] = cnt4[c]; b[c + 1] = b[c] + ((n[c] * (8u >> c) + 31u) & ~31u); }
    }
    // thread -> (packet, lane, group width); false: padding
    __device__ __forceinline__ bool locate(const WfDev& w, uint32_t parity, uint32_t tid, uint32_t& pk, uint32_t& l, uint32_t& G) const {
        const uint32_t c = (tid >= b[1] ? 1u : 0u) + (tid >= b[2] ? 1u : 0u) + (tid >= b[3] ? 1u : 0u);
        const uint32_t base = c == 0u ? 0u : (c == 1u ? b[1] : (c == 2u ? b[2] : b[3]));
        const uint32_t cnt = c == 0u ? n[0] : (c == 1u ? n[1] : (c == 2u ? n[2] : n[3]));
        G = 8u >> c;
        const uint32_t idx = tid - base, pi = idx >> (3u - c);
        l = idx & (G - 1u);
        if (pi >= cnt) return false;
        const uint32_t* list = c == 0u ? w.liveList[parity][0] : (c == 1u ? w.liveList[parity][1] : (c == 2u ? w.liveList[parity][2] : w.liveList[parity][3]));
        pk = list[pi];
        return true;
    }
};

// Pixel enumeration of one render call over the region [y0,yEnd) x [x0,x0+bw).  Region mode (world == 0): the region's pixels
// row-major, written into the full-image RGB32F layout.  Band mode (multi-GPU): the region's rows are cut into bands of
// `bandRows` rows dealt round-robin to `world` ranks; this rank enumerates its bands in order and writes them tile-major into
// a rank-local buffer (row stride bw) whose layout IS the all-gather send buffer (no repack before the collective).
struct PixMap {
    uint32_t x0, y0, bw;          // region origin and width
    uint32_t bandRows, rank, world, yEnd;   // world == 0: region mode
    uint32_t outStride, outX0;    // output row stride in pixels and the image column of output column 0
    __host__ __device__ bool locate(uint32_t pix, uint32_t& px, uint32_t& py, uint32_t& outRow) const {
        uint32_t lr = pix / bw;
        px = x0 + pix % bw;
        if (world == 0) { py = y0 + lr; outRow = py; return true; }
        uint32_t j = lr / bandRows, rib = lr - j * bandRows;
        py = y0 + (j * world + rank) * bandRows + rib;
        outRow = lr;
        return py < yEnd;
    }
    __host__ __device__ size_t outIndex(uint32_t px, uint32_t outRow) const { return (size_t)outRow * outStride + (px - outX0); }
};

struct WavefrontBuffers {
    WfDev d[4] = {};              // lanes: consecutive batches rotate over them, one stream each
    uint32_t lanes = 0;
    uint32_t capacityPackets = 0;
    std::vector<void*> allocs;
};

int foldWavefrontCounters(prt_ctx* ctx) {
    if (ctx->wfCountersPending == 0) return PRT_OK;
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t b = 0; b < ctx->wfCountersPending; b++) {
        const uint32_t* c = ctx->h_wfCounters + b * kCounterWords;
        for (uint32_t st = 1; st < kStages; st++) {
            ctx->counters.extendRays += c[4 * st + 2]; ctx->counters.shadowRays += c[4 * st + 3]; ctx->counters.shadowRaysSkipped += c[8 * kStages + st];
        }
    }
    ctx->wfCountersPending = 0;
    return PRT_OK;
}

void destroyWavefront(prt_ctx* ctx) {
    if (ctx->h_wfCounters) { cudaFreeHost(ctx->h_wfCounters); ctx->h_wfCounters = nullptr; ctx->wfCountersCap = 0; ctx->wfCountersPending = 0; }
    if (!ctx->wf) return;
    for (auto p : ctx->wf->allocs) cudaFree(p);
    delete ctx->wf;
    ctx->wf = nullptr;
}

}  // namespace prt_b200

namespace {

__device__ __forceinline__ float4 pack(F3 v, float w) { return make_float4(v.x, v.y, v.z, w); }

__device__ __forceinline__ uint32_t packetSeedDev(uint32_t seed, uint32_t pixel, uint32_t ppp, uint32_t packet) {
    uint32_t h = seed * 0x9E3779B9u + pixel * ppp + packet;
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h | 1u;
}

// position of the n-th (0-based) set bit of an 8-bit mask (n < popc(mask)); three popc steps instead of __fns' loop
__device__ __forceinline__ uint32_t nthSetBit8(uint32_t m, uint32_t n) {
    uint32_t pos = 0, c = __popc(m & 0xfu);
    if (n >= c) { n -= c; pos = 4; m >>= 4; }
    c = __popc(m & 0x3u);
    if (n >= c) { n -= c; pos += 2; m >>= 2; }
    if (n >= (m & 1u)) pos += 1;
    return pos;
}

// ---- stage 0: primary rays (one thread per sample; 8 consecutive lanes = one packet) ----
// One ray per thread from start to end, no queue: camera rays of neighbouring pixels walk the same nodes, and the static assignment
// keeps them together.  Round 2 measured the alternative — a generation kernel followed by a persistent dynamic-fetch trace kernel
// with shared-memory stacks, as the batch kernels do — and it lost on every config: C1 595 -> 576, C3 1253 -> 1237, C4 619 -> 618,
// C5 450 -> 340 Msamples/s (profiles/r2/b6_split_primary_ab.log); so did the wide-node variant of that kernel in the fast mode (b5).
__global__ void __launch_bounds__(kWfBlock) wfPrimary(DScene sc, WfDev w, PixMap pm, uint32_t pix0, uint32_t ppp,
                                                      uint32_t seed, uint32_t nPackets) {
    uint32_t stack[kStackSize];
    float stackT[kStackSize];
    uint32_t s = blockIdx.x * kWfBlock + threadIdx.x;
    uint32_t pk = s >> 3, l = s & 7u;
    if (pk >= nPackets) return;                                   // whole 8-lane groups leave together
    const unsigned lane = threadIdx.x & 31u, gbase = lane & ~7u, gmask = 0xffu << gbase;
    uint32_t pix = pix0 + pk / ppp, p = pk % ppp;
    uint32_t px, py, outRow;
    if (!pm.locate(pix, px, py, outRow)) {                        // padding rows of the last band: empty packet
        if (l == 0) { w.alive[pk] = kDone; w.packetResult[pk] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); }
        w.extFlag[s] = 0; w.shFlag[s] = 0;
        return;
    }
    uint32_t st = packetSeedDev(seed, py * sc.cam.width + px, ppp, p);
    st = xsAdvance(st, 2 * l);
    st = xs32(st); float u0 = rngFloat(st);
    st = xs32(st); float u1 = rngFloat(st);
    if (l == 7) w.rng[pk] = st;                                   // stream position after the 16 jitter draws
    const DCamera& c = sc.cam;
    const float kScreenScale = 0.6f;
    const float kAspect = (float)c.width / (float)c.height;
    float jx = (2.0f * u0 - 1.0f) * (0.5f * c.invW);
    float jy = (2.0f * u1 - 1.0f) * (0.5f * c.invH);
    float nx = 2.0f * ((float)px * c.invW - 0.5f + jx) * kScreenScale * kAspect;
    float ny = -2.0f * ((float)py * c.invH - 0.5f + jy) * kScreenScale;
    F3 d = normalize3(nx * f3(c.right[0], c.right[1], c.right[2]) + ny * f3(c.up[0], c.up[1], c.up[2]) + f3(c.dir[0], c.dir[1], c.dir[2]));
    float ax = 0.0f, ay = 0.0f, az = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        ax = ax + __shfl_sync(gmask, d.x, gbase + k);
        ay = ay + __shfl_sync(gmask, d.y, gbase + k);
        az = az + __shfl_sync(gmask, d.z, gbase + k);
    }
    ax = ax / 8.0f; ay = ay / 8.0f; az = az / 8.0f;
    uint32_t rev = (ax < 0.0f ? 1u : 0u) | (ay < 0.0f ? 2u : 0u) | (az < 0.0f ? 4u : 0u);
    RayD r = makeRay(c.pos[0], c.pos[1], c.pos[2], d.x, d.y, d.z, 100000.0f, true);
    Hit h = sceneIntersectPacket(sc, r, rev, stack, stackT);
    float2* o = reinterpret_cast<float2*>(w.hits + s);
    o[0] = make_float2(h.t, h.i); o[1] = make_float2(h.j, h.k);
    o[2] = make_float2(__uint_as_float(h.tri), __uint_as_float(h.meshId));
    w.extDir[s] = pack(d, 0.0f);
    w.pos[s] = make_float4(c.pos[0], c.pos[1], c.pos[2], 0.0f);
    w.result[s] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    w.extFlag[s] = 1;                                             // "traced"
    w.shFlag[s] = 0;
    if (l == 0) w.alive[pk] = 8u;
}

// ---- stage d+1: compaction of the previous stage's hits + shading at depth d ----
// depth 0 runs over every packet; depth >= 1 over the compacted list of packets that are still alive (wfCompact), so
// warps stay fully populated as packets finish (v5 ran this kernel at 16.4 / 32 lanes, profiles/r1/r1_v5_*).
// G = width of the lane group that shades this packet (8, 4, 2 or 1; warp-uniform, >= the packet's alive count): lanes l < G
// of the packet take part, slots l >= G hold finished paths whose flags are already clear.
// LIGHT / TEX / SPEC are scene-wide facts fixed when the scene is committed (light type; any material with a texture; any
// material that is not diffuse): the kernel is instantiated per combination, so a scene without textures carries no bilinear /
// powf / bump code, one without an infinite-area light no CDF search, one with diffuse materials only no reflection branch
// — fewer registers and a smaller instruction footprint (round 1: 72 registers, 10-16 % of stall samples on instruction fetch).
template <int LIGHT, bool TEX, bool SPEC>
__device__ __forceinline__ void shadeLane(const DScene& sc, const WfDev& w, uint32_t depth, uint32_t pk, uint32_t l, uint32_t G) {
    const uint32_t s = pk * 8u + l;
    const uint32_t aliveOld = w.alive[pk];
    if (aliveOld == kDone) return;
    // every load of the previous stage's outputs is issued before anything depends on it
    const float4 resIn = w.result[s];
    const float4 pe = w.pend[s];
    const uint8_t shf = w.shFlag[s], occ = w.occl[s], exf = w.extFlag[s];
    const float hitT = w.hits[s].t;
    const unsigned lane = threadIdx.x & 31u, gbase = lane & ~(G - 1u), gbits = (1u << G) - 1u, gmask = gbits << gbase;
    const uint32_t s0 = pk * 8u;
    const bool isPrimary = depth == 0;

    // -- A. consume the previous stage: shadow resolve (path_tracer.cpp:225-232 / 246-249), hit mask --
    bool valid = false;
    float4 res = resIn;
    if (l < aliveOld) {
        if (!isPrimary && (shf == 1 || shf == 2) && occ == 0) { res.x = res.x + pe.x; res.y = res.y + pe.y; res.z = res.z + pe.z; }
        valid = exf != 0 && hitT != -1.0f;
    }
    const uint32_t vmask = (__ballot_sync(gmask, valid) >> gbase) & gbits;
    const uint32_t newAlive = __popc(vmask);
    if (newAlive == 0 || depth >= 14) {                           // ci == 0 -> break (:295), or the loop bound (:124)
        F3 sum = f3(0.0f, 0.0f, 0.0f);
#pragma unroll
        for (uint32_t k = 0; k < 8; k++) {                       // res = res + result[i], i = 0..7 (:303-307)
            if (k < G) {
                sum.x = sum.x + __shfl_sync(gmask, res.x, gbase + k);
                sum.y = sum.y + __shfl_sync(gmask, res.y, gbase + k);
                sum.z = sum.z + __shfl_sync(gmask, res.z, gbase + k);
            } else {                                              // a path that ended earlier: its slot holds the final value
                const float4 rk = w.result[s0 + k];
                sum.x = sum.x + rk.x; sum.y = sum.y + rk.y; sum.z = sum.z + rk.z;
            }
        }
        if (l == 0) { w.packetResult[pk] = pack(sum, 0.0f); w.alive[pk] = kDone; }
        w.extFlag[s] = 0; w.shFlag[s] = 0;                        // nothing of this packet is pending any more
        return;
    }

    // -- compaction: new slot c <- old path src = (c+1)-th set bit of vmask (ci in path_tracer.cpp:254-293) --
    const bool live = l < newAlive;
    const uint32_t src = live ? nthSetBit8(vmask, l) : l;
    const float2* hp = reinterpret_cast<const float2*>(w.hits + s0 + src);
    const float2 h0 = hp[0], h1 = hp[1], h2 = hp[2];
    const float4 ed = w.extDir[s0 + src];
    const float4 po = w.pos[s0 + src];
    float4 obp = w.betaP[s0 + src];                               // .w = props[src].material before this stage
    // beta[ci]: the roulette-scaled copy of path src once depth - 1 > 4 (:256-262), else whatever the slot itself held
    const bool rrBeta = !isPrimary && depth - 1 > 4;
    float4 brr = make_float4(0.0f, 0.0f, 0.0f, 0.0f), ownBeta = brr;
    if (rrBeta) brr = w.betaRR[s0 + src];
    else if (!isPrimary) ownBeta = w.betaP[s];
    __syncwarp(gmask);                                            // all reads of old slots precede the writes below
    F3 pos = f3(0.0f, 0.0f, 0.0f), rayDir = pos, normal = pos, beta = pos;
    uint32_t matRef = 0;
    float su = 0.0f, sv = 0.0f;
    if (live) {
        Hit h; h.t = h0.x; h.i = h0.y; h.j = h1.x; h.k = h1.y; h.tri = __float_as_uint(h2.x); h.meshId = __float_as_uint(h2.y);
        uint32_t newMat = surfaceMaterial(sc, h);
        matRef = (isPrimary || src == l) ? newMat : __float_as_uint(obp.w);      // props[i].material (:287)
        const DMaterial& mt = sc.materials[matRef];
        Surf sp = surfaceProps(sc, h, TEX && mt.bumpTex >= 0);    // props[ci] (:281); bump frame only if it will be read
        rayDir = xyz(ed);
        pos = h.t * rayDir + xyz(po);                             // hit.t*ray.dir + ray.org (:283 / :101)
        normal = TEX ? sampleBump(sc, mt, sp) : sp.normal;        // materials[ci]->sampleBump(props[ci]) (:288)
        if (isPrimary) beta = f3(1.0f, 1.0f, 1.0f);
        else if (rrBeta) beta = xyz(brr);                         // beta[ci] = beta[i]/(1-q) only when depth > 4 (:256-262)
        else beta = xyz(ownBeta);                                 // slot keeps whatever it held
        su = sp.u; sv = sp.v;
        obp.w = __uint_as_float(newMat);                          // props[ci].material
    }

    // -- B. shade at `depth` (path_tracer.cpp:132-190) --
    uint32_t draws = 0, refl = 2;
    DMaterial mat;
    if (live) {
        mat = sc.materials[matRef];
        refl = SPEC ? mat.refl : 0u;
        if (refl <= 1u) draws = 2u + ((refl == 0u && LIGHT == 2) ? 2u : 0u);
    }
    uint32_t off = 0, total = 0;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) {
        if (k < G) {
            uint32_t dk = __shfl_sync(gmask, draws, gbase + k);
            if (k < l) off += dk;
            total += dk;
        }
    }
    const uint32_t base = w.rng[pk];
    const float kPi = 3.14159265358979323846f;
    F3 nextDir = f3(0.0f, 0.0f, 0.0f), lightDir = f3(0.0f, 0.0f, 0.0f), pending = f3(0.0f, 0.0f, 0.0f);
    bool shadowable = false, survive = false;
    F3 betaOut = f3(0.0f, 0.0f, 0.0f);
    uint32_t st = base;
    if (live) {
        if (mat.emissive[0] != 0.0f) {                            // :137-139 (does not terminate the path)
            F3 e = beta * f3(mat.emissive[0], mat.emissive[1], mat.emissive[2]);
            res.x = res.x + e.x; res.y = res.y + e.y; res.z = res.z + e.z;
        }
        st = xsAdvance(base, off);
        if (refl <= 1u) {
            st = xs32(st); float r2 = rngFloat(st);
            float r2sq = sqrtf(r2);
            st = xs32(st); float r1 = rngFloat(st);
            F3 u = (fabsf(normal.x) > 0.1f) ? f3(0.0f, 1.0f, 0.0f) : f3(1.0f, 0.0f, 0.0f);
            F3 tangent = normalize3(cross3(normal, u));
            F3 binormal = normalize3(cross3(tangent, normal));
            float theta = 2.0f * kPi * r1;
            F3 diffuseDir = (r2sq * cosRef(theta)) * binormal + (r2sq * sinRef(theta)) * tangent + (1.0f - r2) * normal;   // :153
            if (refl == 0u) {
                nextDir = diffuseDir;
                beta = beta * (TEX ? sampleDiffuse(sc, mat, su, sv) : f3(mat.diffuse[0], mat.diffuse[1], mat.diffuse[2]));      // :162
                F3 lightIntensity = f3(0.0f, 0.0f, 0.0f);
                if (LIGHT == 2) {
                    st = xs32(st); float ux = rngFloat(st);
                    st = xs32(st); float uy = rngFloat(st);
                    sampleInfinite(sc, ux, uy, lightDir, lightIntensity);
                    shadowable = true;
                } else if (LIGHT == 1) {
                    lightDir = f3(sc.lightDir[0], sc.lightDir[1], sc.lightDir[2]);
                    lightIntensity = f3(sc.lightIntensity[0], sc.lightIntensity[1], sc.lightIntensity[2]);
                    shadowable = true;
                }
                if (shadowable) {
                    float cs = dot3(lightDir, normal);
                    cs = (cs < 0.0f) ? 0.0f : cs;                                                                      // std::max(dot, 0.0f)
                    F3 lightRadiance = divs(lightIntensity * cs, kPi);                                                  // :229 / :247
                    pending = beta * lightRadiance;
                }
            } else {
                float dn = dot3(normal, rayDir);
                F3 reflectDir = rayDir - (normal * 2.0f) * dn;                                                          // :186
                nextDir = 0.9f * reflectDir + 0.1f * diffuseDir;                                                        // :188
            }
        }
        // Russian roulette (:256-262): one draw per alive path, after all of this depth's shading draws
        survive = true;
        betaOut = beta;
        if (depth > 4) {
            st = xsAdvance(st, total - off - draws + l);
            st = xs32(st); float rr = rngFloat(st);
            float lb = 1.0f - length3(beta);
            float q = (0.05f < lb) ? lb : 0.05f;                                                                        // std::max(0.05f, x)
            if (rr < q) survive = false;
            else betaOut = divs(beta, 1.0f - q);
        }
        if (refl > 1u) survive = false;          // kRefraction: nextRayDir stays zero (P3) -> NaN ray -> miss
        if (depth >= 13) survive = false;        // the last depth's extend result is never used (:124, :295-300)
    }
    // The packet's stream position after this depth = the position of whichever lane consumed the last draw:
    // depth > 4: the last alive path's roulette draw; else the last path that sampled.  No recomputation.
    {
        uint32_t lastMask = (__ballot_sync(gmask, live && (depth > 4 || draws != 0u)) >> gbase) & gbits;
        uint32_t fin = __shfl_sync(gmask, st, gbase + (lastMask ? 31 - __clz(lastMask) : 0));
        if (l == 0) w.rng[pk] = lastMask ? fin : base;
    }

    // -- publish the slot's rays: the trace kernel compacts them on the fly --
    const bool grouping = (newAlive & 0xfu) > 2u;                                                                       // :198
    if (live) {
        F3 edir = f3(0.0f, 0.0f, 0.0f);
        if (survive) edir = normalize3(nextDir);                                                                        // :265
        w.pos[s] = pack(pos, 0.0f);
        if (res.x != resIn.x || res.y != resIn.y || res.z != resIn.z) w.result[s] = res;      // only emission / a resolved shadow ray change it
        w.betaP[s] = pack(beta, obp.w);
        if (depth > 4) w.betaRR[s] = pack(betaOut, 0.0f);         // read by the next depth only once the roulette runs
        // A shadow ray whose radiance is exactly +0 cannot change result[] whatever it hits (x + 0 == x; a NaN or a negative
        // zero does not compare equal here / cannot arise from non-negative factors), and tracing consumes no random
        // numbers: it is counted (shFlag 3) but not traced.  Light below the surface's horizon (cs == 0) is the common case.
        const bool zeroShadow = shadowable && pending.x == 0.0f && pending.y == 0.0f && pending.z == 0.0f
                                && !signbit(pending.x) && !signbit(pending.y) && !signbit(pending.z);
        if (shadowable && !zeroShadow) w.pend[s] = pack(pending, 0.0f);   // read only under shFlag 1 / 2
        w.extDir[s] = pack(edir, 0.0f);
        if (shadowable && !zeroShadow) w.lightD[s] = pack(lightDir, 0.0f);
        w.extFlag[s] = survive ? 1 : 0;
        w.shFlag[s] = shadowable ? (zeroShadow ? 3 : (grouping ? 2 : 1)) : 0;
    } else {
        if (l < aliveOld && (res.x != resIn.x || res.y != resIn.y || res.z != resIn.z)) w.result[s] = res;              // shadow resolve of a path that just died
        w.extFlag[s] = 0;
        w.shFlag[s] = 0;
    }
    if (l == 0) w.alive[pk] = newAlive;
}

// Grid-stride over the packets to shade: a fixed, machine-sized grid instead of one block per 16 packets of the batch — at
// depth >= 6 only a few per cent of the packets are alive and the 65 536 mostly-empty blocks cost more than the work
// (49-120 us per launch in profiles/r1/r1_v5_c3_16spp_launches.csv).
// 8 blocks / SM (64 registers) for every untextured variant — the infinite-light ones spill 8 bytes for it and still gain: C3 1250 ->
// 1277 Msamples/s (fast mode 1351 -> 1380); the textured variants would spill 60 - 100 bytes for nothing (C4 640 -> 641) and keep 72.
template <int LIGHT, bool TEX, bool SPEC>
__global__ void __launch_bounds__(kWfBlock, TEX ? 7 : 8) wfShade(DScene sc, WfDev w, uint32_t depth, uint32_t nPackets) {
    if (depth == 0) {
        for (uint32_t tid = blockIdx.x * kWfBlock + threadIdx.x; tid < nPackets * 8u; tid += gridDim.x * kWfBlock) shadeLane<LIGHT, TEX, SPEC>(sc, w, 0, tid >> 3, tid & 7u, 8u);
        return;
    }
    const LiveMap map(w.liveCnt + 4 * depth);
    for (uint32_t tid = blockIdx.x * kWfBlock + threadIdx.x; tid < map.b[4]; tid += gridDim.x * kWfBlock) {   // whole lane groups leave together
        uint32_t pk, l, G;
        if (map.locate(w, depth & 1u, tid, pk, l, G)) shadeLane<LIGHT, TEX, SPEC>(sc, w, depth, pk, l, G);
    }
}

using ShadeKernel = void (*)(DScene, WfDev, uint32_t, uint32_t);
template <int LIGHT>
ShadeKernel shadeKernelFor(bool tex, bool spec) {
    return tex ? (spec ? wfShade<LIGHT, true, true> : wfShade<LIGHT, true, false>) : (spec ? wfShade<LIGHT, false, true> : wfShade<LIGHT, false, false>);
}
ShadeKernel shadeKernelFor(uint32_t light, bool tex, bool spec) {
    return light == 2u ? shadeKernelFor<2>(tex, spec) : (light == 1u ? shadeKernelFor<1>(tex, spec) : shadeKernelFor<0>(tex, spec));
}

// ---- stream compaction of the slots' pending rays into dense slot-id lists: warp ballot/prefix, one atomicAdd per
// block and list (wfShade itself stays free of atomics; 29 % of its stall samples were queue atomics before) ----
// Stage 1 scans every slot of the batch; stage >= 2 only the slots of the packets that were alive before the previous shade
// (the previous live list) — every other slot's flags are already zero.  Grid-stride, machine-sized grid.
__global__ void __launch_bounds__(256) wfCompact(WfDev w, uint32_t stage, uint32_t nSlots) {
    __shared__ uint32_t warpCnt[7][8];                            // 0: extend rays, 1: shadow rays, 2..5: live packets of class 0..3, 6: shadow rays not traced
    __shared__ uint32_t blockBase[7];
    const unsigned lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const LiveMap prev(w.liveCnt + 4 * (stage <= 1u ? 1u : stage - 1u));      // unused for stage <= 1
    const uint32_t limit = stage <= 1u ? nSlots : prev.b[4];
    for (uint32_t base = blockIdx.x * 256u; base < limit; base += gridDim.x * 256u) {      // uniform per block: barriers are safe
    const uint32_t t = base + threadIdx.x;
    uint32_t s = 0, cls = 4;                                      // cls < 4: first lane of a packet that is still alive, and its class
    uint8_t ef = 0, sf = 0;
    if (t < limit) {
        bool in = true;
        if (stage <= 1u) s = t;
        else { uint32_t pk, l, G; in = prev.locate(w, (stage - 1u) & 1u, t, pk, l, G); s = pk * 8u + l; }
        if (in) {
            ef = w.extFlag[s]; sf = w.shFlag[s];
            if ((s & 7u) == 0u) {
                const uint32_t a = w.alive[s >> 3];
                if (a != kDone) cls = a > 4u ? 0u : (a > 2u ? 1u : (a == 2u ? 2u : 3u));
            }
        }
    }
    unsigned m[7];
    m[0] = __ballot_sync(0xffffffffu, ef != 0); m[1] = __ballot_sync(0xffffffffu, sf == 1 || sf == 2); m[6] = __ballot_sync(0xffffffffu, sf == 3);
#pragma unroll
    for (uint32_t c = 0; c < 4; c++) m[2 + c] = __ballot_sync(0xffffffffu, cls == c);
    if (lane < 7) warpCnt[lane][wid] = __popc(lane == 0 ? m[0] : lane == 1 ? m[1] : lane == 2 ? m[2] : lane == 3 ? m[3] : lane == 4 ? m[4] : lane == 5 ? m[5] : m[6]);
    __syncthreads();
    uint32_t* cnt = w.counters + 4 * stage;
    if (threadIdx.x < 7) {
        uint32_t tot = 0;
        for (int k = 0; k < 8; k++) { uint32_t c = warpCnt[threadIdx.x][k]; warpCnt[threadIdx.x][k] = tot; tot += c; }
        uint32_t* dst = threadIdx.x < 2 ? cnt + 2 + threadIdx.x : (threadIdx.x < 6 ? w.liveCnt + 4 * stage + (threadIdx.x - 2) : w.counters + 8 * kStages + stage);
        blockBase[threadIdx.x] = tot ? atomicAdd(dst, tot) : 0u;
    }
    __syncthreads();
    const unsigned below = (1u << lane) - 1u;
    if (ef) w.extList[blockBase[0] + warpCnt[0][wid] + __popc(m[0] & below)] = s;
    if (sf == 1 || sf == 2) w.shList[blockBase[1] + warpCnt[1][wid] + __popc(m[1] & below)] = s | (sf == 2 ? 0x80000000u : 0u);
    if (cls < 4u) {
        const unsigned mc = cls == 0u ? m[2] : (cls == 1u ? m[3] : (cls == 2u ? m[4] : m[5]));
        uint32_t* list = cls == 0u ? w.liveList[stage & 1u][0] : (cls == 1u ? w.liveList[stage & 1u][1] : (cls == 2u ? w.liveList[stage & 1u][2] : w.liveList[stage & 1u][3]));
        list[blockBase[2 + cls] + warpCnt[2 + cls][wid] + __popc(mc & below)] = s >> 3;
    }
    __syncthreads();                                              // warpCnt / blockBase are reused by the next round
    }
}

// ---- parked rays of one stage (multi-mesh scenes) into dense lists for the second trace launch; clears the extend marks ----
__global__ void __launch_bounds__(256) wfCompactParked(WfDev w, uint32_t stage) {
    __shared__ uint32_t warpCnt[8];
    __shared__ uint32_t blockBase;
    const unsigned lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t* cnt = w.counters + 4 * stage;
    uint32_t* cnt2 = w.counters + 9 * kStages + 4 * stage;
    for (uint32_t kind = 0; kind < 2u; kind++) {
        const uint32_t n = cnt[2 + kind];
        const uint32_t* list = kind == 0u ? w.extList : w.shList;
        uint32_t* out = kind == 0u ? w.extList2 : w.shList2;
        for (uint32_t base = blockIdx.x * 256u; base < n; base += gridDim.x * 256u) {      // uniform per block: barriers are safe
            const uint32_t i = base + threadIdx.x;
            uint32_t e = 0; bool parked = false;
            if (i < n) {
                e = list[i];
                const uint32_t s = e & 0x7fffffffu;
                if (kind == 0u) { parked = w.extPark[s] != 0; if (parked) w.extPark[s] = 0; }
                else parked = w.occl[s] == 2;
            }
            const unsigned m = __ballot_sync(0xffffffffu, parked);
            if (lane == 0) warpCnt[wid] = __popc(m);
            __syncthreads();
            if (threadIdx.x == 0) {
                uint32_t tot = 0;
                for (int k = 0; k < 8; k++) { uint32_t c = warpCnt[k]; warpCnt[k] = tot; tot += c; }
                blockBase = tot ? atomicAdd(cnt2 + 2 + kind, tot) : 0u;
            }
            __syncthreads();
            if (parked) out[blockBase + warpCnt[wid] + __popc(m & ((1u << lane) - 1u))] = e;
            __syncthreads();
        }
    }
}

// ---- persistent trace kernel for one stage: closest hit for the extend list, any hit for the shadow list.  PHASE 0: every mesh
// in one launch; 1: mesh 0, rays that go on are parked; 2: the parked rays, meshes 1.. (prt_trace.cuh) ----
template <int PHASE>
__global__ void __launch_bounds__(128) wfTrace(DScene sc, WfDev w, uint32_t stage) {
    extern __shared__ uint32_t smemStack[];                       // [depth][128]: entry k of thread t at smemStack[k * 128 + t]
    uint32_t* stack = smemStack + threadIdx.x;
    float dummyT[1];
    uint32_t* cnt = w.counters + (PHASE == 2 ? 9 * kStages : 0) + 4 * stage;
    const uint32_t nExt = cnt[2], nSh = cnt[3];
    const uint32_t* extList = PHASE == 2 ? w.extList2 : w.extList;
    const uint32_t* shList = PHASE == 2 ? w.shList2 : w.shList;
    const float kFar = 2.0f * sc.radius;                                                                                // :192
    {
        auto fetch = [&](uint32_t i, RayD& r, uint32_t& tag, bool& pkt) {
            uint32_t s = __ldcs(extList + i);                     // path state is streamed (read once per stage): evict-first keeps L2 for the tree
            float4 p = __ldcs(w.pos + s), d = __ldcs(w.extDir + s);
            r = makeRay(p.x, p.y, p.z, d.x, d.y, d.z, kFar, false);                                                     // :267-271
            tag = s; pkt = false;
            return true;
        };
        auto commit = [&](uint32_t i, const Hit& h, bool, uint32_t s) {
            float2* o = reinterpret_cast<float2*>(w.hits + s);
            __stcs(o, make_float2(h.t, h.i)); __stcs(o + 1, make_float2(h.j, h.k));
            __stcs(o + 2, make_float2(__uint_as_float(h.tri), __uint_as_float(h.meshId)));
        };
        auto aux = [&](uint32_t s, Hit& h) {
            float2* o = reinterpret_cast<float2*>(w.hits + s);
            if (PHASE == 1) {                                     // park: the hit so far (t = maxT: none yet), marked for wfCompactParked
                o[0] = make_float2(h.t, h.i); o[1] = make_float2(h.j, h.k); o[2] = make_float2(__uint_as_float(h.tri), __uint_as_float(h.meshId));
                w.extPark[s] = 1;
            } else if (PHASE == 2) {
                const float2 a = o[0], b = o[1], c = o[2];
                h.t = a.x; h.i = a.y; h.j = b.x; h.k = b.y; h.tri = __float_as_uint(c.x); h.meshId = __float_as_uint(c.y);
            }
        };
        traceQueue<kClosestSingle, PHASE>(sc, nExt, cnt + 0, stack, dummyT, fetch, commit, aux);
    }
    {
        auto fetch = [&](uint32_t i, RayD& r, uint32_t& tag, bool& pkt) {
            uint32_t e = __ldcs(shList + i), s = e & 0x7fffffffu;
            float4 p = __ldcs(w.pos + s), L = __ldcs(w.lightD + s);
            pkt = (e >> 31) != 0u;
            // org = pos + kFar*lightDir ; dir = -lightDir ; maxT = kFar - 0.0008 (:209-212 / :236-238)
            r = makeRay(p.x + kFar * L.x, p.y + kFar * L.y, p.z + kFar * L.z, -L.x, -L.y, -L.z, kFar - 0.0008f, pkt);
            tag = s;
            return true;
        };
        auto commit = [&](uint32_t i, const Hit&, bool occ, uint32_t s) { __stcs(w.occl + s, (uint8_t)(occ ? 1 : 0)); };
        auto aux = [&](uint32_t s, Hit&) { if (PHASE == 1) w.occl[s] = 2; };       // park: no state beyond "not occluded by mesh 0"
        traceQueue<kOcclude, PHASE>(sc, nSh, cnt + 1, stack, dummyT, fetch, commit, aux);
    }
}

// ---- the same stage in the fast traversal mode (prt_wide.cuh): extension and shadow rays over the Wide8 trees, one launch for
// every mesh.  Same lists, same ray construction, same result slots as wfTrace. ----
__global__ void __launch_bounds__(128, 7) wfTraceWide(DScene sc, WfDev w, uint32_t stage) {
    extern __shared__ uint2 smemWide[];                           // [kWideCap][128]
    uint2 deep[kWideDeep];
    const WideStack stack{smemWide + threadIdx.x, deep};
    uint32_t* cnt = w.counters + 4 * stage;
    const uint32_t nExt = cnt[2], nSh = cnt[3];
    const float kFar = 2.0f * sc.radius;                                                                                // :192
    {
        auto fetch = [&](uint32_t i, RayD& r, uint32_t& tag, bool& pkt) {
            uint32_t s = __ldcs(w.extList + i);
            float4 p = __ldcs(w.pos + s), d = __ldcs(w.extDir + s);
            r = makeRay(p.x, p.y, p.z, d.x, d.y, d.z, kFar, false);                                                     // :267-271
            tag = s; pkt = false;
            return true;
        };
        auto commit = [&](uint32_t, const Hit& h, bool, uint32_t s) {
            float2* o = reinterpret_cast<float2*>(w.hits + s);
            __stcs(o, make_float2(h.t, h.i)); __stcs(o + 1, make_float2(h.j, h.k));
            __stcs(o + 2, make_float2(__uint_as_float(h.tri), __uint_as_float(h.meshId)));
        };
        traceQueueWide<false>(sc, nExt, cnt + 0, stack, fetch, commit);
    }
    {
        auto fetch = [&](uint32_t i, RayD& r, uint32_t& tag, bool& pkt) {
            uint32_t e = __ldcs(w.shList + i), s = e & 0x7fffffffu;
            float4 p = __ldcs(w.pos + s), L = __ldcs(w.lightD + s);
            pkt = (e >> 31) != 0u;
            r = makeRay(p.x + kFar * L.x, p.y + kFar * L.y, p.z + kFar * L.z, -L.x, -L.y, -L.z, kFar - 0.0008f, pkt);   // :209-212 / :236-238
            tag = s;
            return true;
        };
        auto commit = [&](uint32_t, const Hit&, bool occ, uint32_t s) { __stcs(w.occl + s, (uint8_t)(occ ? 1 : 0)); };
        traceQueueWide<true>(sc, nSh, cnt + 1, stack, fetch, commit);
    }
}

// ---- per pixel: colour = sum of its packets in order; colour/spp; exposure*colour (path_tracer.cpp:28, image.cpp:45) ----
__global__ void wfAccumulate(WfDev w, PixMap pm, uint32_t pix0, uint32_t nPix, uint32_t ppp, uint32_t spp,
                             float exposure, float* rgb) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPix) return;
    F3 c = f3(0.0f, 0.0f, 0.0f);
    for (uint32_t p = 0; p < ppp; p++) c = c + xyz(w.packetResult[(size_t)i * ppp + p]);
    c = divs(c, (float)spp);
    c = exposure * c;
    uint32_t px, py, outRow;
    pm.locate(pix0 + i, px, py, outRow);                          // padding rows come out as zeros
    float* o = rgb + 3 * pm.outIndex(px, outRow);
    o[0] = c.x; o[1] = c.y; o[2] = c.z;
}

template <typename T>
int wfAlloc(prt_ctx* ctx, T** p, size_t count) {
    void* d = nullptr;
    PRT_CUDA_CHECK(ctx, cudaMalloc(&d, count * sizeof(T)));
    ctx->wf->allocs.push_back(d);
    *p = (T*)d;
    return PRT_OK;
}

int ensureWavefront(prt_ctx* ctx, uint32_t packets, uint32_t lanes, bool parked) {
    if (ctx->wf && ctx->wf->capacityPackets >= packets && ctx->wf->lanes >= lanes && (!parked || ctx->wf->d[0].extPark)) return PRT_OK;
    destroyWavefront(ctx);
    ctx->wf = new WavefrontBuffers;
    size_t S = (size_t)packets * 8;
    int r = 0;
    for (uint32_t k = 0; k < lanes; k++) {
        WfDev& d = ctx->wf->d[k];
        r |= wfAlloc(ctx, &d.pos, S); r |= wfAlloc(ctx, &d.betaP, S); r |= wfAlloc(ctx, &d.result, S); r |= wfAlloc(ctx, &d.pend, S);
        r |= wfAlloc(ctx, &d.betaRR, S); r |= wfAlloc(ctx, &d.extDir, S); r |= wfAlloc(ctx, &d.lightD, S);
        r |= wfAlloc(ctx, &d.hits, S); r |= wfAlloc(ctx, &d.occl, S); r |= wfAlloc(ctx, &d.extFlag, S); r |= wfAlloc(ctx, &d.shFlag, S);
        r |= wfAlloc(ctx, &d.extList, S); r |= wfAlloc(ctx, &d.shList, S);
        if (parked) {
            r |= wfAlloc(ctx, &d.extList2, S); r |= wfAlloc(ctx, &d.shList2, S); r |= wfAlloc(ctx, &d.extPark, S);
            if (!r && cudaMemset(d.extPark, 0, S) != cudaSuccess) r = 1;
        }
        r |= wfAlloc(ctx, &d.rng, (size_t)packets); r |= wfAlloc(ctx, &d.alive, (size_t)packets); r |= wfAlloc(ctx, &d.packetResult, (size_t)packets);
        r |= wfAlloc(ctx, &d.counters, (size_t)kCounterWords);
        for (int p = 0; p < 2; p++) for (int c = 0; c < 4; c++) r |= wfAlloc(ctx, &d.liveList[p][c], (size_t)packets);
        d.liveCnt = d.counters ? d.counters + 4 * kStages : nullptr;
    }
    if (r) { destroyWavefront(ctx); return PRT_ERR_CUDA; }
    ctx->wf->capacityPackets = packets;
    ctx->wf->lanes = lanes;
    return PRT_OK;
}

}  // namespace

extern "C" {

static int renderImpl(prt_ctx* ctx, const PixMap& pm, uint32_t nPixTotal, uint32_t spp, uint32_t seed, float exposure, float* d_rgb) {
    int r;
    const uint32_t ppp = spp / 8u;                                // path_tracer.cpp:65: samples < 8 renders black
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    if (ppp == 0) {
        uint32_t rows = nPixTotal / pm.bw;
        for (uint32_t lr = 0; lr < rows; lr++) {
            uint32_t px, py, outRow;
            pm.locate(lr * pm.bw, px, py, outRow);
            PRT_CUDA_CHECK(ctx, cudaMemsetAsync(d_rgb + 3 * pm.outIndex(px, outRow), 0, 12 * (size_t)pm.bw, ctx->stream));
        }
        PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
        return PRT_OK;
    }
    // Packets per wavefront batch.  Every stage launch ends with a tail (the last long rays of its queue), so bigger batches
    // amortise 29 tails over more work: measured on B200 1 Mi -> 8 Mi packets: C3 818 -> 917, C5 311 -> 362 Msamples/s
    // (2 Mi: 872 / 339, 4 Mi: 902 / 354).  State is ~1.9 KB per packet: 8 Mi packets = 15 GB of the 180 GB; capped at a
    // quarter of the free memory, and the frame is cut into EQUAL batches (no small straggler batch).
    uint64_t kBatchPackets = 8u << 20;
    if (const char* e = getenv("PRT_BATCH_PACKETS")) { long v = atol(e); if (v >= 1024) kBatchPackets = (uint64_t)v; }   // profiling override
    // Two lanes: consecutive batches run on two streams with their own path state, so one batch's kernels fill the SMs the other
    // batch's kernel leaves idle in its tail (a persistent wfTrace ends with a few long rays) and the small late-depth launches
    // of one batch hide behind the other's work.  PRT_WF_LANES=1 restores the single stream.
    static const uint32_t laneCfg = getenv("PRT_WF_LANES") ? (uint32_t)atoi(getenv("PRT_WF_LANES")) : 2u;
    const uint64_t totalPackets = (uint64_t)nPixTotal * ppp;
    // frames of at least 256 Ki packets are cut into two batches for the two lanes (C1, 512 Ki packets: 562 -> 577 Msamples/s; three / four
    // lanes 558 / 530; profiles/r1/z16_lane_min.log, z17_c1_lanes.log); PRT_WF_LANE_MIN overrides the threshold
    static const uint64_t laneMin = getenv("PRT_WF_LANE_MIN") ? (uint64_t)atoll(getenv("PRT_WF_LANE_MIN")) : (1u << 18);
    const uint32_t lanes = (laneCfg >= 2u && totalPackets >= laneMin) ? (laneCfg > 4u ? 4u : laneCfg) : 1u;
    if (!ctx->wf || ctx->wf->capacityPackets < kBatchPackets || ctx->wf->lanes < lanes) {
        size_t freeB = 0, totalB = 0;
        if (cudaMemGetInfo(&freeB, &totalB) == cudaSuccess) {
            if (ctx->wf) freeB += (size_t)ctx->wf->capacityPackets * 2048u * ctx->wf->lanes;      // what a reallocation would give back
            const uint64_t cap = (uint64_t)(freeB / 4) / 2048u / lanes;
            if (cap < kBatchPackets) kBatchPackets = cap < 65536u ? 65536u : cap;
        }
    }
    uint64_t wantBatches = (totalPackets + kBatchPackets - 1) / kBatchPackets;
    wantBatches = (wantBatches + lanes - 1u) / lanes * lanes;                                     // a whole number of rounds of equal batches
    uint32_t pixPerBatch = (uint32_t)((nPixTotal + wantBatches - 1) / wantBatches);
    if (pixPerBatch == 0) pixPerBatch = 1;
    if (pixPerBatch > nPixTotal) pixPerBatch = nPixTotal;
    // Scenes with several meshes: two trace launches per stage (mesh 0, then the rays that go on to later meshes; prt_trace.cuh).
    // Measured on B200 (profiles/r1/z7_park_ab.log): C3 1150 -> 1213 Msamples/s at 256 spp, 1109 -> 1151 at 16 spp; C1 (one batch
    // of 0.5 Mi packets) 560 -> 546 — the 28 extra launches per batch cost about 7 us each, so small batches keep the single
    // launch.  PRT_WF_PARK=0 / 1 forces it off / on.
    const int parkCfg = getenv("PRT_WF_PARK") ? atoi(getenv("PRT_WF_PARK")) : -1;          // read per call: tests toggle it
    const bool fast = ctx->traversal == PRT_TRAVERSAL_FAST;       // opt-in: Wide8 trees for the extension and shadow rays (prt_wide.cuh)
    const bool parked = !fast && ctx->scene.meshCount > 1u && (parkCfg < 0 ? (uint64_t)pixPerBatch * ppp >= (1u << 20) : parkCfg != 0);
    if ((r = ensureWavefront(ctx, pixPerBatch * ppp, lanes, parked))) return r;
    {
        size_t nBatches = (nPixTotal + pixPerBatch - 1) / pixPerBatch;
        if (ctx->wfCountersPending + nBatches > ctx->wfCountersCap) {
            if ((r = foldWavefrontCounters(ctx))) return r;
            if (nBatches > ctx->wfCountersCap) {
                if (ctx->h_wfCounters) cudaFreeHost(ctx->h_wfCounters);
                ctx->h_wfCounters = nullptr;
                size_t cap = std::max<size_t>(nBatches, 256);
                PRT_CUDA_CHECK(ctx, cudaMallocHost((void**)&ctx->h_wfCounters, cap * kCounterWords * sizeof(uint32_t)));
                ctx->wfCountersCap = cap;
            }
        }
    }
    uint32_t depthCap = 1;
    for (auto m : ctx->meshes) depthCap = std::max(depthCap, m->build->bvh.maxDepth + 1u);
    const size_t stackBytes = (size_t)depthCap * 128u * sizeof(uint32_t);
    int traceGrid[3] = {0, 0, 0};
    {
        const void* k[3] = {(const void*)wfTrace<0>, (const void*)wfTrace<1>, (const void*)wfTrace<2>};
        for (int p = 0; p < 3; p++) {
            int perSm = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k[p], 128, stackBytes) != cudaSuccess || perSm <= 0) perSm = 4;
            traceGrid[p] = ctx->smCount * perSm;
        }
    }
    int wideGrid = 0;
    const size_t wideBytes = (size_t)kWideCap * 128u * sizeof(uint2);
    if (fast) {
        uint32_t wd = 0;
        for (auto m : ctx->meshes) wd = std::max(wd, m->fastDepth);
        if (7u * wd + 1u > (uint32_t)(kWideCap + kWideDeep)) return setError(ctx, PRT_ERR_STATE, "render: fast-mode tree of %u levels exceeds the traversal stack", wd);
        int perSm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, (const void*)wfTraceWide, 128, wideBytes) != cudaSuccess || perSm <= 0) perSm = 4;
        wideGrid = ctx->smCount * perSm;
    }
    cudaStream_t laneStream[4] = {ctx->stream, ctx->stream, ctx->stream, ctx->stream};
    if (lanes > 1u) {
        if (!ctx->evLane) PRT_CUDA_CHECK(ctx, cudaEventCreateWithFlags(&ctx->evLane, cudaEventDisableTiming));
        PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->evLane, ctx->stream));          // the other lanes start after whatever the context stream holds
        for (uint32_t k = 1; k < lanes; k++) {
            if (!ctx->laneStreams[k - 1]) PRT_CUDA_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->laneStreams[k - 1], cudaStreamNonBlocking));
            laneStream[k] = ctx->laneStreams[k - 1];
            PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(laneStream[k], ctx->evLane, 0));
        }
    }
    const uint32_t shadeGrid = (uint32_t)ctx->smCount * 32u, compactGrid = (uint32_t)ctx->smCount * 16u;   // a few waves of resident blocks
    bool anyTex = false, anyNonDiffuse = false;
    for (const DMaterial& m : ctx->materials) { anyTex |= m.diffuseTex >= 0 || m.bumpTex >= 0; anyNonDiffuse |= m.refl != 0u; }
    static const bool shadeGeneric = getenv("PRT_WF_SHADE_GENERIC") && atoi(getenv("PRT_WF_SHADE_GENERIC")) != 0;   // A/B: the one-size kernel of round 1
    const ShadeKernel shade = shadeGeneric ? shadeKernelFor(ctx->scene.lightType, true, true) : shadeKernelFor(ctx->scene.lightType, anyTex, anyNonDiffuse);
    uint32_t batch = 0;
    for (uint32_t pix0 = 0; pix0 < nPixTotal; pix0 += pixPerBatch, batch++) {
        WfDev& w = ctx->wf->d[batch % lanes];
        cudaStream_t st = laneStream[batch % lanes];
        uint32_t nPix = std::min(pixPerBatch, nPixTotal - pix0);
        uint32_t nPackets = nPix * ppp;
        uint32_t grid = (nPackets * 8u + kWfBlock - 1) / kWfBlock;
        PRT_CUDA_CHECK(ctx, cudaMemsetAsync(w.counters, 0, sizeof(uint32_t) * kCounterWords, st));
        wfPrimary<<<grid, kWfBlock, 0, st>>>(ctx->scene, w, pm, pix0, ppp, seed, nPackets);
        for (uint32_t depth = 0; depth <= 14; depth++) {
            shade<<<std::min(grid, shadeGrid), kWfBlock, 0, st>>>(ctx->scene, w, depth, nPackets);
            if (depth < 14) {
                wfCompact<<<std::min((nPackets * 8u + 255u) / 256u, compactGrid), 256, 0, st>>>(w, depth + 1, nPackets * 8u);
                if (fast) wfTraceWide<<<wideGrid, 128, wideBytes, st>>>(ctx->scene, w, depth + 1);
                else if (!parked) wfTrace<0><<<traceGrid[0], 128, stackBytes, st>>>(ctx->scene, w, depth + 1);
                else {
                    wfTrace<1><<<traceGrid[1], 128, stackBytes, st>>>(ctx->scene, w, depth + 1);
                    wfCompactParked<<<std::min((nPackets * 8u + 255u) / 256u, compactGrid), 256, 0, st>>>(w, depth + 1);
                    wfTrace<2><<<traceGrid[2], 128, stackBytes, st>>>(ctx->scene, w, depth + 1);
                }
            }
        }
        wfAccumulate<<<(nPix + 127) / 128, 128, 0, st>>>(w, pm, pix0, nPix, ppp, spp, exposure, d_rgb);
        PRT_CUDA_CHECK(ctx, cudaGetLastError());
        PRT_CUDA_CHECK(ctx, cudaMemcpyAsync(ctx->h_wfCounters + ctx->wfCountersPending * kCounterWords, w.counters, kCounterWords * sizeof(uint32_t),
                                            cudaMemcpyDeviceToHost, st));
        ctx->wfCountersPending++;
        ctx->counters.kernelLaunches += 1 + 15 + 14 + 14 + 1 + (parked ? 28 : 0);
        ctx->counters.primaryRays += (uint64_t)nPackets * 8u;
    }
    for (uint32_t k = 1; k < lanes; k++) {                                       // the context stream is the one callers synchronise on
        PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->evLane, laneStream[k]));
        PRT_CUDA_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->evLane, 0));
    }
    PRT_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    return PRT_OK;
}

static int renderLocal(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed, float exposure, float* d_rgb) {
    const uint32_t W = ctx->scene.cam.width;
    PixMap pm{x0, y0, x1 - x0, 0, 0, 0, y1, W, 0};
    return renderImpl(ctx, pm, (x1 - x0) * (y1 - y0), spp, seed, exposure, d_rgb);
}

}  // extern "C"

namespace prt_b200 {

// This rank's bands of the region into ctx->d_bandLocal (multi_gpu.cu gathers them).  Runs on the context's own thread
// (the caller for the leader / a single-GPU rank, the worker thread for a peer).
int renderBandsLocal(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t bandRows, uint32_t spp, uint32_t seed, float exposure) {
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    const uint32_t bw = x1 - x0, regionH = y1 - y0, world = (uint32_t)ctx->world;
    const uint32_t nb = (regionH + bandRows - 1) / bandRows, nbPerRank = (nb + world - 1) / world;
    if ((r = ensureBandBuffers(ctx, (size_t)nbPerRank * bandRows * bw * 3))) return r;
    PixMap pm{x0, y0, bw, bandRows, (uint32_t)ctx->rank, world, y1, bw, x0};
    return renderImpl(ctx, pm, nbPerRank * bandRows * bw, spp, seed, exposure, ctx->d_bandLocal);
}

}  // namespace prt_b200

extern "C" {

int prt_cuda_render_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed, float exposure, float* d_rgb) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (!ctx->hasCamera) return setError(ctx, PRT_ERR_STATE, "render: no camera set");
    if (!d_rgb || x1 <= x0 || y1 <= y0 || x1 > ctx->scene.cam.width || y1 > ctx->scene.cam.height) return setError(ctx, PRT_ERR_ARG, "render: bad region");
    if (ctx->world <= 1) return renderLocal(ctx, x0, y0, x1, y1, spp, seed, exposure, d_rgb);
    // sharded: every GPU renders its bands (leader + worker threads / this rank), then ONE collective + de-tile
    const uint32_t bandRows = bandRowsDefault();
    r = forEachDevice(ctx, [&](prt_ctx* c, int) -> int { return renderBandsLocal(c, x0, y0, x1, y1, bandRows, spp, seed, exposure); });
    if (r) return r;
    return gatherBands(ctx, x0, y0, x1, y1, bandRows, ctx->scene.cam.width, d_rgb);
}

int prt_cuda_render_bands_dev(prt_ctx* ctx, uint32_t bandRows, uint32_t rank, uint32_t world, uint32_t spp, uint32_t seed, float exposure, float* d_local) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (!ctx->hasCamera) return setError(ctx, PRT_ERR_STATE, "render_bands: no camera set");
    if (!d_local || bandRows == 0 || world == 0 || rank >= world) return setError(ctx, PRT_ERR_ARG, "render_bands: bad argument");
    const uint32_t W = ctx->scene.cam.width, H = ctx->scene.cam.height;
    const uint32_t nb = (H + bandRows - 1) / bandRows, nbPerRank = (nb + world - 1) / world;
    PixMap pm{0, 0, W, bandRows, rank, world, H, W, 0};
    return renderImpl(ctx, pm, nbPerRank * bandRows * W, spp, seed, exposure, d_local);
}

int prt_cuda_render(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed, float exposure, float* rgb) {
    if (!ctx) return PRT_ERR_ARG;
    DeviceGuard g(ctx->device);
    int r = commitScene(ctx);
    if (r) return r;
    if (!rgb || !ctx->hasCamera) return setError(ctx, PRT_ERR_ARG, "render: bad argument");
    const uint32_t W = ctx->scene.cam.width, H = ctx->scene.cam.height;
    if (x1 <= x0 || y1 <= y0 || x1 > W || y1 > H) return setError(ctx, PRT_ERR_ARG, "render: bad region");
    if ((r = ensureScratch(ctx, 3, (size_t)W * H * 12))) return r;
    if ((r = prt_cuda_render_dev(ctx, x0, y0, x1, y1, spp, seed, exposure, (float*)ctx->d_scratch[3]))) return r;
    // only the region's pixels cross PCIe; everything else in the caller's image is left alone
    const size_t pitch = (size_t)W * 12, off = ((size_t)y0 * W + x0) * 12;
    PRT_CUDA_CHECK(ctx, cudaMemcpy2DAsync((char*)rgb + off, pitch, (const char*)ctx->d_scratch[3] + off, pitch, (size_t)(x1 - x0) * 12, y1 - y0,
                                          cudaMemcpyDeviceToHost, ctx->stream));
    PRT_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0.0f;
    PRT_CUDA_CHECK(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->counters.lastKernelMs = ms;
    if (ctx->world > 1) { PRT_CUDA_CHECK(ctx, cudaEventElapsedTime(&ms, ctx->evG0, ctx->evG1)); ctx->counters.lastGatherMs = ms; ctx->gatherRecorded = false; }
    return PRT_OK;
}

}  // extern "C"
