// Fast traversal mode (opt-in): single-ray closest hit / any hit over the Wide8 layout (bvh_build.h) — up to eight
// children per 96-byte record, boxes quantised to 8 bits on a power-of-two grid.  Same persistent while-while loop with dynamic
// ray fetch as prt_trace.cuh, same exact triangle test (prt_device.cuh, triTest), a different node step:
//
//   * one node = three 256-bit loads for eight children (the exact layout needs one 64-byte record per two children: on the scenes
//     that overflow L1 the kernel is bound by the L1 data pipe, i.e. by load instructions x distinct lines, profiles/r2);
//   * the fast mode's tree is its own (leaves of <= 4 triangles instead of the reference's 8: half the triangle fetches per ray);
//   * children are stored in ascending order along the node's longest axis; the step walks them back to front for the ray's
//     direction and pushes the ones it enters with their entry distance, so the nearest is popped first and an entry that a
//     closer hit has made pointless is dropped at pop without touching memory;
//   * slab distances use FMA (no parity contract here: the quantised boxes are conservative by a margin far above the rounding).
//
// What it does NOT promise (hence opt-in): the reference's choice between two triangles at exactly the same t, and its rejections
// at grazing incidence (a ray passing exactly along a box face).  One rejection rule IS reproduced because whole surfaces hang on
// it: the reference's single-ray any-hit test is strict, so a box that is flat along an axis (an axis-aligned wall's leaf) is
// never entered and the triangles under it never occlude; they carry a flag here (triTest).  tests/test_gpu_fast_mode.py
// enumerates every differing ray.
#pragma once
#include "prt_device.cuh"

namespace prt_b200 {

struct WideNode {
    float ox, oy, oz;
    uint32_t hdr;                 // exp[3] | info << 24
    uint32_t qlo[3][2], qhi[3][2];
    uint32_t ref[8];
};

__device__ __forceinline__ WideNode loadWide(const DScene& sc, uint32_t idx) {
    const float4* p = sc.fastNodes + 6 * (size_t)idx;
    WideNode n;
    uint32_t w[24];
#pragma unroll
    for (int k = 0; k < 3; k++)
        asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
            : "=r"(w[8 * k]), "=r"(w[8 * k + 1]), "=r"(w[8 * k + 2]), "=r"(w[8 * k + 3]), "=r"(w[8 * k + 4]), "=r"(w[8 * k + 5]), "=r"(w[8 * k + 6]), "=r"(w[8 * k + 7])
            : "l"(p + 2 * k));
    n.ox = __uint_as_float(w[0]); n.oy = __uint_as_float(w[1]); n.oz = __uint_as_float(w[2]); n.hdr = w[3];
#pragma unroll
    for (int d = 0; d < 3; d++) { n.qlo[d][0] = w[4 + 2 * d]; n.qlo[d][1] = w[5 + 2 * d]; n.qhi[d][0] = w[10 + 2 * d]; n.qhi[d][1] = w[11 + 2 * d]; }
#pragma unroll
    for (int k = 0; k < 8; k++) n.ref[k] = w[16 + k];
    return n;
}

// byte `k` (compile-time) of a 64-bit plane array as the float 1 + q * 2^-15: PRMT drops the byte into bits 8..15 of 0x3F800000.
// The slab distance q * s + b is then one FMA, (1 + q * 2^-15) * (2^15 * s) + (b - 2^15 * s), with the bias folded into the
// per-node constants: no int-to-float conversion and no subtraction per plane.  The folded constant is rounded once, an error of
// at most 2^-24 * (|b| + 2^15 |s|), i.e. 2^-9 of a grid cell plus the usual few ulps of the distance to the ray origin — inside
// the margin the builder quantises the boxes with (bvh_build.cpp, buildWide).
template <int K>
__device__ __forceinline__ float planeByte(const uint32_t (&q)[2], uint32_t one) {
    // `one` = 0x3F800000 in a register, the selector an immediate: PRMT has a single immediate slot, and with both operands constant
    // the compiler spent it on 0x3F800000 and rebuilt the four selectors in registers at every use (80 moves per node step)
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(q[K >> 2]), "r"(one), "n"(0x7604 | ((K & 3) << 4)));
    return __uint_as_float(d);
}

// A hit lying exactly on a face of a flat box (an axis-aligned wall) has slab distances equal up to rounding: both culls
// (child box vs limit, stacked entry vs hit.t) leave a relative slack of 2^-18 so that rounding can only cost an extra visit.
constexpr float kWideSlack = 1.0f + 1.0f / 262144.0f;

// Traversal stack of one thread: entry k = (ref, entry distance) at sm[k * 128] (shared memory, one column per thread, kWideCap
// entries).  A node step pushes up to eight entries with nothing but predicated shared-memory stores; the rare ray that would run
// out of room parks its eight OLDEST entries in local memory before the step (wideSpill) and takes them back when the shared
// part has drained (wideRefill).  kWideCap + kWideDeep entries cover (kWideCap + kWideDeep - 1) / 7 levels of wide nodes.
#ifndef PRT_WIDE_CAP
#define PRT_WIDE_CAP 24                                           // scripts/build_wide_cap_variant.sh builds with 16: spills as soon as a stack passes 8
#endif
constexpr int kWideCap = PRT_WIDE_CAP, kWideDeep = 128 - PRT_WIDE_CAP;
static_assert(kWideCap >= 16 && kWideDeep % 8 == 0, "a spill moves the 8 oldest entries: it must find at least 8 (trigger at kWideCap - 7 entries)");
struct WideStack {
    uint2* sm;
    uint2* deep;
};
static __device__ __noinline__ void wideSpill(uint2* sm, uint2* deep, int sp, int spilled) {
    for (int k = 0; k < 8; k++) deep[spilled + k] = sm[k * 128];
    for (int k = 8; k < sp; k++) sm[(k - 8) * 128] = sm[k * 128];
}
static __device__ __noinline__ void wideRefill(uint2* sm, const uint2* deep, int spilled) {
    for (int k = 0; k < 8; k++) sm[k * 128] = deep[spilled - 8 + k];
}

// One child slot K (compile-time) of the oriented arrays: pushed with its entry distance when its quantised box is crossed before
// `limit`.  No control flow: the comparison feeds the stores' predicate and the stack pointer.
template <int K>
__device__ __forceinline__ void wideChild(const WideNode& n, bool rev, const uint32_t (&nx)[2], const uint32_t (&ny)[2], const uint32_t (&nz)[2],
                                          const uint32_t (&fx)[2], const uint32_t (&fy)[2], const uint32_t (&fz)[2],
                                          float sx, float sy, float sz, float bx, float by, float bz, float limit, uint32_t one, uint32_t& top) {
    const float tn = fmaxf(fmaxf(fmaf(planeByte<K>(nx, one), sx, bx), fmaf(planeByte<K>(ny, one), sy, by)), fmaxf(fmaf(planeByte<K>(nz, one), sz, bz), 0.0f));
    const float tf = fminf(fminf(fmaf(planeByte<K>(fx, one), sx, bx), fmaf(planeByte<K>(fy, one), sy, by)), fminf(fmaf(planeByte<K>(fz, one), sz, bz), limit));
    if (tn <= tf * kWideSlack) {                                  // empty slots have inverted boxes: never
        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(top), "r"(rev ? n.ref[7 - K] : n.ref[K]), "r"(__float_as_uint(tn)) : "memory");
        top += 128u * 8u;
    }
}

// reverses the order of the eight bytes of a plane array when `rev`
__device__ __forceinline__ void orient(uint32_t (&q)[2], bool rev) {
    const uint32_t a = rev ? q[1] : q[0], b = rev ? q[0] : q[1], sel = rev ? 0x0123u : 0x3210u;
    q[0] = __byte_perm(a, 0u, sel); q[1] = __byte_perm(b, 0u, sel);
}

// (Measured and dropped: moving the entered child with the smallest entry distance to the top of the stack after the eight tests,
// instead of trusting the ordering axis — C5 1085 -> 1117, C4 1184 -> 1169, C3 1380 -> 1362 Msamples/s, profiles/r2/b9_wide_nearest_ab.log.)
//
// One wide node: tests the eight children far to near along the node's ordering axis and pushes the entered ones with their entry
// distances, so that the caller's pop takes the nearest first and drops, without touching memory, what a closer hit has made
// pointless by then.  limit: hit.t (closest hit) or ray.maxT (any hit).  The caller guarantees room for eight entries.
__device__ __forceinline__ void wideStep(const DScene& sc, const RayD& r, uint32_t nodeIdx, float limit, uint32_t one, uint32_t& top) {
    const WideNode n = loadWide(sc, nodeIdx);
    // grid -> ray parameter: t = q * (cell * invDir) + (origin - rayOrg) * invDir, evaluated on planeByte's 1 + q * 2^-15
    const float cx = __uint_as_float(((n.hdr & 0xffu) + 15u) << 23), cy = __uint_as_float((((n.hdr >> 8) & 0xffu) + 15u) << 23),
                cz = __uint_as_float((((n.hdr >> 16) & 0xffu) + 15u) << 23);                                  // 2^15 * cell
    const float sx = cx * r.ix, sy = cy * r.iy, sz = cz * r.iz;
    const float bx = (n.ox - r.ox) * r.ix - sx, by = (n.oy - r.oy) * r.iy - sy, bz = (n.oz - r.oz) * r.iz - sz;
    const bool negx = r.ix < 0.0f, negy = r.iy < 0.0f, negz = r.iz < 0.0f;
    // children are stored in ascending order along `axis`: a ray travelling in -axis meets them in reverse, so the arrays are
    // reversed for it and the code below always walks slots 7..0 (far to near), one code path for every lane
    const uint32_t axis = (n.hdr >> 28) & 3u;
    const bool rev = axis == 0u ? negx : (axis == 1u ? negy : negz);
    uint32_t nx[2], ny[2], nz[2], fx[2], fy[2], fz[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {                                 // near / far plane arrays per axis by the ray's sign
        nx[h] = negx ? n.qhi[0][h] : n.qlo[0][h]; fx[h] = negx ? n.qlo[0][h] : n.qhi[0][h];
        ny[h] = negy ? n.qhi[1][h] : n.qlo[1][h]; fy[h] = negy ? n.qlo[1][h] : n.qhi[1][h];
        nz[h] = negz ? n.qhi[2][h] : n.qlo[2][h]; fz[h] = negz ? n.qlo[2][h] : n.qhi[2][h];
    }
    orient(nx, rev); orient(ny, rev); orient(nz, rev); orient(fx, rev); orient(fy, rev); orient(fz, rev);
#define PRT_WIDE_CHILD(K) wideChild<K>(n, rev, nx, ny, nz, fx, fy, fz, sx, sy, sz, bx, by, bz, limit, one, top)
    PRT_WIDE_CHILD(7); PRT_WIDE_CHILD(6); PRT_WIDE_CHILD(5); PRT_WIDE_CHILD(4); PRT_WIDE_CHILD(3); PRT_WIDE_CHILD(2); PRT_WIDE_CHILD(1); PRT_WIDE_CHILD(0);
#undef PRT_WIDE_CHILD
}

// Rays with a non-finite invDir component (a zero direction component) keep the exact routines, as in prt_trace.cuh.
// TAG: every kernel of a translation unit instantiates the loop with its own tag.  Two entry points calling the SAME out-of-line
// instance crash ptxas 12.9 (segmentation fault on the file as a whole; each entry alone compiles).
template <bool OCCLUDE, int TAG>
__device__ __noinline__ bool wideSlowTrace(const DScene& sc, const RayD& r, Hit& hit, bool pkt) {
    uint32_t stack[kStackSize];
    if (!OCCLUDE) { hit = sceneIntersectSingle(sc, r, stack); return false; }
    return pkt ? sceneOccluded<true>(sc, r, stack) : sceneOccluded<false>(sc, r, stack);
}

// Next mesh whose root box the ray passes (the meshes' exact root boxes and the reference's root tests, scene.cpp:47-63 / 69-94;
// bvh.cpp:458 strict for single-ray rules, the SoA test for packet-rule rays, bvh.cpp:480-485 / 601-605).
template <bool OCCLUDE>
__device__ __forceinline__ bool wideEnterNextMesh(const DScene& sc, const RayD& r, uint32_t& mesh, uint32_t& cur, float limit, bool pkt) {
    for (;;) {
        mesh++;
        if (mesh >= sc.meshCount) return false;
        const DMesh& m = sc.meshes[mesh];
        const Slab s = slab<false>(r, m.rootLo[0], m.rootLo[1], m.rootLo[2], m.rootHi[0], m.rootHi[1], m.rootHi[2]);
        float e;
        const bool soa = boxSoa<false>(s, &e) && e < limit, strict = boxStrict<false>(s, limit);
        if (pkt ? soa : strict) { cur = m.fastRootRef; return true; }
    }
}

// Measured and dropped: the renderer's PRIMARY rays through this loop (generation kernel + a third persistent launch, packet-rule rays).
// Coherent camera rays are not bound by the L1 data pipe in the exact kernel, so the wide nodes bought nothing and the extra pass
// over the ray state cost: C5 908 -> 850, C3 1213 -> 1175, C4 1130 -> 1140 Msamples/s (profiles/r2/f3_fast_primary_summary.log); again
// with the hand-over rule below in place: C5 1092 -> 1017, C3 1351 -> 1312, C4 1190 -> 1198 (profiles/r2/b5_fast_primary_ab.log).
// wfPrimary stays exact in both modes.
//
// Persistent loop with dynamic ray fetch over the Wide8 trees.  Fetch / Commit as in traceQueue (prt_trace.cuh).
template <bool OCCLUDE, int TAG = 0, typename Fetch, typename Commit>
__device__ __forceinline__ void traceQueueWide(const DScene& sc, uint32_t n, uint32_t* counter, const WideStack& stack, Fetch fetch, Commit commit) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned full = 0xffffffffu;
    bool has = false, exhausted = false;
    uint32_t idx = 0, tag = 0, cur = 0, mesh = 0;
    // shared-space byte address of the next free entry (32-bit arithmetic, st.shared / ld.shared); empty when top == bottom
    const uint32_t bottom = (uint32_t)__cvta_generic_to_shared(stack.sm);
    uint32_t top = bottom;
    int spilled = 0;
    // 0x3F800000 for planeByte in a register the compiler cannot fold into an immediate (meshCount < 2^31)
    const uint32_t one = 0x3F800000u + (sc.meshCount >> 31);
    bool pktRules = false;
    RayD r; Hit hit;
    hit.t = 0.0f; hit.i = hit.j = hit.k = 0.0f; hit.tri = hit.meshId = 0;
#ifdef PRT_STATS
    // counting build: the same counters as traceQueue, in the classes of the single-ray rules (closest hit -> kClosestSingle,
    // any hit -> kOcclude); nodeSteps counts 96-byte Wide8 records here
    TravStats tsv; tsv.rays = tsv.nodeSteps = tsv.leafVisits = tsv.triTests = tsv.uvLoads = 0u;
    TravStats* const ts = &tsv;
#else
    TravStats* const ts = nullptr;
#endif
    auto finish = [&]() { if (!OCCLUDE && hit.t == r.maxT) hit.t = -1.0f; commit(idx, hit, false, tag); has = false; };
    // pops the next entry still worth visiting, moving on to the next mesh when the stack is empty; ends the ray when nothing is left
    auto pop = [&]() {
        for (;;) {
            if (top == bottom) {
                if (spilled) {
#ifdef PRT_WIDE_CHECK
                    if (spilled < 8 || (spilled & 7)) __trap();
#endif
                    wideRefill(stack.sm, stack.deep, spilled); spilled -= 8; top = bottom + 8u * 1024u; continue;
                }
                if (!wideEnterNextMesh<OCCLUDE>(sc, r, mesh, cur, OCCLUDE ? r.maxT : hit.t, pktRules)) finish();
                return;
            }
            top -= 1024u;
            uint2 e;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e.x), "=r"(e.y) : "r"(top) : "memory");
            if (OCCLUDE || __uint_as_float(e.y) < hit.t * kWideSlack) { cur = e.x; return; }
        }
    };
    for (;;) {
        unsigned idle = __ballot_sync(full, !has);
        if (idle != 0u && !exhausted && (__popc(idle) >= (int)((sc.traceFlags >> 8) & 0xffu) || idle == full)) {
            uint32_t cnt = __popc(idle), base = 0;
            if (lane == (unsigned)(__ffs(idle) - 1)) base = atomicAdd(counter, cnt);
            base = __shfl_sync(full, base, __ffs(idle) - 1);
            if (base + cnt >= n) exhausted = true;
            if (!has) {
                const uint32_t my = base + __popc(idle & ((1u << lane) - 1u));
                if (my < n) {
                    idx = my;
                    if (!fetch(idx, r, tag, pktRules)) {
                    } else if (r.nanSafe) {
                        // copies: the out-of-line call takes addresses, and `r` / `hit` themselves must stay in registers (taking
                        // theirs cost 10 local stores per fetched ray: 19 M STL and 3.7 GB of L2 writes per C4 launch, profiles/r2 n2)
                        const RayD r2 = r; Hit h2;
                        const bool occ = wideSlowTrace<OCCLUDE, TAG>(sc, r2, h2, pktRules);
                        commit(idx, h2, occ, tag);
                    } else {
                        hit.t = r.maxT; hit.i = hit.j = hit.k = 0.0f; hit.tri = 0; hit.meshId = 0;
                        if (ts) ts->rays++;
                        mesh = 0xffffffffu; has = true; spilled = 0;
                        top = bottom;
                        if (!wideEnterNextMesh<OCCLUDE>(sc, r, mesh, cur, r.maxT, pktRules)) finish();
                    }
                }
            }
            idle = __ballot_sync(full, !has);
        }
        if (idle == full && exhausted) break;
        // ---- interior nodes ----
        // Warp-uniform loop: the lanes sitting on a wide node step it; the loop hands over to the leaf phase when no lane is left on
        // a node or — traceFlags bits 16-23 — when too few are: 1 = fewer than the lanes waiting at a leaf, k >= 2 = fewer than k
        // while any lane waits at a leaf (a node step is ~300 instructions whatever the number of lanes that execute it).
        {
            const uint32_t wy = (sc.traceFlags >> 16) & 0xffu;
            for (;;) {
                const bool inNode = has && !(cur & kLeafBit);
                const unsigned nm = __ballot_sync(full, inNode);
                if (nm == 0u) break;
                if (wy) {
                    const unsigned lm = __ballot_sync(full, has && (cur & kLeafBit));
                    if (lm != 0u && (wy == 1u ? __popc(nm) < __popc(lm) : __popc(nm) < (int)wy)) break;
                }
                if (inNode) {
                    if (top > bottom + (uint32_t)(kWideCap - 8) * 1024u) {
#ifdef PRT_WIDE_CHECK
                        if (spilled + 8 > kWideDeep) __trap();      // checking build (scripts/build_wide_cap_variant.sh): the local part would overflow
#endif
                        wideSpill(stack.sm, stack.deep, (int)((top - bottom) >> 10), spilled); top -= 8u * 1024u; spilled += 8;
                    }
                    if (ts) ts->nodeSteps++;
#ifdef PRT_WIDE_CHECK
                    if (top < bottom || top > bottom + (uint32_t)(kWideCap - 8) * 1024u || ((top - bottom) & 1023u)) __trap();   // room for eight pushes
#endif
                    wideStep(sc, r, cur, OCCLUDE ? r.maxT : hit.t, one, top);
#ifdef PRT_WIDE_CHECK
                    if (top > bottom + (uint32_t)kWideCap * 1024u) __trap();
#endif
                    pop();
                }
            }
        }
        // ---- leaf: the exact triangle test over the fast tree's triangle records ----
        // (Measured and dropped: one triangle per lane and iteration, popping inside the loop so that a lane landing on another leaf
        // stays in it — C4 1190 -> 1138, C5 1091 -> 1001, C3 1383 -> 1256 Msamples/s, profiles/r2/b10_wide_leaf_loop_ab.log.)
        if (has && (cur & kLeafBit)) {
            const DMesh& m = sc.meshes[mesh];
            if (OCCLUDE) {
                // single-ray-rule any hit never sees triangles under the reference tree's flat boxes (triTest); packet-rule any hit does
                if (leafTest<true, true>(sc, m, mesh, cur, r, hit, r.maxT, ts, !pktRules)) { has = false; commit(idx, hit, true, tag); }
            } else {
                leafTest<false, true>(sc, m, mesh, cur, r, hit, 0.0f, ts);
            }
            if (has) pop();
        }
    }
#ifdef PRT_STATS
    if (sc.stats) {                                               // one atomic per warp and counter
        const uint32_t v[5] = {tsv.rays, tsv.nodeSteps, tsv.leafVisits, tsv.triTests, tsv.uvLoads};
#pragma unroll
        for (int f = 0; f < 5; f++) {
            unsigned long long x = v[f];
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(full, x, o);
            if (lane == 0 && x) atomicAdd(sc.stats + f * 3 + (OCCLUDE ? 2 : 0), x);
        }
    }
#endif
}

}  // namespace prt_b200
