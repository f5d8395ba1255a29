// Persistent "while-while" traversal with dynamic ray fetch (after Aila & Laine, "Understanding the Efficiency of Ray
// Traversal on GPUs"): each lane owns one ray at a time; lanes whose ray has terminated pull the next ray from the
// global queue inside the SAME loop that steps the other lanes' traversals, so a warp is not held hostage by its
// longest ray.  ncu on the first version of these kernels (one ray per lane until the whole warp finished) showed
// 3.3 / 32 active lanes per issued instruction on occlusion rays and 13.5 / 32 on primary rays (profiles/r1_*).
//
// Tried and rejected (v6, profiles/r1/r1_v6_vote_*): a majority-vote scheduler that runs ONE node step or ONE triangle
// test per iteration for whichever state holds more lanes.  Lanes per instruction rose 11.4 -> 15.1 on C3's bounce
// rays, but the three ballots per step and the lost leaf-loop ILP cost more (+26 % warp instructions, wfTrace
// 1.02 -> 1.39 ms, C2 7.3 -> 5.7 Grays/s), so the nested while-while form stays.
//
// Also tried (gpurun A/B, round 1): an "if-if" loop where every lane advances one unit (node step or one triangle) per
// round — slower on every workload (C2 7.2 -> 5.8 Grays/s, C5 258 -> 250 Msamples/s): the leaf loop's overlapped
// triangle loads are worth more than the idle lanes.  Kept: traceFlags bit 0, a node loop that yields to the leaf
// phase once the lanes parked at a leaf outnumber the lanes still descending (+7 % on C5, -2 % on C3; on for big scenes).
//
// Round 2: warp-local queue chunks (a warp reserves 64 / 128 / 256 consecutive queue entries and refills its idle lanes from them,
// so that warp-mates stay neighbours in pixel order) — slower everywhere: c3 1251 -> 1222 / 1225 / 1202 Msamples/s, c5 446 -> 443 /
// 430 / 428, c2 16.2 -> 13.9 / 12.6 / 11.1 Grays/s (profiles/r2/b1_ab.log): warps that still own entries keep running half-empty
// while others have nothing left to take.  The single shared queue head stays.
//
// Round 2, also without effect: only the first 6 / 8 / 12 / 16 stack entries in shared memory and the rest in local memory (shared
// memory and L1 share one 256 KB array per SM, and full-depth columns take 129 KB of it on the 4.9 M-triangle scene): c5 435.6 ->
// 429.7 / 433.4 / 435.4 / 435.5 Msamples/s, c4 and c3 within 0.5 %, and the extra branch per push / pop cost the batch kernels
// 2 - 6 % (profiles/r2/b2_stack_cap_ab.log).  Full-depth shared columns stay.
//
// Speculative traversal (park the first leaf, keep descending to the next one, re-cull what was found with the shrunk
// hit.t) was the fourth scheduling variant tried: C2 7.58 -> 5.99 Grays/s, C5 367 -> 268 Msamples/s — rejected too.
//
// The per-ray arithmetic and visiting rules are exactly those of prt_device.cuh (same slab / leaf routines), only the
// control structure differs: traversal state (current node, stack pointer, mesh index) lives in registers so it can be
// suspended between steps.
#pragma once
#include "prt_device.cuh"

#ifndef PRT_STACK_STRIDE
#define PRT_STACK_STRIDE 1      // > 1: the caller's stack pointer addresses entry k at stack[k * stride] (shared-memory stack, one column per thread)
#endif

namespace prt_b200 {

enum : int { kClosestSingle = 0, kClosestPacket = 1, kOcclude = 2 };

struct TravState {
    uint32_t cur;        // current child reference (leaf or interior)
    int sp;              // stack pointer
    uint32_t mesh;       // index of the BVH being traversed (Scene::intersect walks them in order, scene.cpp:55-59)
};

// Rays whose invDir has a non-finite component (a zero or denormal direction component: 0*inf = NaN can reach the slab
// min/max) never enter this loop: traceQueue resolves them at fetch time with the exact SSE-select routines of
// prt_device.cuh (slowTrace below).  Everything here may therefore use FMNMX, which agrees with min_ps/max_ps whenever
// no operand is NaN.
__device__ __forceinline__ Slab slabOf(const RayD& r, float lx, float ly, float lz, float hx, float hy, float hz) { return slab<false>(r, lx, ly, lz, hx, hy, hz); }
__device__ __forceinline__ float entryOf(const RayD&, const Slab& s) { return boxEntry<false>(s); }
__device__ __forceinline__ bool strictOf(const RayD&, const Slab& s, float maxT) { return boxStrict<false>(s, maxT); }
__device__ __forceinline__ bool soaOf(const RayD&, const Slab& s, float* e) { return boxSoa<false>(s, e); }

// Out-of-line so that the rare path costs the hot loop no registers.  Returns the occlusion bit for kOcclude.
template <int KIND>
__device__ bool slowTrace(const DScene& sc, const RayD& r, uint32_t revMask, bool pktRules, Hit& hit, uint32_t* callerStack, float* stackT) {
#if PRT_STACK_STRIDE == 1
    uint32_t* stack = callerStack;
#else
    uint32_t stack[kStackSize];                                   // the caller's stacks are strided: use private contiguous ones
    float stackTPrivate[KIND == kClosestPacket ? kStackSize : 1];
    (void)callerStack;
    stackT = stackTPrivate;
#endif
    if (KIND == kClosestSingle) { hit = sceneIntersectSingle(sc, r, stack); return false; }
    if (KIND == kClosestPacket) { hit = sceneIntersectPacket(sc, r, revMask, stack, stackT); return false; }
    return pktRules ? sceneOccluded<true>(sc, r, stack) : sceneOccluded<false>(sc, r, stack);
}

// Moves on to the next BVH whose root box the ray passes (bvh.cpp:458 / :480-485 / :601-605).  limit = current hit.t
// for closest hit, ray.maxT for occlusion.  Returns false when every BVH has been visited.
template <int KIND>
__device__ __forceinline__ bool enterNextMesh(const DScene& sc, const RayD& r, TravState& st, float limit, bool pktRules) {
    for (;;) {
        st.mesh++;
        if (st.mesh >= sc.meshCount) return false;
        const DMesh& m = sc.meshes[st.mesh];
        Slab s = slabOf(r, m.rootLo[0], m.rootLo[1], m.rootLo[2], m.rootHi[0], m.rootHi[1], m.rootHi[2]);
        bool pass;
        if (KIND == kClosestSingle) pass = strictOf(r, s, limit);
        else if (KIND == kClosestPacket) { float e; pass = soaOf(r, s, &e) && e < limit; }
        else { if (pktRules) { float e; pass = soaOf(r, s, &e) && e < limit; } else pass = strictOf(r, s, limit); }
        if (pass) { st.cur = m.rootRef; st.sp = 0; return true; }
    }
}

// One interior-node step.  Returns false when the ray has finished all BVHs.
template <int KIND>
__device__ __forceinline__ bool nodeStep(const DScene& sc, const RayD& r, TravState& st, Hit& hit, uint32_t revMask, bool pktRules,
                                         uint32_t* stack, float* stackT, TravStats* ts = nullptr) {
    if (ts) ts->nodeSteps++;
    const NodeD n = loadNode(sc, st.cur);
    Slab s0 = slabOf(r, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y);
    Slab s1 = slabOf(r, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w);
    bool popNeeded = false;
    if (KIND == kClosestSingle) {                                 // bvh.cpp:499-521
        // both hit: nearer child first, child 1 on ties, the other one pushed; written with selects — as branches
        // the two push arms ran at 2.3 / 32 lanes and were 6.8 % of C5's instructions (profiles/r1/r1_v7_c5_*)
        float t0 = entryOf(r, s0), t1 = entryOf(r, s1);
        const bool h0 = t0 < hit.t, h1 = t1 < hit.t;
        const bool first0 = h0 && (!h1 || t0 < t1);
        const uint32_t nearRef = first0 ? n.ref0 : n.ref1, farRef = first0 ? n.ref1 : n.ref0;
        if (h0 && h1) stack[(st.sp++) * PRT_STACK_STRIDE] = farRef;
        if (h0 || h1) st.cur = nearRef; else popNeeded = true;
    } else if (KIND == kClosestPacket) {                          // bvh.cpp:480-485, 523-534 (see prt_device.cuh)
        float e0, e1;
        bool h0 = soaOf(r, s0, &e0) && e0 < hit.t;
        bool h1 = soaOf(r, s1, &e1) && e1 < hit.t;
        bool rev = (revMask >> n.axis) & 1u;
        uint32_t firstRef = rev ? n.ref1 : n.ref0, secondRef = rev ? n.ref0 : n.ref1;
        bool hf = rev ? h1 : h0, hs = rev ? h0 : h1;
        float es = rev ? e0 : e1;
        if (hs) { stack[st.sp * PRT_STACK_STRIDE] = secondRef; stackT[st.sp * PRT_STACK_STRIDE] = es; st.sp++; }
        if (hf) st.cur = firstRef; else popNeeded = true;
    } else {                                                      // bvh.cpp:601-618
        bool h0, h1;
        if (pktRules) { float e; h0 = soaOf(r, s0, &e) && e < r.maxT; h1 = soaOf(r, s1, &e) && e < r.maxT; }
        else { h0 = strictOf(r, s0, r.maxT); h1 = strictOf(r, s1, r.maxT); }
        // Any hit is a boolean over the same reachable set whatever the visiting order, so the order is free: traceFlags bits 1-2
        // pick child0-first (0, the reference's order), front-to-back along the split axis (1) or back-to-front (2).
        const uint32_t order = (sc.traceFlags >> 1) & 3u;
        bool swap = false;
        if (order) {
            const float inv = n.axis == 0u ? r.ix : (n.axis == 1u ? r.iy : r.iz);
            swap = (order == 1u) == (inv < 0.0f);
        }
        const uint32_t firstRef = swap ? n.ref1 : n.ref0, secondRef = swap ? n.ref0 : n.ref1;
        const bool hf = swap ? h1 : h0, hs = swap ? h0 : h1;
        if (hf && hs) stack[(st.sp++) * PRT_STACK_STRIDE] = secondRef;
        if (hf || hs) st.cur = hf ? firstRef : secondRef; else popNeeded = true;
    }
    if (!popNeeded) return true;
    // pop (closest-packet re-applies "max_t0 < hit.t" with the current hit.t)
    for (;;) {
        if (st.sp == 0) return enterNextMesh<KIND>(sc, r, st, KIND == kOcclude ? r.maxT : hit.t, pktRules);
        --st.sp;
        if (KIND != kClosestPacket || stackT[st.sp * PRT_STACK_STRIDE] < hit.t) { st.cur = stack[st.sp * PRT_STACK_STRIDE]; return true; }
    }
}

template <int KIND>
__device__ __forceinline__ bool popAfterLeaf(const DScene& sc, const RayD& r, TravState& st, Hit& hit, bool pktRules, uint32_t* stack, float* stackT) {
    for (;;) {
        if (st.sp == 0) return enterNextMesh<KIND>(sc, r, st, KIND == kOcclude ? r.maxT : hit.t, pktRules);
        --st.sp;
        if (KIND != kClosestPacket || stackT[st.sp * PRT_STACK_STRIDE] < hit.t) { st.cur = stack[st.sp * PRT_STACK_STRIDE]; return true; }
    }
}

// Generic persistent loop.  Fetch: bool (uint32_t index, RayD& ray, uint32_t& revMask, bool& pktRules) loads ray `index`; false: the
// index carries no ray (the fetch has stored whatever result it stands for) and the lane takes another one next round.
// Commit: (uint32_t index, const Hit& hit, bool occluded, uint32_t tag) stores the result; tag is whatever Fetch put
// into its third argument (the packet's avgDir sign mask for packet-rule closest hit, a slot id for the wavefront).  `refill` = minimum number of idle lanes
// before the warp goes back to the queue (amortises the atomic and keeps fetched rays contiguous).
//
// PHASE splits the walk over the scene's BVHs (Scene::intersect / Scene::occluded visit them in insertion order, scene.cpp:55-59,
// :80-90) into two launches for scenes with several meshes.  PHASE 1 traverses mesh 0 only: a ray that goes on to a later mesh
// (its root box passes) is PARKED — aux(tag, hit) stores the hit found so far and marks the ray — and the lane takes a new ray.
// PHASE 2 runs over the parked rays: aux(tag, hit) reloads the partial hit and the ray resumes behind mesh 0 (the same root
// tests, now against the same hit.t, select the same next mesh).  The visiting order of every ray is unchanged; what changes is
// which rays share a warp: on Cornell + teapot the few lanes descending the 5 279-node teapot tree used to run the node loop
// at 6.7 / 32 lanes while the lanes that only had the 11-node box to cross waited (profiles/r1/r1_f2_c3_wftrace_lines.txt).
// PHASE 0 = one launch, no parking (single-mesh scenes, the batch entry points).
struct NoAux { __device__ __forceinline__ void operator()(uint32_t, Hit&) const {} };

template <int KIND, int PHASE = 0, typename Fetch, typename Commit, typename Aux = NoAux>
__device__ __forceinline__ void traceQueue(const DScene& sc, uint32_t n, uint32_t* counter, uint32_t* stack, float* stackT, Fetch fetch, Commit commit, Aux aux = Aux()) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned full = 0xffffffffu;
    bool has = false, exhausted = false;
    uint32_t idx = 0, revMask = 0;
    bool pktRules = false;
    RayD r; Hit hit; TravState st;
    st.cur = 0; st.sp = 0; st.mesh = 0;
    hit.t = 0.0f; hit.i = hit.j = hit.k = 0.0f; hit.tri = hit.meshId = 0;
#ifdef PRT_STATS
    TravStats tsv; tsv.rays = tsv.nodeSteps = tsv.leafVisits = tsv.triTests = tsv.uvLoads = 0u;
    TravStats* const ts = &tsv;
#else
    TravStats* const ts = nullptr;
#endif
    for (;;) {
        // ---- refill idle lanes ----
        unsigned idle = __ballot_sync(full, !has);
        if (idle != 0u && !exhausted && (__popc(idle) >= (int)((sc.traceFlags >> 8) & 0xffu) || idle == full)) {
            uint32_t cnt = __popc(idle), base = 0;
            if (lane == (unsigned)(__ffs(idle) - 1)) base = atomicAdd(counter, cnt);
            base = __shfl_sync(full, base, __ffs(idle) - 1);
            if (base + cnt >= n) exhausted = true;
            if (!has) {
                uint32_t my = base + __popc(idle & ((1u << lane) - 1u));
                if (my < n) {
                    idx = my;
                    if (!fetch(idx, r, revMask, pktRules)) {      // nothing to trace for this index (the fetch wrote its result itself)
                    } else if (r.nanSafe) {                       // exact SSE-select path, resolved here (see slabOf)
                        const bool occ = slowTrace<KIND>(sc, r, revMask, pktRules, hit, stack, stackT);
                        commit(idx, hit, occ, revMask);
                    } else {
                        if (ts) ts->rays++;
                        hit.t = r.maxT; hit.i = hit.j = hit.k = 0.0f; hit.tri = 0; hit.meshId = 0;
                        st.mesh = 0xffffffffu;
                        if (PHASE == 2) { aux(revMask, hit); st.mesh = 0u; }          // resume behind mesh 0 with what it found
                        has = enterNextMesh<KIND>(sc, r, st, KIND == kOcclude ? r.maxT : hit.t, pktRules);
                        if (!has) { if (KIND != kOcclude && hit.t == r.maxT) hit.t = -1.0f; commit(idx, hit, false, revMask); }
                        else if (PHASE == 1 && st.mesh != 0u) { aux(revMask, hit); has = false; }
                    }
                }
            }
            idle = __ballot_sync(full, !has);
        }
        if (idle == full && exhausted) break;
        // ---- descend: interior nodes ----
        if (sc.traceFlags & 1u) {
            // warp-uniform loop: step the lanes that sit on an interior node until none is left OR the lanes waiting
            // at a leaf outnumber them (then serve the leaves first; the node lanes resume next round).  Round 2 tried the fast
            // mode's rule here too (hand over when fewer than k = 6 / 10 / 14 / 20 lanes sit on a node): nothing on C4 / C5 (the
            // exact kernels are bound by the L1 data pipe there), -2 .. -15 % on C3 and c2 (profiles/r2/b4_exact_yield_ab.log).
            for (;;) {
                const bool inNode = has && !(st.cur & kLeafBit);
                const unsigned nm = __ballot_sync(full, inNode);
                if (nm == 0u) break;
                const unsigned lm = __ballot_sync(full, has && (st.cur & kLeafBit));
                if (__popc(nm) < __popc(lm)) break;
                if (inNode) {
                    has = nodeStep<KIND>(sc, r, st, hit, revMask, pktRules, stack, stackT, ts);
                    if (!has) { if (KIND != kOcclude && hit.t == r.maxT) hit.t = -1.0f; commit(idx, hit, false, revMask); }
                    else if (PHASE == 1 && st.mesh != 0u) { aux(revMask, hit); has = false; }
                }
            }
        } else {
            while (has && !(st.cur & kLeafBit)) {
                has = nodeStep<KIND>(sc, r, st, hit, revMask, pktRules, stack, stackT, ts);
                if (!has) { if (KIND != kOcclude && hit.t == r.maxT) hit.t = -1.0f; commit(idx, hit, false, revMask); }
                else if (PHASE == 1 && st.mesh != 0u) { aux(revMask, hit); has = false; }
            }
        }
        // ---- leaf ----
        if (has && (st.cur & kLeafBit)) {
            const DMesh& m = sc.meshes[st.mesh];
            if (KIND == kOcclude) {
                if (leafTest<true>(sc, m, st.mesh, st.cur, r, hit, r.maxT, ts)) { has = false; commit(idx, hit, true, revMask); }
            } else {
                leafTest<false>(sc, m, st.mesh, st.cur, r, hit, 0.0f, ts);
            }
            if (has) {
                has = popAfterLeaf<KIND>(sc, r, st, hit, pktRules, stack, stackT);
                if (!has) { if (KIND != kOcclude && hit.t == r.maxT) hit.t = -1.0f; commit(idx, hit, false, revMask); }
                else if (PHASE == 1 && st.mesh != 0u) { aux(revMask, hit); has = false; }
            }
        }
    }
#ifdef PRT_STATS
    if (sc.stats) {                                               // one atomic per warp and counter
        const uint32_t v[5] = {tsv.rays, tsv.nodeSteps, tsv.leafVisits, tsv.triTests, tsv.uvLoads};
#pragma unroll
        for (int f = 0; f < 5; f++) {
            unsigned long long x = v[f];
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(full, x, o);
            if (lane == 0 && x) atomicAdd(sc.stats + f * 3 + KIND, x);
        }
    }
#endif
}

}  // namespace prt_b200
