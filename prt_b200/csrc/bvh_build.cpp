// Host BVH builder: the reference's binned-SAH tree (bvh.cpp:31-171), computed with per-triangle data cached
// once, a physically partitioned reference array (sequential memory), prefix/suffix bucket sweeps and subtree
// tasks on all host cores; then flattened to the reference layout (bvh.cpp:230-299) and to the device layout.
//
// Compiled with -ffp-contract=off: every float operation below is individually rounded, in the reference's
// order, so split decisions (and therefore the tree) are bit-identical to the reference builder's.
#include "bvh_build.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <deque>
#include <limits>
#include <mutex>
#include <thread>

namespace prt_b200 {
namespace {

struct PrimRef {          // 40 bytes, moved as a unit by the partition step
    float c[3];           // centroid: (1.0f/3) * (((0+v0)+v1)+v2)   (bvh.cpp:78-87)
    float lo[3], hi[3];   // triangle bounds
    uint32_t id;
};

struct Box {
    float lo[3], hi[3];
    void reset() { for (int k = 0; k < 3; k++) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; } }   // BBox::init, vecmath.cpp:46-52
    void merge(const float* l, const float* h) {                                            // BBox::merge, vecmath.cpp:63-73
        for (int k = 0; k < 3; k++) { lo[k] = std::fmin(lo[k], l[k]); hi[k] = std::fmax(hi[k], h[k]); }
    }
    float area() const {                                                                    // vecmath.cpp:75-79
        float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        return 2.0f * (ex * ey + ey * ez + ez * ex);
    }
};

struct BNode {
    Box box;
    BNode* child[2] = {nullptr, nullptr};
    int32_t primIndex = -1;   // -1: interior
    int32_t primCount = 0;
    int32_t splitAxis = 0;
};

constexpr uint32_t kBuckets = 32;    // bvh.cpp:57

struct Builder {
    PrimRef* refs;
    int maxThreads;
    int kMaxLeaf = 8;                // kLaneCount on the reference's AVX build (bvh.h:21); smaller for the fast-mode tree
    std::atomic<int> activeThreads{1};
    std::mutex arenaMutex;
    std::deque<std::deque<BNode>> arenas;   // one arena per task; deque never relocates

    std::deque<BNode>* newArena() {
        std::lock_guard<std::mutex> g(arenaMutex);
        arenas.emplace_back();
        return &arenas.back();
    }

    void build(BNode* n, int32_t start, int32_t end, std::deque<BNode>* arena) {
        n->box.reset();
        for (int32_t i = start; i <= end; i++) n->box.merge(refs[i].lo, refs[i].hi);      // calculateBounds, bvh.cpp:21-29
        const int32_t primCount = end - start + 1;
        if (primCount <= kMaxLeaf) { n->primIndex = start; n->primCount = primCount; return; }

        float extent[3];
        for (int k = 0; k < 3; k++) extent[k] = n->box.hi[k] - n->box.lo[k];

        // One pass bins all three axes (the reference makes one pass per axis; bucket contents are identical).
        uint32_t count[3][kBuckets];
        Box bounds[3][kBuckets];
        float splitExtent[3], lowerPos[3];
        for (int d = 0; d < 3; d++) {
            splitExtent[d] = extent[d] == 0.0f ? 0.0001f : extent[d];                        // bvh.cpp:67
            lowerPos[d] = n->box.lo[d];
            for (uint32_t b = 0; b < kBuckets; b++) { count[d][b] = 0; bounds[d][b].reset(); }
        }
        for (int32_t i = start; i <= end; i++) {
            const PrimRef& p = refs[i];
            for (int d = 0; d < 3; d++) {
                int32_t b = static_cast<int32_t>(kBuckets * (p.c[d] - lowerPos[d]) / splitExtent[d]);   // bvh.cpp:90
                if (b >= (int32_t)kBuckets) b = kBuckets - 1;
                if (b < 0) b = 0;
                count[d][b]++;
                bounds[d][b].merge(p.lo, p.hi);
            }
        }

        float lowestCost = std::numeric_limits<float>::max();
        uint32_t lowestDim = 0;
        int32_t lowestSplit = -1;
        for (uint32_t d = 0; d < 3; d++) {
            // suffix[i] = buckets i..31 merged; min/max merges are order independent, so these equal the
            // reference's per-split re-merges (bvh.cpp:97-109) exactly.
            Box suffix[kBuckets]; uint32_t suffixCount[kBuckets];
            Box acc; acc.reset(); uint32_t cnt = 0;
            for (int32_t b = kBuckets - 1; b >= 0; b--) { acc.merge(bounds[d][b].lo, bounds[d][b].hi); cnt += count[d][b]; suffix[b] = acc; suffixCount[b] = cnt; }
            Box left; left.reset(); uint32_t leftCount = 0;
            for (uint32_t i = 0; i < kBuckets - 1; i++) {
                left.merge(bounds[d][i].lo, bounds[d][i].hi); leftCount += count[d][i];
                float cost = 0.125f + (leftCount * left.area() + suffixCount[i + 1] * suffix[i + 1].area());   // bvh.cpp:111
                if (lowestCost > cost) { lowestDim = d; lowestCost = cost; lowestSplit = (int32_t)i; }          // NaN (0*inf) never wins
            }
        }

        const uint32_t dim = lowestDim;
        const float splitPos = lowerPos[dim] + (lowestSplit + 1) * splitExtent[dim] / kBuckets;               // bvh.cpp:124

        // Partition with the reference's exact swap sequence (bvh.cpp:137-149) so remap order matches.
        int32_t cursor;
        for (cursor = start; cursor <= end; cursor++) if (!(refs[cursor].c[dim] < splitPos)) break;
        for (int32_t i = cursor + 1; i <= end; i++) {
            if (refs[i].c[dim] < splitPos) { std::swap(refs[i], refs[cursor]); cursor++; }
        }
        int32_t mid = cursor - 1;
        if (mid <= start || end <= mid) mid = (start + end) / 2;                                               // bvh.cpp:151-154
        n->splitAxis = (int32_t)dim;

        arena->emplace_back(); BNode* c0 = &arena->back();
        arena->emplace_back(); BNode* c1 = &arena->back();
        n->child[0] = c0; n->child[1] = c1;

        // Subtree tasks: the two ranges are disjoint, the result does not depend on scheduling.
        const int32_t leftCount = mid - start + 1;
        bool spawn = false;
        if (leftCount > 16384 && (end - mid) > 16384) {
            int cur = activeThreads.load();
            while (cur < maxThreads) { if (activeThreads.compare_exchange_weak(cur, cur + 1)) { spawn = true; break; } }
        }
        if (spawn) {
            std::deque<BNode>* a2 = newArena();
            std::thread t([=]() { build(c0, start, mid, a2); activeThreads.fetch_sub(1); });
            build(c1, mid + 1, end, arena);
            t.join();
        } else {
            build(c0, start, mid, arena);
            build(c1, mid + 1, end, arena);
        }
    }
};

void flatten32(std::vector<Node32>& out, const BNode* n, uint32_t& triVec, uint32_t depth, uint32_t& maxDepth) {
    uint32_t me = (uint32_t)out.size();
    out.emplace_back();
    if (depth > maxDepth) maxDepth = depth;
    memcpy(out[me].lower, n->box.lo, 12); memcpy(out[me].upper, n->box.hi, 12);
    if (n->primIndex < 0) {
        out[me].bits = (0xfu << 26) | ((uint32_t)n->splitAxis << 30);
        flatten32(out, n->child[0], triVec, depth + 1, maxDepth);
        out[me].primOrSecond = (uint32_t)out.size();
        flatten32(out, n->child[1], triVec, depth + 1, maxDepth);
    } else {
        out[me].primOrSecond = (uint32_t)n->primIndex;
        out[me].bits = (triVec & 0x3ffffffu) | ((uint32_t)n->primCount << 26) | ((uint32_t)n->splitAxis << 30);
        triVec++;
    }
}

uint32_t childRef(const std::vector<Node32>& n32, uint32_t idx, const std::vector<uint32_t>& interiorIndex) {
    const Node32& c = n32[idx];
    if (c.primCount() == 0xf) return interiorIndex[idx];
    return makeLeafRef(c.primOrSecond, c.primCount());
}

}  // namespace

void buildBvh(BvhHost& out, const uint32_t* indices, uint32_t primCount, const float* positions, int nthreads, int maxLeaf) {
    if (nthreads <= 0) nthreads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<PrimRef> refs(primCount);
    auto fill = [&](uint32_t a, uint32_t b) {
        for (uint32_t i = a; i < b; i++) {
            PrimRef& r = refs[i];
            r.id = i;
            float t[3] = {0.0f, 0.0f, 0.0f};
            for (int k = 0; k < 3; k++) { r.lo[k] = FLT_MAX; r.hi[k] = -FLT_MAX; }
            for (int j = 0; j < 3; j++) {
                const float* v = positions + 3 * (size_t)indices[3 * (size_t)i + j];
                for (int k = 0; k < 3; k++) { t[k] = t[k] + v[k]; r.lo[k] = std::fmin(r.lo[k], v[k]); r.hi[k] = std::fmax(r.hi[k], v[k]); }
            }
            for (int k = 0; k < 3; k++) r.c[k] = (1.0f / 3) * t[k];
        }
    };
    if (primCount > 100000 && nthreads > 1) {
        std::vector<std::thread> th;
        uint32_t per = (primCount + nthreads - 1) / nthreads;
        for (int t = 0; t < nthreads; t++) { uint32_t a = t * per, b = std::min(primCount, a + per); if (a < b) th.emplace_back(fill, a, b); }
        for (auto& t : th) t.join();
    } else fill(0, primCount);

    Builder b;
    b.refs = refs.data();
    b.maxThreads = nthreads;
    b.kMaxLeaf = maxLeaf < 1 ? 1 : (maxLeaf > 8 ? 8 : maxLeaf);
    std::deque<BNode>* arena = b.newArena();
    arena->emplace_back();
    BNode* root = &arena->back();
    b.build(root, 0, (int32_t)primCount - 1, arena);

    out.remap.resize(primCount);
    for (uint32_t i = 0; i < primCount; i++) out.remap[i] = refs[i].id;

    out.nodes32.clear();
    uint32_t triVec = 0, maxDepth = 0;
    flatten32(out.nodes32, root, triVec, 0, maxDepth);
    out.maxDepth = maxDepth + 2;   // the reference pushes both children before popping one (bvh.cpp:509-534)

    // Device layout: one Node64 per interior node, numbered in DFS order.
    const auto& n32 = out.nodes32;
    std::vector<uint32_t> interiorIndex(n32.size(), 0);
    uint32_t ni = 0;
    for (size_t i = 0; i < n32.size(); i++) if (n32[i].primCount() == 0xf) interiorIndex[i] = ni++;
    out.nodes64.assign(ni, Node64{});
    for (size_t i = 0; i < n32.size(); i++) {
        if (n32[i].primCount() != 0xf) continue;
        Node64& d = out.nodes64[interiorIndex[i]];
        uint32_t c0 = (uint32_t)i + 1, c1 = n32[i].primOrSecond;
        memcpy(d.box + 0, n32[c0].lower, 12); memcpy(d.box + 3, n32[c0].upper, 12);
        memcpy(d.box + 6, n32[c1].lower, 12); memcpy(d.box + 9, n32[c1].upper, 12);
        d.ref0 = childRef(n32, c0, interiorIndex);
        d.ref1 = childRef(n32, c1, interiorIndex);
        d.axis = n32[i].splitAxis();
        d.pad = 0;
    }
    out.rootRef = n32[0].primCount() == 0xf ? 0u : makeLeafRef(n32[0].primOrSecond, n32[0].primCount());
    memcpy(out.rootLower, n32[0].lower, 12); memcpy(out.rootUpper, n32[0].upper, 12);
}

// ---- Wide8 collapse (fast traversal mode) ----
namespace {

struct WideBuild {
    const std::vector<Node32>& n32;
    std::vector<Wide8>& out;
    uint32_t maxDepth = 0;

    static float area(const Node32& n) {
        float ex = n.upper[0] - n.lower[0], ey = n.upper[1] - n.lower[1], ez = n.upper[2] - n.lower[2];
        return 2.0f * (ex * ey + ey * ez + ez * ex);
    }
    bool interior(uint32_t i) const { return n32[i].primCount() == 0xf; }

    // Emits the wide node for binary interior node `root` at out[self] (already reserved) and recurses.
    void emit(uint32_t root, uint32_t self, uint32_t depth) {
        if (depth > maxDepth) maxDepth = depth;
        // children set: start with the two binary children, repeatedly open the interior child with the largest box
        uint32_t kids[8]; int nk = 0;
        kids[nk++] = root + 1; kids[nk++] = n32[root].primOrSecond;
        while (nk < 8) {
            int best = -1; float bestA = -1.0f;
            for (int k = 0; k < nk; k++) if (interior(kids[k])) { float a = area(n32[kids[k]]); if (a > bestA) { bestA = a; best = k; } }
            if (best < 0) break;
            uint32_t open = kids[best];
            kids[best] = open + 1; kids[nk++] = n32[open].primOrSecond;
        }
        const Node32& me = n32[root];
        // order the children along the axis of the node's largest extent by ascending centroid: traversal visits them front to
        // back for a ray travelling in +axis and back to front otherwise (the kernel reverses the order for -axis rays)
        int axis = 0;
        { float e[3] = {me.upper[0] - me.lower[0], me.upper[1] - me.lower[1], me.upper[2] - me.lower[2]}; if (e[1] > e[axis]) axis = 1; if (e[2] > e[axis]) axis = 2; }
        std::sort(kids, kids + nk, [&](uint32_t a, uint32_t b) {
            float ca = n32[a].lower[axis] + n32[a].upper[axis], cb = n32[b].lower[axis] + n32[b].upper[axis];
            return ca < cb || (ca == cb && a < b);
        });
        Wide8 w{};
        for (int d = 0; d < 3; d++) {
            // grid anchored a little below the node's lower corner, cell = power of two covering the padded extent in <= 254 steps
            float ext = me.upper[d] - me.lower[d];
            float pad = ext * (1.0f / 4096.0f) + std::fabs(me.lower[d]) * 1e-6f + 1e-30f;
            float org = me.lower[d] - pad;
            float span = (me.upper[d] + pad) - org;
            int e = 0;
            std::frexp(span / 254.0f, &e);                        // span/254 = m * 2^e, m in [0.5, 1) -> cell 2^e >= span/254
            int biased = e + 127;
            if (biased < 1) biased = 1;
            if (biased > 230) biased = 230;                       // the kernel scales the cell by 2^15 (prt_wide.cuh, planeByte)
            w.origin[d] = org;
            w.exp[d] = (uint8_t)biased;
            const float cell = std::ldexp(1.0f, biased - 127);
            for (int k = 0; k < 8; k++) {
                if (k >= nk) { w.qlo[d][k] = 255; w.qhi[d][k] = 0; continue; }        // empty slot: inverted box, never entered
                const Node32& c = n32[kids[k]];
                // margin: 2 % of a cell plus a few ulps of the coordinates' magnitude — the kernel evaluates the slab distances as
                // q * (cell * invDir) + (origin - rayOrg) * invDir, whose rounding error scales with the distance to the ray origin
                // (shadow rays start 2 scene radii away), not with the cell
                const double mag = std::fmax(std::fmax(std::fabs((double)me.lower[d]), std::fabs((double)me.upper[d])), (double)span);
                // + 2^-9 cell for the bias the kernel folds into its per-node constant (prt_wide.cuh, planeByte)
                const double m = 0.025 + 8e-6 * mag / (double)cell;
                double lo = ((double)c.lower[d] - (double)org) / (double)cell - m, hi = ((double)c.upper[d] - (double)org) / (double)cell + m;
                long ql = (long)std::floor(lo), qh = (long)std::ceil(hi);
                if (ql < 0) ql = 0;
                if (qh > 255) qh = 255;
                if (qh < ql) qh = ql;
                w.qlo[d][k] = (uint8_t)ql; w.qhi[d][k] = (uint8_t)qh;
            }
        }
        w.info = (uint8_t)(nk | (axis << 4));
        // interior children get consecutive records
        uint32_t childIndex[8];
        for (int k = 0; k < nk; k++) if (interior(kids[k])) { childIndex[k] = (uint32_t)out.size(); out.emplace_back(); }
        for (int k = 0; k < 8; k++) {
            // unused slot: its inverted box is never entered; should rounding ever let one through, the reference must still be
            // a valid one (the kernel pushes without looking at it) — child 0's, a second visit of which changes nothing
            if (k >= nk) { w.ref[k] = w.ref[0]; continue; }
            w.ref[k] = interior(kids[k]) ? childIndex[k] : makeLeafRef(n32[kids[k]].primOrSecond, n32[kids[k]].primCount());
        }
        out[self] = w;
        for (int k = 0; k < nk; k++) if (interior(kids[k])) emit(kids[k], childIndex[k], depth + 1);
    }
};

}  // namespace

void buildWide(WideHost& out, const BvhHost& bin) {
    out.nodes.clear();
    out.maxDepth = 0;
    const auto& n32 = bin.nodes32;
    if (n32.empty()) return;
    if (n32[0].primCount() != 0xf) { out.rootRef = makeLeafRef(n32[0].primOrSecond, n32[0].primCount()); return; }
    WideBuild wb{n32, out.nodes};
    out.nodes.emplace_back();
    wb.emit(0, 0, 1);
    out.rootRef = 0;
    out.maxDepth = wb.maxDepth;
}

}  // namespace prt_b200
