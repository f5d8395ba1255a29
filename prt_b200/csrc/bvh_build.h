// Host-side BVH builder + device-layout flattening.
//
// The tree is the reference's: centroid-binned SAH, 32 buckets x 3 axes, cost 0.125 + Nl*Al + Nr*Ar, leaves of
// <= 8 triangles, median fallback (BvhBuildNode::build, bvh.cpp:31-171), flattened depth-first with child0 at
// node+1 (Bvh::buildLinearBvhNodes, bvh.cpp:230-299).  It must come out node-for-node identical to the
// reference's because traversal order decides ties; tests/test_oracle_golden.py (test_bvh_build_matches_reference) and tests/test_oracle_vs_ref.py check that against the oracle and
// the reference.  What is new is how it is computed (per-triangle centroids/bounds computed once, prefix/suffix
// bucket sweeps, subtree tasks on all host cores) and the second, GPU-facing layout produced from it.
#pragma once
#include <cstdint>
#include <vector>

namespace prt_b200 {

// Reference-layout node (LinearBvhNode, bvh.h:49-62), kept for parity checks and as the flatten input.
struct Node32 {
    float lower[3];
    float upper[3];
    uint32_t primOrSecond;   // leaf: first index into remap; interior: index of child1
    uint32_t bits;           // triVectorIndex:26 | primCount:4 (0xf = interior) | splitAxis:2
    uint32_t primCount() const { return (bits >> 26) & 0xf; }
    uint32_t splitAxis() const { return bits >> 30; }
};
static_assert(sizeof(Node32) == 32, "Node32 must match LinearBvhNode");

// Device node: one record per INTERIOR reference node holding BOTH children's boxes, so a traversal step is one
// 64-byte fetch (four 128-bit loads) and leaves never cost a node fetch.
//   f[0..5]  child0 lower.xyz, upper.xyz      f[6..11] child1 lower.xyz, upper.xyz
//   ref0, ref1: child reference (see makeLeafRef / interior index); axis: reference splitAxis of THIS node
struct Node64 {
    float box[12];
    uint32_t ref0, ref1, axis, pad;
};
static_assert(sizeof(Node64) == 64, "Node64 must be 64 bytes");

constexpr uint32_t kLeafFlag = 0x80000000u;
// leaf ref: bit31 | (count-1) << 28 | first triangle slot (28 bits) in the mesh's leaf-ordered triangle array
inline uint32_t makeLeafRef(uint32_t first, uint32_t count) { return kLeafFlag | ((count - 1u) << 28) | first; }

struct BvhHost {
    std::vector<Node32> nodes32;     // reference DFS layout
    std::vector<uint32_t> remap;     // m_primRemapping: leaf-ordered triangle slot -> primitive id
    std::vector<Node64> nodes64;     // device layout, interior nodes in DFS order
    uint32_t rootRef = 0;            // interior index 0 or a leaf ref when the whole mesh is one leaf
    float rootLower[3], rootUpper[3];
    uint32_t maxDepth = 0;           // deepest stack the reference traversal can need
};

// positions: 3 floats per vertex; indices: 3 per primitive.  nthreads <= 0: all host cores.  maxLeaf = 8 is the reference's
// tree (kLaneCount on its AVX build, bvh.h:21); the fast traversal mode builds a second tree with smaller leaves.
void buildBvh(BvhHost& out, const uint32_t* indices, uint32_t primCount, const float* positions, int nthreads, int maxLeaf = 8);

// ---- opt-in fast traversal mode (flatten-time re-encoding to a wide, quantised layout) ----
// Wide8: up to eight children per record, their boxes quantised to 8 bits per plane on a power-of-two grid anchored at the
// record's origin — 96 bytes = three 256-bit loads for eight children, against 64 bytes for two in the exact layout.  The
// quantised boxes CONTAIN the exact ones (with a margin far above fp32 rounding), so a ray enters every child it enters in
// the exact tree; triangle tests are the exact ones, so hit t / i / j / k of a given triangle are bit-identical.  What can
// differ from the reference: which of two triangles at exactly the same t is reported (visiting order), and rays the
// reference's box tests reject at grazing incidence (flat boxes under its strict test).  Hence opt-in, never the default.
struct Wide8 {
    float origin[3];
    uint8_t exp[3];          // grid cell per axis = 2^(exp - 127) (the float's exponent byte)
    uint8_t info;            // bits 0-3: children used (1..8); bits 4-5: axis the children are ordered along (ascending centroid)
    uint8_t qlo[3][8];       // child boxes on the grid, lower corner (axis-major)
    uint8_t qhi[3][8];       // upper corner
    uint32_t ref[8];         // child references: Wide8 index, or makeLeafRef(first, count) into the fast triangle array; unused: a copy of ref[0] (box inverted, never entered)
};
static_assert(sizeof(Wide8) == 96, "Wide8 must be 96 bytes");

struct WideHost {
    std::vector<Wide8> nodes;
    uint32_t rootRef = 0;            // node 0, or a leaf ref when the mesh is one leaf
    uint32_t maxDepth = 0;           // levels of wide nodes
};
// Collapses the binary tree of `bin` (its leaves become the wide tree's leaves) into Wide8 records.
void buildWide(WideHost& out, const BvhHost& bin);

}  // namespace prt_b200
