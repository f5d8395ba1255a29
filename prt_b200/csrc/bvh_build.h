// Host-side BVH builder + device-layout flattening.
//
// The tree is the reference's: centroid-binned SAH, 32 buckets x 3 axes, cost 0.125 + Nl*Al + Nr*Ar, leaves of
// <= 8 triangles, median fallback (BvhBuildNode::build, bvh.cpp:31-171), flattened depth-first with child0 at
// node+1 (Bvh::buildLinearBvhNodes, bvh.cpp:230-299).  It must come out node-for-node identical to the
// reference's because traversal order decides ties; tests/test_bvh_build.py checks that against the oracle and
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

// positions: 3 floats per vertex; indices: 3 per primitive.  nthreads <= 0: all host cores.
void buildBvh(BvhHost& out, const uint32_t* indices, uint32_t primCount, const float* positions, int nthreads);

}  // namespace prt_b200
