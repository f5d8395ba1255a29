"""Host side of the opt-in fast traversal mode (prt_b200/csrc/bvh_build.cpp buildWide, prt_cuda.cu add_mesh), checked without a GPU on
a host-only context: the Wide8 tree covers every triangle exactly once in leaves of <= 4, every child's quantised box CONTAINS the
triangles under it with the margin the kernel's arithmetic needs (prt_wide.cuh: slab distances evaluated as one FMA on
1 + q * 2^-15 with the bias folded into a rounded per-node constant), and the "hidden from single-ray any-hit" flags are exactly
the triangles under a flat box of the reference tree (bvh.cpp:576-657 tests every box with the strict test)."""
import numpy as np
import pytest

import prt_b200
from prt_b200 import scenes

LEAF = 0x80000000


def _decode(rec):
    """one Wide8 record (24 uint32) -> origin[3], cell[3], count, axis, qlo[3][8], qhi[3][8], ref[8]"""
    b = rec.view(np.uint8)
    origin = rec[0:3].view(np.float32).astype(np.float64)
    exp = b[12:15].astype(np.int64)
    info = int(b[15])
    cell = np.ldexp(1.0, exp - 127)
    qlo = b[16:40].reshape(3, 8).astype(np.float64)
    qhi = b[40:64].reshape(3, 8).astype(np.float64)
    return origin, cell, info & 15, (info >> 4) & 3, qlo, qhi, rec[16:24]


def _walk(root, nodes, tris_pos, visit_leaf):
    """depth-first over the wide tree; returns per node-or-leaf reference the bounds of the triangles under it"""
    def bounds_of(ref, depth):
        if ref & LEAF:
            first, count = int(ref & 0x0fffffff), int((ref >> 28) & 7) + 1
            visit_leaf(first, count, depth)
            p = tris_pos[first:first + count].reshape(-1, 3)
            return p.min(axis=0), p.max(axis=0)
        origin, cell, count, axis, qlo, qhi, refs = _decode(nodes[ref])
        assert 2 <= count <= 8
        lo_all, hi_all = np.full(3, np.inf), np.full(3, -np.inf)
        cent = []
        for k in range(8):
            if k >= count:
                assert (qlo[:, k] == 255).all() and (qhi[:, k] == 0).all(), "unused slot must carry an inverted box"
                assert refs[k] == refs[0], "unused slot must repeat child 0's reference"
                continue
            lo, hi = bounds_of(int(refs[k]), depth + 1)
            blo, bhi = origin + qlo[:, k] * cell, origin + qhi[:, k] * cell
            # containment with the kernel's margin: 2^-9 cell for the folded bias + 2^-22 of the coordinates' magnitude for fp32 rounding
            mag = np.maximum(np.abs(lo), np.abs(hi))               # per axis, as the builder's margin
            slack = cell * 2.0 ** -9 + mag * 2.0 ** -21
            assert (blo <= lo).all() and (bhi >= hi).all(), f"node {ref} child {k}: quantised box does not contain its triangles"
            # ... and with room for the kernel's rounding wherever the plane is not pinned to the edge of the node's grid (there the
            # node's own padding is the margin, and flat boxes rely on the kernel's relative slack, prt_wide.cuh kWideSlack)
            coarse = cell * 254 >= mag * 2.0 ** -12                # a (nearly) flat axis of the NODE: its whole grid is thinner than any margin
            inner_lo, inner_hi = (qlo[:, k] > 0) & coarse, (qhi[:, k] < 255) & coarse
            assert (blo + slack <= lo)[inner_lo].all() and (bhi - slack >= hi)[inner_hi].all(), f"node {ref} child {k}: no margin for rounding"
            assert ((bhi - blo) <= (hi - lo) + 2.2 * cell + 8e-6 * 4 * np.maximum(mag, np.abs(origin) + 255 * cell)).all(), "quantised box is looser than two cells + margin"
            lo_all, hi_all = np.minimum(lo_all, lo), np.maximum(hi_all, hi)
            cent.append(0.5 * (lo[axis] + hi[axis]))
        return lo_all, hi_all
    return bounds_of(root, 0)


@pytest.mark.parametrize("name,make", [
    ("cornell + teapot", lambda: scenes.cornell_teapot(64, 64, teapot_refl=scenes.DIFFUSE, light="directional")),
    ("bunny stand-in", lambda: scenes.bunny_standin(64, 64, subdiv=4)),
    ("atrium (quarter detail)", lambda: scenes.atrium(64, 64, tex_size=32, detail=0.25)),
])
def test_wide_tree_invariants(name, make):
    desc = make()
    ctx = prt_b200.load_scene(desc, device=-1, fast=True)
    for mi, mesh in enumerate(desc["meshes"]):
        idx = np.asarray(mesh["indices"]).reshape(-1, 3)
        P = idx.shape[0]
        root, nodes, tris = ctx.bvh_fast(mi, P)
        pos = tris.reshape(P, 3, 4)[:, :, 0:3].astype(np.float64)
        w = tris.view(np.uint32).reshape(P, 3, 4)[:, :, 3]
        slots = w[:, 0]
        # every triangle exactly once, and the record IS that triangle (positions of primitive remap[slot])
        assert sorted(slots.tolist()) == list(range(P))
        _, remap = ctx.bvh(mi, P)
        want = np.asarray(mesh["positions"], np.float32)[idx[remap[slots]]]
        assert np.array_equal(pos.astype(np.float32), want)
        seen = np.zeros(P, np.int32)
        depth_max = [0]

        def leaf(first, count, depth):
            assert 1 <= count <= 4
            seen[first:first + count] += 1
            depth_max[0] = max(depth_max[0], depth)
        _walk(root, nodes, pos, leaf)
        assert (seen == 1).all(), "a triangle is under no leaf or under two"
        assert 7 * depth_max[0] + 1 <= 128, "deeper than the traversal stack"
        # hidden flags == triangles under a flat box of the REFERENCE tree
        n32, _ = ctx.bvh(mi, P)
        lo = n32[:, 0:3].view(np.float32); hi = n32[:, 3:6].view(np.float32)
        second = n32[:, 6]; bits = n32[:, 7]
        pc = (bits >> 26) & 15
        hidden_slot = np.zeros(P, bool)
        todo = [(0, False)]
        while todo:
            ni, inh = todo.pop()
            flat = inh or bool((lo[ni] == hi[ni]).any())
            if pc[ni] == 15:
                todo.append((ni + 1, flat)); todo.append((int(second[ni]), flat))
            elif flat:
                hidden_slot[int(second[ni]):int(second[ni]) + int(pc[ni])] = True
        got_hidden = (w[:, 1] >> 31).astype(bool)
        assert np.array_equal(got_hidden, hidden_slot[slots]), f"{name} mesh {mi}: hidden flags differ from the reference tree's flat boxes"
        if name.startswith("cornell") and mi == 0:
            assert got_hidden.any(), "the Cornell box has axis-aligned walls: some leaves of the reference tree must be flat"
    ctx.close()


def test_fast_structure_needs_the_build_option():
    ctx = prt_b200.load_scene(scenes.cornell_teapot(32, 32), device=-1)
    with pytest.raises(prt_b200.PrtError):
        ctx.bvh_fast(0, 36)
    ctx.close()
