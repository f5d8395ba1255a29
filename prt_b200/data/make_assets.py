"""Generates prt_b200/data/{teapot,cornell}.npz from the reference's own loader / sample model.

Run ONCE in the build container (needs /root/reference and oracle/_ref/libprt_ref.so):
    python prt_b200/data/make_assets.py
* teapot.npz  = Mesh::loadObj(data/teapot/teapot.obj) (mesh.cpp:151-209) after the setupCornellBox transform
  s=0.005, t=(-0.5,0,0.5) (main.cpp:40-44) and Mesh::calculateVertexNormals (mesh.cpp:108-149):
  indices (15704,3) u32, positions/normals (7850,3) f32.
* cornell.npz = SampleModels::getCornellBox(true) (sample_models.cpp:13-209): 36 prims, 72 verts, 18 materials.
The arrays are data, not code; nothing at run time reads /root/reference.
"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import refprt

here = os.path.dirname(os.path.abspath(__file__))
idx, pos, nrm = refprt.load_obj("/root/reference/data/teapot/teapot.obj", 0.005, (-0.5, 0.0, 0.5))
print("teapot", idx.shape, pos.shape)
np.savez_compressed(os.path.join(here, "teapot.npz"), indices=idx, positions=pos, normals=nrm)
cidx, cpos, cpm, mats = refprt.cornell_box()
print("cornell", cidx.shape, cpos.shape, len(mats))
np.savez_compressed(os.path.join(here, "cornell.npz"), indices=cidx, positions=cpos, prim_material=cpm,
                    diffuse=np.array([m["diffuse"] for m in mats], np.float32),
                    emissive=np.array([m["emissive"] for m in mats], np.float32),
                    refl=np.array([m["refl"] for m in mats], np.uint32))
