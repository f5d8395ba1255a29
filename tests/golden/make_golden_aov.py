"""Generates tests/golden/aov.npz from the UNMODIFIED reference (oracle/_ref/libprt_ref.so): G-buffer images
(GbufferVisualizer::Trace, per-pixel seeding) and the output stage (Image::savePpm payload bytes, Image::saveExr
pixels read back with tinyexr).  Run once where /root/reference exists:  python tests/golden/make_golden_aov.py"""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from prt_b200 import scenes
from oracle import refprt

out = {}
c1 = refprt.RefScene(scenes.cornell_teapot(64, 48, teapot_refl=scenes.DIFFUSE, light="directional"))
c4 = refprt.RefScene(scenes.atrium(64, 48, detail=0.25, tex_size=64))
for name, sc in (("c1", c1), ("c4", c4)):
    for t in (0, 1, 2):
        out[f"gbuffer_{name}_{t}"] = sc.gbuffer(t, 7, exposure=1.5 if t == 0 else 1.0, nthreads=8)
# output stage on a rendered image plus hand-made edge values (zero, one, inf, nan, negative, half overflow / denormal range, ties)
img = c1.render(16, 7, nthreads=8)
edge = np.array([0, 1, np.inf, np.nan, -1, 1e-8, 65504, 65520, 70000, 6e-8, 1e-5, 3e-6, 0.33325195, 1.0009765625, 1.00048828125,
                 0.5, 2.0 ** -14, 2.0 ** -15, 2.0 ** -24, 2.0 ** -25, 3 * 2.0 ** -25, 255.0 / 256, 0.21763764, 1e10, -0.0, 0.7, 0.1, 0.01, 1e-3, 4.0], np.float32)
img[0, :10] = edge.reshape(10, 3)
out["image"] = img
out["ppm_tonemap"] = refprt.image_ppm(img, True)
out["ppm_linear"] = refprt.image_ppm(img, False)
out["exr_rgba"] = refprt.image_exr(img)
np.savez_compressed(os.path.join(HERE, "aov.npz"), **out)
print("aov.npz written:", {k: v.shape for k, v in out.items()})
