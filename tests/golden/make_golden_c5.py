"""Generates tests/golden/render_c5_full.npz from the UNMODIFIED reference (oracle/_ref/libprt_ref.so): two full-width rows of
config C5 at its full size (scenes.city(3840, 2160): 4.9 M triangles), 8 spp, per-packet seeding, plus the closest-hit records
of a seeded set of bounce-like rays.  Run once where /root/reference exists:  python tests/golden/make_golden_c5.py"""
import os, sys, time
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from prt_b200 import scenes
from oracle import refprt

Y0, ROWS, SPP, SEED = 1440, 2, 8, 5
t = time.time()
desc = scenes.city(3840, 2160)
ref = refprt.RefScene(desc)
print("reference scene built in %.1f s" % (time.time() - t))
img = ref.render(SPP, SEED, region=(0, Y0, 3840, Y0 + ROWS), nthreads=8)
np.savez_compressed(os.path.join(HERE, "render_c5_full.npz"), rows=img[Y0:Y0 + ROWS], y0=Y0, spp=SPP, seed=SEED)
print("render_c5_full.npz written", img[Y0:Y0 + ROWS].shape, float(img[Y0:Y0 + ROWS].max()))
