"""prt_b200 — B200-native (sm_100a) implementation of amada/PRT's ray-tracing hot path behind a C ABI.

Only what the hot path needs lives here: csrc/ (CUDA kernels, host BVH builder, C ABI -> libprt_cuda.so),
capi.py (ctypes binding of include/prt_cuda.h), scenes.py (scene descriptions for the BASELINE configs).
"""
from . import capi, scenes  # noqa: F401
from .capi import Context, PrtError, load_scene, PRT_MODE_SINGLE, PRT_MODE_PACKET, PRT_JOB_INTERSECT, PRT_JOB_OCCLUDED  # noqa: F401
