"""ctypes binding of the C ABI in include/prt_cuda.h (prt_b200/libprt_cuda.so).

This is the only way Python reaches the hot path: there is no Python or CPU implementation behind it.  If the
shared library is missing, importing this module raises; if no CUDA device is usable, every compute call
raises PrtError — nothing falls back to the CPU.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libprt_cuda.so")
STATS_LIB_PATH = os.path.join(_HERE, "libprt_cuda_stats.so")   # same sources, -DPRT_STATS=1 (make -C prt_b200/csrc stats)

PRT_MODE_SINGLE, PRT_MODE_PACKET = 0, 1
PRT_JOB_INTERSECT, PRT_JOB_OCCLUDED = 0, 1
PRT_RAYS_FULL, PRT_RAYS_SHARED_ORIGIN, PRT_RAYS_SHARED_DIR, PRT_RAYS_PRIMARY, PRT_RAYS_FROM_HITS = 0, 1, 2, 3, 4
PRT_OPTION_BUILD_FAST_TRAVERSAL, PRT_OPTION_TRAVERSAL = 1, 2
PRT_TRAVERSAL_EXACT, PRT_TRAVERSAL_FAST = 0, 1
PRT_COMM_ID_BYTES = 128
PRT_NOALLOC_MAX_RAYS = 262144   # prt_cuda_primary_shadow_noalloc: rays its static scratch holds
RAY_CLASSES = ("closest_single", "closest_packet", "occlude")


class PrtError(RuntimeError):
    pass


class prt_material(C.Structure):
    _fields_ = [("diffuse", C.c_float * 3), ("emissive", C.c_float * 3), ("reflType", C.c_uint32),
                ("diffuseTex", C.c_int32), ("bumpTex", C.c_int32), ("alphaTest", C.c_uint32)]


class prt_mesh_desc(C.Structure):
    _fields_ = [("primCount", C.c_uint32), ("vertexCount", C.c_uint32), ("materialCount", C.c_uint32),
                ("indices", C.POINTER(C.c_uint32)), ("positions", C.POINTER(C.c_float)), ("normals", C.POINTER(C.c_float)),
                ("texcoords", C.POINTER(C.c_float)), ("primMaterial", C.POINTER(C.c_uint32)), ("materials", C.POINTER(prt_material))]


class prt_trace_job(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("n", C.c_uint32), ("mode", C.c_int32), ("layout", C.c_uint32),
                ("rays", C.c_void_p), ("avgDir", C.c_void_p), ("out", C.c_void_p),
                ("shared", C.c_float * 4), ("region", C.c_uint32 * 4), ("perPixel", C.c_uint32), ("seed", C.c_uint32),
                ("fromJob", C.c_int32), ("reserved", C.c_uint32)]


class prt_counters(C.Structure):
    _fields_ = [("extendRays", C.c_uint64), ("shadowRays", C.c_uint64), ("primaryRays", C.c_uint64),
                ("kernelLaunches", C.c_uint64), ("lastKernelMs", C.c_float), ("shadowRaysSkipped", C.c_uint64),
                ("lastGatherMs", C.c_float),
                ("statRays", C.c_uint64 * 3), ("statNodeSteps", C.c_uint64 * 3), ("statLeafVisits", C.c_uint64 * 3),
                ("statTriTests", C.c_uint64 * 3), ("statUvLoads", C.c_uint64 * 3)]


# prt_surface as a numpy record (64 bytes, SurfaceProperties field order, mesh.h:19-28)
SURFACE_DTYPE = np.dtype([("normal", np.float32, 3), ("material", np.uint32), ("uv", np.float32, 2), ("duv01", np.float32, 2),
                          ("duv02", np.float32, 2), ("dp01", np.float32, 3), ("dp02", np.float32, 3)])
assert SURFACE_DTYPE.itemsize == 64


# every symbol include/prt_cuda.h declares: name -> (restype, argtypes)
_VP, _U32, _I, _F = C.c_void_p, C.c_uint32, C.c_int, C.c_float
SYMBOLS = {
    "prt_cuda_create": (_I, [_I, C.POINTER(_VP)]),
    "prt_cuda_destroy": (None, [_VP]),
    "prt_cuda_last_error": (C.c_char_p, [_VP]),
    "prt_cuda_device_count": (_I, []),
    "prt_cuda_create_multi": (_I, [C.POINTER(_I), _I, C.POINTER(_VP)]),
    "prt_cuda_comm_unique_id": (_I, [_VP]),
    "prt_cuda_comm_init_rank": (_I, [_VP, _VP, _I, _I]),
    "prt_cuda_comm_info": (_I, [_VP, C.POINTER(_I), C.POINTER(_I), C.POINTER(_I)]),
    "prt_cuda_host_alloc": (_I, [C.c_size_t, C.POINTER(_VP)]),
    "prt_cuda_host_free": (_I, [_VP]),
    "prt_cuda_host_register": (_I, [_VP, C.c_size_t]),
    "prt_cuda_host_unregister": (_I, [_VP]),
    "prt_cuda_surface_properties": (_I, [_VP, _U32, _VP, _VP]),
    "prt_cuda_surface_properties_dev": (_I, [_VP, _U32, _VP, _VP]),
    "prt_cuda_gather_bands_dev": (_I, [_VP, _U32, _U32, _U32, _U32, _VP]),
    "prt_cuda_set_option": (_I, [_VP, _I, _I]),
    "prt_cuda_get_option": (_I, [_VP, _I, C.POINTER(_I)]),
    "prt_cuda_scene_add_texture": (_I, [_VP, _I, _I, _I, _I, _VP]),
    "prt_cuda_scene_add_mesh": (_I, [_VP, C.POINTER(prt_mesh_desc)]),
    "prt_cuda_scene_set_directional_light": (_I, [_VP, C.POINTER(_F), C.POINTER(_F)]),
    "prt_cuda_scene_set_infinite_light": (_I, [_VP, _I, _I, C.POINTER(_F)]),
    "prt_cuda_scene_get_radius": (_I, [_VP, C.POINTER(_F)]),
    "prt_cuda_set_camera": (_I, [_VP, C.POINTER(_F), C.POINTER(_F), _U32, _U32]),
    "prt_cuda_bvh_node_count": (_I, [_VP, _I, C.POINTER(_U32)]),
    "prt_cuda_bvh_get": (_I, [_VP, _I, C.POINTER(_U32), C.POINTER(_U32)]),
    "prt_cuda_light_get_tables": (_I, [_VP, C.POINTER(_F), C.POINTER(_F)]),
    "prt_cuda_get_camera": (_I, [_VP, C.POINTER(_F)]),
    "prt_cuda_intersect": (_I, [_VP, _U32, _VP, _VP, _I, _VP]),
    "prt_cuda_intersect_dev": (_I, [_VP, _U32, _VP, _VP, _I, _VP]),
    "prt_cuda_occluded": (_I, [_VP, _U32, _VP, _I, _VP]),
    "prt_cuda_occluded_dev": (_I, [_VP, _U32, _VP, _I, _VP]),
    "prt_cuda_trace_jobs": (_I, [_VP, _U32, C.POINTER(prt_trace_job)]),
    "prt_cuda_primary_dev": (_I, [_VP, _U32, _U32, _U32, _U32, _U32, _U32, _I, _VP, _VP, _VP]),
    "prt_cuda_bvh_get_fast": (_I, [_VP, _I, C.POINTER(_U32), C.POINTER(_U32), _VP, _VP]),
    "prt_cuda_primary_shadow_noalloc": (_I, [_VP, _U32, _U32, _U32, _U32, _U32, _U32, C.POINTER(C.c_float), _VP]),
    "prt_cuda_render": (_I, [_VP, _U32, _U32, _U32, _U32, _U32, _U32, _F, _VP]),
    "prt_cuda_render_dev": (_I, [_VP, _U32, _U32, _U32, _U32, _U32, _U32, _F, _VP]),
    "prt_cuda_render_bands_dev": (_I, [_VP, _U32, _U32, _U32, _U32, _U32, _F, _VP]),
    "prt_cuda_gbuffer": (_I, [_VP, _U32, _U32, _U32, _U32, _I, _U32, _F, _VP]),
    "prt_cuda_gbuffer_dev": (_I, [_VP, _U32, _U32, _U32, _U32, _I, _U32, _F, _VP]),
    "prt_cuda_image_to_ppm": (_I, [_VP, _VP, _U32, _U32, _I, _VP]),
    "prt_cuda_image_to_ppm_dev": (_I, [_VP, _VP, _U32, _U32, _I, _VP]),
    "prt_cuda_image_to_half_bgr": (_I, [_VP, _VP, _U32, _U32, _VP]),
    "prt_cuda_image_to_half_bgr_dev": (_I, [_VP, _VP, _U32, _U32, _VP]),
    "prt_cuda_synchronize": (_I, [_VP]),
    "prt_cuda_get_counters": (_I, [_VP, C.POINTER(prt_counters)]),
    "prt_cuda_reset_counters": (_I, [_VP]),
    "prt_cuda_stream": (C.c_uint64, [_VP]),
    "prt_cuda_probe_peaks": (C.c_int, [_VP, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "prt_packet_seed": (_U32, [_U32, _U32, _U32, _U32]),
}

_libs = {}


def lib(stats=False):
    """The product library, or (stats=True) its counting build: same sources with -DPRT_STATS=1, whose
    prt_cuda_get_counters also reports node steps / triangle tests / uv loads per ray class."""
    path = STATS_LIB_PATH if stats else LIB_PATH
    if path not in _libs:
        if not os.path.exists(path):
            raise PrtError(f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C prt_b200/csrc [stats]).  There is no CPU fallback.")
        l = C.CDLL(path)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)      # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _libs[path] = l
    return _libs[path]


_nccl_preloaded = False


def _prefer_bundled_nccl():
    """libprt_cuda.so binds NCCL at its first communicator call and reuses a libnccl.so.2 the process already holds
    (multi_gpu.cu).  In a Python process that may import torch later, make that copy the one PyTorch bundles — the system
    library has the same soname but is older, and torch would fail to import on top of it."""
    global _nccl_preloaded
    if _nccl_preloaded:
        return
    _nccl_preloaded = True
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for loc in (spec.submodule_search_locations if spec else []):
            path = os.path.join(loc, "lib", "libnccl.so.2")
            if os.path.exists(path):
                C.CDLL(path, mode=C.RTLD_GLOBAL)
                return
    except Exception:
        pass


def comm_unique_id():
    """PRT_COMM_ID_BYTES bytes from ncclGetUniqueId (rank 0 calls this and broadcasts the bytes)."""
    _prefer_bundled_nccl()
    buf = (C.c_uint8 * PRT_COMM_ID_BYTES)()
    if lib().prt_cuda_comm_unique_id(buf) != 0:
        raise PrtError(lib().prt_cuda_last_error(None).decode())
    return bytes(buf)


class PinnedBuffer:
    """Page-locked host memory from prt_cuda_host_alloc, viewed as a numpy array."""

    def __init__(self, shape, dtype):
        self.L = lib()
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        if self.L.prt_cuda_host_alloc(max(self.nbytes, 1), C.byref(p)) != 0:
            raise PrtError(self.L.prt_cuda_last_error(None).decode())
        self.ptr = p.value
        self.array = np.frombuffer((C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr), dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            self.L.prt_cuda_host_free(self.ptr)
            self.ptr = None


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32)) if a is not None else None


class Context:
    """One prt_ctx (one GPU).  device=-1 gives a host-only context: scene building (BVH, light tables, camera)
    works, every compute entry point raises."""

    def __init__(self, device=0, stats=False):
        """device: one CUDA device index, -1 (host-only), or a list of indices (one context driving several GPUs,
        prt_cuda_create_multi)."""
        self.L = lib(stats)
        h = C.c_void_p()
        if isinstance(device, (list, tuple)):
            _prefer_bundled_nccl()
            devs = (C.c_int * len(device))(*[int(d) for d in device])
            r = self.L.prt_cuda_create_multi(devs, len(device), C.byref(h))
        else:
            r = self.L.prt_cuda_create(int(device), C.byref(h))
        if r != 0:
            raise PrtError(self.L.prt_cuda_last_error(None).decode())
        self.h = h
        self.device = device

    def set_option(self, option, value):
        self.check(self.L.prt_cuda_set_option(self.h, int(option), int(value)))

    def get_option(self, option):
        v = C.c_int()
        self.check(self.L.prt_cuda_get_option(self.h, int(option), C.byref(v)))
        return v.value

    def set_traversal(self, fast):
        """Switches single-ray-rule rays between the exact structure (default) and the opt-in fast mode (needs load_scene(..., fast=True))."""
        self.set_option(PRT_OPTION_TRAVERSAL, PRT_TRAVERSAL_FAST if fast else PRT_TRAVERSAL_EXACT)

    def comm_init_rank(self, unique_id, rank, world):
        _prefer_bundled_nccl()
        buf = (C.c_uint8 * PRT_COMM_ID_BYTES).from_buffer_copy(unique_id)
        self.check(self.L.prt_cuda_comm_init_rank(self.h, buf, int(rank), int(world)))

    def comm_info(self):
        r, w, n = C.c_int(), C.c_int(), C.c_int()
        self.check(self.L.prt_cuda_comm_info(self.h, C.byref(r), C.byref(w), C.byref(n)))
        return r.value, w.value, n.value

    def close(self):
        if getattr(self, "h", None):
            self.L.prt_cuda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, r):
        if r < 0:
            raise PrtError(f"prt_cuda error {r}: {self.L.prt_cuda_last_error(self.h).decode()}")
        return r

    # ---- scene ----
    def add_texture(self, width, height, comp, texels):
        t = np.ascontiguousarray(texels)
        ch = 1 if t.dtype == np.float32 else 0
        return self.check(self.L.prt_cuda_scene_add_texture(self.h, int(width), int(height), int(comp), ch, t.ctypes.data_as(C.c_void_p)))

    def add_mesh(self, m):
        idx = np.ascontiguousarray(m["indices"], np.uint32)
        pos = np.ascontiguousarray(m["positions"], np.float32)
        nrm = np.ascontiguousarray(m["normals"], np.float32) if m.get("normals") is not None else None
        uv = np.ascontiguousarray(m["texcoords"], np.float32) if m.get("texcoords") is not None else None
        pm = np.ascontiguousarray(m["prim_material"], np.uint32)
        mats = (prt_material * len(m["materials"]))()
        for i, mm in enumerate(m["materials"]):
            mats[i].diffuse[:] = [float(x) for x in mm["diffuse"]]
            mats[i].emissive[:] = [float(x) for x in mm.get("emissive", (0, 0, 0))]
            mats[i].reflType = int(mm.get("refl", 0))
            mats[i].diffuseTex = int(mm.get("diffuse_tex", -1))
            mats[i].bumpTex = int(mm.get("bump_tex", -1))
        d = prt_mesh_desc(idx.size // 3, pos.size // 3, len(mats), _up(idx), _fp(pos), _fp(nrm), _fp(uv), _up(pm), mats)
        mid = self.check(self.L.prt_cuda_scene_add_mesh(self.h, C.byref(d)))
        return mid, [int(mats[i].alphaTest) for i in range(len(mats))]

    def set_directional_light(self, dirv, intensity):
        d = np.ascontiguousarray(dirv, np.float32); it = np.ascontiguousarray(intensity, np.float32)
        self.check(self.L.prt_cuda_scene_set_directional_light(self.h, _fp(d), _fp(it)))

    def set_infinite_light(self, image):
        img = np.ascontiguousarray(image, np.float32)
        assert img.ndim == 3 and img.shape[2] == 4
        self.check(self.L.prt_cuda_scene_set_infinite_light(self.h, img.shape[1], img.shape[0], _fp(img)))
        self._light_wh = (img.shape[1], img.shape[0])

    def set_camera(self, pos, dirv, width, height):
        p = np.ascontiguousarray(pos, np.float32); d = np.ascontiguousarray(dirv, np.float32)
        self.check(self.L.prt_cuda_set_camera(self.h, _fp(p), _fp(d), int(width), int(height)))
        self.width, self.height = int(width), int(height)

    def radius(self):
        r = C.c_float()
        self.check(self.L.prt_cuda_scene_get_radius(self.h, C.byref(r)))
        return r.value

    def camera_basis(self):
        out = np.zeros(14, np.float32)
        self.check(self.L.prt_cuda_get_camera(self.h, _fp(out)))
        return out

    def bvh(self, mesh, nprim):
        n = C.c_uint32()
        self.check(self.L.prt_cuda_bvh_node_count(self.h, mesh, C.byref(n)))
        nodes = np.zeros((n.value, 8), np.uint32); remap = np.zeros(nprim, np.uint32)
        self.check(self.L.prt_cuda_bvh_get(self.h, mesh, _up(nodes), _up(remap)))
        return nodes, remap

    def bvh_fast(self, mesh, nprim):
        """-> (rootRef, Wide8 records as (n, 24) uint32, fast-order triangle records as (nprim, 12) float32)"""
        n, root = C.c_uint32(), C.c_uint32()
        r = self.L.prt_cuda_bvh_get_fast(self.h, mesh, C.byref(n), C.byref(root), None, None)
        if r != 0:
            raise PrtError(f"prt_cuda_bvh_get_fast: status {r} (mesh added without PRT_OPTION_BUILD_FAST_TRAVERSAL?)")
        nodes = np.zeros((n.value, 24), np.uint32); tris = np.zeros((nprim, 12), np.float32)
        r = self.L.prt_cuda_bvh_get_fast(self.h, mesh, None, None, nodes.ctypes.data, tris.ctypes.data)
        if r != 0:
            raise PrtError(f"prt_cuda_bvh_get_fast: status {r} (device contexts drop the host copy after the upload)")
        return root.value, nodes, tris

    def light_tables(self):
        w, h = self._light_wh
        v = np.zeros(h, np.float32); hz = np.zeros((h, w), np.float32)
        self.check(self.L.prt_cuda_light_get_tables(self.h, _fp(v), _fp(hz)))
        return v, hz

    # ---- host-buffer hot-path calls (the e2e path: H2D + kernel + D2H inside the call) ----
    def intersect(self, rays, mode, avg_dir=None):
        rays = np.ascontiguousarray(rays, np.float32); n = rays.shape[0]
        hits = np.zeros((n, 6), np.float32)
        avg = np.ascontiguousarray(avg_dir, np.float32) if avg_dir is not None else None
        self.check(self.L.prt_cuda_intersect(self.h, n, rays.ctypes.data_as(C.c_void_p), avg.ctypes.data_as(C.c_void_p) if avg is not None else None,
                                             int(mode), hits.ctypes.data_as(C.c_void_p)))
        return hits

    def occluded(self, rays, mode):
        rays = np.ascontiguousarray(rays, np.float32); n = rays.shape[0]
        out = np.zeros(n, np.uint8)
        self.check(self.L.prt_cuda_occluded(self.h, n, rays.ctypes.data_as(C.c_void_p), int(mode), out.ctypes.data_as(C.c_void_p)))
        return out

    def trace_jobs(self, jobs):
        """One prt_cuda_trace_jobs call.  jobs: list of dicts with kind, mode and ONE ray source:
             rays=(n,8) float32                              PRT_RAYS_FULL  (+ avg_dir for packet-mode closest hit)
             dirs=(n,3), origin=(3,), max_t                  PRT_RAYS_SHARED_ORIGIN  (+ avg_dir)
             origins=(n,3), direction=(3,), max_t            PRT_RAYS_SHARED_DIR
             primary=(region, per_pixel, seed)               PRT_RAYS_PRIMARY (closest hit)
             from_job=j, light_dir=(3,)                      PRT_RAYS_FROM_HITS (any hit)
           (the round-1 tuple form (kind, rays, mode, avg_dir) is accepted too); returns the list of result arrays."""
        arr = (prt_trace_job * len(jobs))()
        keep, outs, counts = [], [], []
        for i, job in enumerate(jobs):
            if isinstance(job, (tuple, list)):
                kind, rays, mode, avg = job
                job = dict(kind=kind, mode=mode, rays=rays, avg_dir=avg)
            kind, mode = int(job["kind"]), int(job["mode"])
            avg = np.ascontiguousarray(job["avg_dir"], np.float32) if job.get("avg_dir") is not None else None
            J = arr[i]
            J.kind, J.mode, J.fromJob = kind, mode, -1
            data = None
            if job.get("rays") is not None:
                data = np.ascontiguousarray(job["rays"], np.float32); n = data.shape[0]; J.layout = PRT_RAYS_FULL
            elif job.get("dirs") is not None:
                data = np.ascontiguousarray(job["dirs"], np.float32); n = data.shape[0]; J.layout = PRT_RAYS_SHARED_ORIGIN
                J.shared[:] = [float(x) for x in job["origin"]] + [float(job["max_t"])]
            elif job.get("origins") is not None:
                data = np.ascontiguousarray(job["origins"], np.float32); n = data.shape[0]; J.layout = PRT_RAYS_SHARED_DIR
                J.shared[:] = [float(x) for x in job["direction"]] + [float(job["max_t"])]
            elif job.get("primary") is not None:
                (x0, y0, x1, y1), pp, seed = job["primary"]
                J.layout = PRT_RAYS_PRIMARY
                J.region[:] = [x0, y0, x1, y1]; J.perPixel = int(pp); J.seed = int(seed)
                n = (x1 - x0) * (y1 - y0) * int(pp) * (8 if mode == PRT_MODE_PACKET else 1)
            else:
                J.layout = PRT_RAYS_FROM_HITS; J.fromJob = int(job["from_job"])
                n = counts[J.fromJob] if 0 <= J.fromJob < len(counts) else 8      # a bad index is the library's to reject
                J.shared[:] = [float(x) for x in job["light_dir"]] + [0.0]
            out = np.zeros((n, 6), np.float32) if kind == PRT_JOB_INTERSECT else np.zeros(n, np.uint8)
            J.n = n
            J.rays = data.ctypes.data if data is not None and n else None
            J.avgDir = avg.ctypes.data if avg is not None and n else None
            J.out = out.ctypes.data if n else None
            keep += [data, avg]; outs.append(out); counts.append(n)
        self.check(self.L.prt_cuda_trace_jobs(self.h, len(jobs), arr))
        return outs

    def surface_properties(self, hits):
        """Scene::getSurfaceProperties over (n,6) hit records -> structured array of prt_surface (SURFACE_DTYPE)."""
        hits = np.ascontiguousarray(hits, np.float32); n = hits.shape[0]
        out = np.zeros(n, SURFACE_DTYPE)
        self.check(self.L.prt_cuda_surface_properties(self.h, n, hits.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def surface_properties_dev(self, n, d_hits, d_out):
        self.check(self.L.prt_cuda_surface_properties_dev(self.h, int(n), d_hits, d_out))

    def gather_bands_dev(self, region, d_rgb):
        x0, y0, x1, y1 = region
        self.check(self.L.prt_cuda_gather_bands_dev(self.h, x0, y0, x1, y1, d_rgb))

    def render(self, spp, seed, region=None, exposure=1.0, out=None):
        if region is None:
            region = (0, 0, self.width, self.height)
        x0, y0, x1, y1 = region
        if out is None:
            out = np.zeros((self.height, self.width, 3), np.float32)
        self.check(self.L.prt_cuda_render(self.h, x0, y0, x1, y1, int(spp), int(seed), C.c_float(exposure), out.ctypes.data_as(C.c_void_p)))
        return out

    def gbuffer(self, gtype, seed, region=None, exposure=1.0, out=None):
        if region is None:
            region = (0, 0, self.width, self.height)
        x0, y0, x1, y1 = region
        if out is None:
            out = np.zeros((self.height, self.width, 3), np.float32)
        self.check(self.L.prt_cuda_gbuffer(self.h, x0, y0, x1, y1, int(gtype), int(seed), C.c_float(exposure), out.ctypes.data_as(C.c_void_p)))
        return out

    def image_to_ppm(self, rgb, tonemap=True):
        rgb = np.ascontiguousarray(rgb, np.float32); h, w = rgb.shape[:2]
        out = np.zeros((h, w, 3), np.uint8)
        self.check(self.L.prt_cuda_image_to_ppm(self.h, rgb.ctypes.data_as(C.c_void_p), w, h, 1 if tonemap else 0, out.ctypes.data_as(C.c_void_p)))
        return out

    def image_to_half_bgr(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.float32); h, w = rgb.shape[:2]
        out = np.zeros((3, h, w), np.uint16)
        self.check(self.L.prt_cuda_image_to_half_bgr(self.h, rgb.ctypes.data_as(C.c_void_p), w, h, out.ctypes.data_as(C.c_void_p)))
        return out

    # ---- device-pointer calls (ints are raw device addresses, e.g. torch.Tensor.data_ptr()) ----
    def intersect_dev(self, n, d_rays, d_avg, mode, d_hits):
        self.check(self.L.prt_cuda_intersect_dev(self.h, int(n), d_rays, d_avg, int(mode), d_hits))

    def occluded_dev(self, n, d_rays, mode, d_out):
        self.check(self.L.prt_cuda_occluded_dev(self.h, int(n), d_rays, int(mode), d_out))

    def primary_dev(self, region, ppp, seed, mode, d_rays=None, d_avg=None, d_hits=None):
        x0, y0, x1, y1 = region
        self.check(self.L.prt_cuda_primary_dev(self.h, x0, y0, x1, y1, int(ppp), int(seed), int(mode), d_rays, d_avg, d_hits))

    def primary_shadow_noalloc(self, region, per_pixel, seed, light_dir, d_out):
        """prt_cuda_primary_shadow_noalloc: no device allocation inside the call; d_out is a device pointer (1 byte per ray)."""
        x0, y0, x1, y1 = region
        L3 = (C.c_float * 3)(*[float(v) for v in light_dir])
        self.check(self.L.prt_cuda_primary_shadow_noalloc(self.h, x0, y0, x1, y1, int(per_pixel), int(seed), L3, d_out))

    def render_dev(self, region, spp, seed, exposure, d_rgb):
        x0, y0, x1, y1 = region
        self.check(self.L.prt_cuda_render_dev(self.h, x0, y0, x1, y1, int(spp), int(seed), C.c_float(exposure), d_rgb))

    def render_bands_dev(self, band_rows, rank, world, spp, seed, exposure, d_local):
        self.check(self.L.prt_cuda_render_bands_dev(self.h, int(band_rows), int(rank), int(world), int(spp), int(seed), C.c_float(exposure), d_local))

    def synchronize(self):
        self.check(self.L.prt_cuda_synchronize(self.h))

    def counters(self):
        c = prt_counters()
        self.check(self.L.prt_cuda_get_counters(self.h, C.byref(c)))
        d = dict(extendRays=c.extendRays, shadowRays=c.shadowRays, primaryRays=c.primaryRays,
                 kernelLaunches=c.kernelLaunches, lastKernelMs=c.lastKernelMs, shadowRaysSkipped=c.shadowRaysSkipped,
                 lastGatherMs=c.lastGatherMs)
        d["stats"] = {name: dict(rays=int(c.statRays[k]), nodeSteps=int(c.statNodeSteps[k]), leafVisits=int(c.statLeafVisits[k]),
                                 triTests=int(c.statTriTests[k]), uvLoads=int(c.statUvLoads[k])) for k, name in enumerate(RAY_CLASSES)}
        return d

    def reset_counters(self):
        self.check(self.L.prt_cuda_reset_counters(self.h))

    def stream(self):
        return int(self.L.prt_cuda_stream(self.h))

    def probe_peaks(self):
        """(L2 read GB/s, FP32 warp-instruction issue rate in 1e9/s) measured on this context's device."""
        a, b = C.c_double(0.0), C.c_double(0.0)
        self.check(self.L.prt_cuda_probe_peaks(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value


def load_scene(desc, device=0, stats=False, fast=False):
    """Uploads a prt_b200.scenes description; returns the Context (mesh ids == list order, scene.cpp:21).  fast=True also builds
    the structure of the opt-in fast traversal mode (the context still starts in exact mode: ctx.set_traversal(True) switches)."""
    ctx = Context(device, stats=stats)
    if fast:
        ctx.set_option(PRT_OPTION_BUILD_FAST_TRAVERSAL, 1)
    for t in desc.get("textures", []):
        ctx.add_texture(t["width"], t["height"], t["comp"], t["texels"])
    ctx.alpha_flags = []
    for m in desc["meshes"]:
        _, flags = ctx.add_mesh(m)
        ctx.alpha_flags.append(flags)
    light = desc.get("light")
    if light is not None:
        if light["type"] == "directional":
            ctx.set_directional_light(light["dir"], light["intensity"])
        elif light["type"] == "infinite":
            ctx.set_infinite_light(light["image"])
    cam = desc.get("camera")
    if cam is not None:
        ctx.set_camera(cam["pos"], cam["dir"], cam["width"], cam["height"])
    return ctx
