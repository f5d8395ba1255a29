"""ctypes binding of oracle/_ref/libprt_ref.so — the UNMODIFIED reference (amada/PRT) behind
oracle/ref_harness.cpp.  TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.  prt_b200/ never does.

A scene is the plain-python description produced by prt_b200.scenes (dict of numpy arrays); the same
description is uploaded to the CUDA path through the C ABI, so both sides see identical inputs.
"""
import ctypes as C
import os
import sys
import contextlib
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libprt_ref.so")
REF_STATS_SO = os.path.join(_HERE, "_ref", "libprt_ref_stats.so")


class RefMaterial(C.Structure):
    _fields_ = [("diffuse", C.c_float * 3), ("emissive", C.c_float * 3), ("reflType", C.c_uint32),
                ("diffuseTex", C.c_int32), ("bumpTex", C.c_int32), ("alphaTest", C.c_uint32)]


def available(stats=False):
    return os.path.exists(REF_STATS_SO if stats else REF_SO)


@contextlib.contextmanager
def quiet_stdout():
    """The reference prints from ThreadPool::create / Bvh::build (thread_pool.cpp:13, bvh.cpp:209-225);
    keep that off our stdout (bench.py must print exactly one JSON line)."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    try:
        os.dup2(devnull, 1)
        yield
    finally:
        # flush the C stdio buffer of the library before restoring fd 1
        try:
            C.CDLL(None).fflush(None)
        except Exception:
            pass
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


_libs = {}


def _load(stats=False):
    path = REF_STATS_SO if stats else REF_SO
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise RuntimeError(f"{path} missing: run `make -C oracle ref` where /root/reference exists")
    lib = C.CDLL(path)
    lib.ref_create.restype = C.c_void_p
    lib.ref_get_radius.restype = C.c_float
    lib.ref_get_radius.argtypes = [C.c_void_p]
    lib.ref_bvh_node_count.restype = C.c_uint32
    lib.ref_bvh_node_count.argtypes = [C.c_void_p, C.c_int]
    _libs[path] = lib
    return lib


ORACLE_SO = os.path.join(_HERE, "_build", "libprt_oracle.so")


class _Prefixed:
    """lib.<prefix><name> lookup so the reference harness (ref_*) and our restatement (orc_*) share one wrapper."""

    def __init__(self, lib, prefix):
        self._lib, self._prefix = lib, prefix

    def __getattr__(self, name):
        return getattr(self._lib, self._prefix + name)


def _load_oracle():
    if ORACLE_SO in _libs:
        return _libs[ORACLE_SO]
    if not os.path.exists(ORACLE_SO):
        raise RuntimeError(f"{ORACLE_SO} missing: run `make -C oracle oracle`")
    lib = C.CDLL(ORACLE_SO)
    lib.orc_create.restype = C.c_void_p
    lib.orc_get_radius.restype = C.c_float
    lib.orc_get_radius.argtypes = [C.c_void_p]
    lib.orc_bvh_node_count.restype = C.c_uint32
    lib.orc_bvh_node_count.argtypes = [C.c_void_p, C.c_int]
    _libs[ORACLE_SO] = lib
    return lib


class RefScene:
    """Reference Scene + Camera built from a prt_b200.scenes description (kind='ref'), or the same API over
    our scalar restatement oracle/prt_oracle.cpp (kind='oracle' — see OracleScene)."""

    def __init__(self, desc, stats=False, kind="ref"):
        self.kind = kind
        self.lib = _Prefixed(_load(stats), "ref_") if kind == "ref" else _Prefixed(_load_oracle(), "orc_")
        L = self.lib
        with quiet_stdout():
            self.h = C.c_void_p(L.create())
            self.tex_ids = []
            for t in desc.get("textures", []):
                tex = np.ascontiguousarray(t["texels"])
                ch = 1 if tex.dtype == np.float32 else 0
                self.tex_ids.append(L.add_texture(self.h, int(t["width"]), int(t["height"]), int(t["comp"]), ch,
                                                      tex.ctypes.data_as(C.c_void_p)))
            self.alpha_flags = []
            for m in desc["meshes"]:
                idx = _u32(m["indices"]); pos = _f32(m["positions"])
                nrm = _f32(m["normals"]) if m.get("normals") is not None else None
                uv = _f32(m["texcoords"]) if m.get("texcoords") is not None else None
                pm = _u32(m["prim_material"])
                mats = (RefMaterial * len(m["materials"]))()
                for i, mm in enumerate(m["materials"]):
                    mats[i].diffuse[:] = [float(x) for x in mm["diffuse"]]
                    mats[i].emissive[:] = [float(x) for x in mm.get("emissive", (0, 0, 0))]
                    mats[i].reflType = int(mm.get("refl", 0))
                    mats[i].diffuseTex = int(mm.get("diffuse_tex", -1))
                    mats[i].bumpTex = int(mm.get("bump_tex", -1))
                L.add_mesh(self.h, C.c_uint32(idx.size // 3), C.c_uint32(pos.size // 3), _p(idx, C.c_uint32),
                               _p(pos, C.c_float), _p(nrm, C.c_float), _p(uv, C.c_float), _p(pm, C.c_uint32),
                               C.c_uint32(len(mats)), mats)
                self.alpha_flags.append([int(mats[i].alphaTest) for i in range(len(mats))])
            light = desc.get("light")
            if light is not None:
                if light["type"] == "directional":
                    d = _f32(light["dir"]); it = _f32(light["intensity"])
                    L.set_directional_light(self.h, _p(d, C.c_float), _p(it, C.c_float))
                elif light["type"] == "infinite":
                    img = _f32(light["image"])
                    assert img.ndim == 3 and img.shape[2] == 4
                    r = L.set_infinite_light(self.h, img.shape[1], img.shape[0], _p(img, C.c_float))
                    assert r == 0
                    self.light_wh = (img.shape[1], img.shape[0])
            cam = desc.get("camera")
            if cam is not None:
                self.set_camera(cam["pos"], cam["dir"], cam["width"], cam["height"])

    def set_camera(self, pos, dirv, w, h):
        p = _f32(pos); d = _f32(dirv)
        self.lib.set_camera(self.h, _p(p, C.c_float), _p(d, C.c_float), C.c_uint32(w), C.c_uint32(h))
        self.width, self.height = int(w), int(h)

    def camera_basis(self):
        out = np.zeros(14, np.float32)
        self.lib.get_camera(self.h, _p(out, C.c_float))
        return out

    def radius(self):
        return float(self.lib.get_radius(self.h))

    def infinite_light_tables(self):
        w, h = self.light_wh
        v = np.zeros(h, np.float32); hz = np.zeros(w * h, np.float32); img = np.zeros((h, w, 4), np.float32)
        self.lib.get_infinite_light(self.h, _p(v, C.c_float), _p(hz, C.c_float), _p(img, C.c_float))
        return v, hz.reshape(h, w), img

    def light_sample(self, u):
        u = _f32(u); n = u.shape[0]
        d = np.zeros((n, 3), np.float32); c = np.zeros((n, 3), np.float32)
        self.lib.light_sample(self.h, C.c_uint32(n), _p(u, C.c_float), _p(d, C.c_float), _p(c, C.c_float))
        return d, c

    # ---- ray batches: rays (n,8) float32 {org,dir,maxT,pad}; hits (n,6) float32 view, cols 4,5 are u32 bits
    def intersect_single(self, rays, nthreads=1):
        rays = _f32(rays); n = rays.shape[0]
        hits = np.zeros((n, 6), np.float32)
        self.lib.intersect_single(self.h, C.c_uint32(n), _p(rays, C.c_float), _p(hits, C.c_float), C.c_int(nthreads))
        return hits

    def intersect_packet(self, rays, avg_dir, nthreads=1):
        rays = _f32(rays); n = rays.shape[0]; assert n % 8 == 0
        avg = _f32(avg_dir); assert avg.shape[0] == n // 8
        hits = np.zeros((n, 6), np.float32)
        self.lib.intersect_packet(self.h, C.c_uint32(n // 8), _p(rays, C.c_float), _p(avg, C.c_float), _p(hits, C.c_float), C.c_int(nthreads))
        return hits

    def occluded_single(self, rays, nthreads=1):
        rays = _f32(rays); n = rays.shape[0]
        out = np.zeros(n, np.uint8)
        self.lib.occluded_single(self.h, C.c_uint32(n), _p(rays, C.c_float), _p(out, C.c_uint8), C.c_int(nthreads))
        return out

    def occluded_packet(self, rays, mask_bits, nthreads=1):
        """Returns per-ray 0/1 (n,) for active lanes; inactive lanes report 1 like the reference."""
        rays = _f32(rays); n = rays.shape[0]; assert n % 8 == 0
        mb = np.ascontiguousarray(mask_bits, dtype=np.uint8)
        out = np.zeros(n // 8, np.uint8)
        self.lib.occluded_packet(self.h, C.c_uint32(n // 8), _p(rays, C.c_float), _p(mb, C.c_uint8), _p(out, C.c_uint8), C.c_int(nthreads))
        return np.unpackbits(out[:, None], axis=1, bitorder="little").reshape(-1)

    def primary_packets(self, region, ppp, seed, mode=0, nthreads=1, want_rays=True):
        x0, y0, x1, y1 = region
        npk = (x1 - x0) * (y1 - y0) * ppp
        rays = np.zeros((npk * 8, 8), np.float32) if want_rays else None
        avg = np.zeros((npk, 3), np.float32) if want_rays else None
        hits = np.zeros((npk * 8, 6), np.float32) if (mode == 0 and want_rays) else None
        self.lib.primary_packets(self.h, C.c_uint32(x0), C.c_uint32(y0), C.c_uint32(x1), C.c_uint32(y1), C.c_uint32(ppp),
                                     C.c_uint32(seed), C.c_int(mode), _p(rays, C.c_float), _p(avg, C.c_float), _p(hits, C.c_float), C.c_int(nthreads))
        return rays, avg, hits

    def primary_single(self, region, spp, seed, nthreads=1, want_rays=True):
        x0, y0, x1, y1 = region
        n = (x1 - x0) * (y1 - y0) * spp
        rays = np.zeros((n, 8), np.float32) if want_rays else None
        hits = np.zeros((n, 6), np.float32) if want_rays else None
        self.lib.primary_single(self.h, C.c_uint32(x0), C.c_uint32(y0), C.c_uint32(x1), C.c_uint32(y1), C.c_uint32(spp),
                                    C.c_uint32(seed), _p(rays, C.c_float), _p(hits, C.c_float), C.c_int(nthreads))
        return rays, hits

    def render(self, spp, seed, region=None, exposure=1.0, nthreads=1, out=None):
        if region is None:
            region = (0, 0, self.width, self.height)
        x0, y0, x1, y1 = region
        if out is None:
            out = np.zeros((self.height, self.width, 3), np.float32)
        self.lib.render(self.h, C.c_uint32(x0), C.c_uint32(y0), C.c_uint32(x1), C.c_uint32(y1), C.c_uint32(spp), C.c_uint32(seed),
                            C.c_float(exposure), _p(out, C.c_float), C.c_int(nthreads))
        return out

    def render_tiles(self, tiles, spp, seed, exposure=1.0, nthreads=1, out=None):
        """TraceBlock over a list of tiles [(x0,y0,x1,y1) exclusive], one tile per task (the reference's parallel unit,
        main.cpp:121-158); pixels outside the tiles are left alone."""
        t = np.ascontiguousarray(tiles, np.uint32).reshape(-1, 4)
        if out is None:
            out = np.zeros((self.height, self.width, 3), np.float32)
        self.lib.render_tiles(self.h, C.c_uint32(t.shape[0]), _p(t, C.c_uint32), C.c_uint32(spp), C.c_uint32(seed), C.c_float(exposure),
                              _p(out, C.c_float), C.c_int(nthreads))
        return out

    def gbuffer(self, gtype, seed, region=None, exposure=1.0, nthreads=1, out=None):
        """GbufferVisualizer::TraceBlock (gbuffer_visualizer.cpp:17-51); gtype 0 diffuse, 1 mesh normal, 2 normal."""
        if region is None:
            region = (0, 0, self.width, self.height)
        x0, y0, x1, y1 = region
        if out is None:
            out = np.zeros((self.height, self.width, 3), np.float32)
        self.lib.gbuffer(self.h, C.c_uint32(x0), C.c_uint32(y0), C.c_uint32(x1), C.c_uint32(y1), C.c_int(gtype), C.c_uint32(seed),
                         C.c_float(exposure), _p(out, C.c_float), C.c_int(nthreads))
        return out

    def surface_properties(self, hits):
        """Scene::getSurfaceProperties (scene.cpp:30-45) over (n,6) hit records -> (n,16) float32 in prt_surface order
        (normal, material index as u32 bits, uv, duv01, duv02, dp01, dp02); misses give zeros."""
        hits = _f32(hits); n = hits.shape[0]
        out = np.zeros((n, 16), np.float32)
        self.lib.surface_properties(self.h, C.c_uint32(n), _p(hits, C.c_float), _p(out, C.c_float))
        return out

    def bvh_dump(self, mesh, nprim):
        n = int(self.lib.bvh_node_count(self.h, mesh))
        nodes = np.zeros((n, 8), np.uint32); dec = np.zeros((n, 4), np.uint32); remap = np.zeros(nprim, np.uint32)
        self.lib.bvh_dump(self.h, mesh, _p(nodes, C.c_uint32), _p(dec, C.c_uint32), _p(remap, C.c_uint32))
        return nodes, dec, remap


class OracleScene(RefScene):
    def __init__(self, desc):
        super().__init__(desc, kind="oracle")


def hits_ids(hits):
    """(n,6) float32 hit records -> (t,i,j,k float32 (n,4), primId uint32, meshId uint32)."""
    u = hits.view(np.uint32)
    return hits[:, :4], u[:, 4].copy(), u[:, 5].copy()


def intersect_triangle_soa(rays8, tri9, kind="ref"):
    lib = _Prefixed(_load(), "ref_") if kind == "ref" else _Prefixed(_load_oracle(), "orc_")
    r = _f32(rays8); t = _f32(tri9); out = np.zeros((8, 4), np.float32)
    lib.intersect_triangle_soa(_p(r, C.c_float), _p(t, C.c_float), _p(out, C.c_float))
    return out


def load_obj(path, scale, translate):
    lib = _load()
    tr = _f32(translate); np_ = C.c_uint32(); nv = C.c_uint32()
    with quiet_stdout():
        r = lib.ref_load_obj(path.encode(), C.c_float(scale), _p(tr, C.c_float), C.byref(np_), C.byref(nv), None, None, None)
        assert r == 0, path
        idx = np.zeros((np_.value, 3), np.uint32); pos = np.zeros((nv.value, 3), np.float32); nrm = np.zeros((nv.value, 3), np.float32)
        lib.ref_load_obj(path.encode(), C.c_float(scale), _p(tr, C.c_float), C.byref(np_), C.byref(nv), _p(idx, C.c_uint32), _p(pos, C.c_float), _p(nrm, C.c_float))
    return idx, pos, nrm


def calc_vertex_normals(idx, pos, kind="ref"):
    lib = _Prefixed(_load(), "ref_") if kind == "ref" else _Prefixed(_load_oracle(), "orc_")
    idx = _u32(idx); pos = _f32(pos); out = np.zeros_like(pos)
    lib.calc_vertex_normals(C.c_uint32(idx.size // 3), C.c_uint32(pos.size // 3), _p(idx, C.c_uint32), _p(pos, C.c_float), _p(out, C.c_float))
    return out


def cornell_box():
    lib = _load()
    counts = np.zeros(3, np.uint32)
    lib.ref_cornell_box(_p(counts, C.c_uint32), None, None, None, None)
    nprim, nvert, nmat = [int(x) for x in counts]
    idx = np.zeros((nprim, 3), np.uint32); pos = np.zeros((nvert, 3), np.float32); pm = np.zeros(nprim, np.uint32)
    mats = (RefMaterial * nmat)()
    lib.ref_cornell_box(_p(counts, C.c_uint32), _p(idx, C.c_uint32), _p(pos, C.c_float), _p(pm, C.c_uint32), mats)
    materials = [dict(diffuse=tuple(m.diffuse), emissive=tuple(m.emissive), refl=int(m.reflType)) for m in mats]
    return idx, pos, pm, materials


def ref_stats(reset=False):
    """(rays, nodesTraversed, primsTraversed) accumulated by libprt_ref_stats.so closest-hit calls."""
    lib = _load(True)
    out = np.zeros(3, np.uint64)
    lib.ref_stats_get(out.ctypes.data_as(C.POINTER(C.c_uint64)))
    if reset:
        lib.ref_stats_reset()
    return [int(x) for x in out]


def oracle_stats(enable=None, reset=False):
    lib = _load_oracle()
    if enable is not None:
        lib.orc_stats_enable(1 if enable else 0)
    out = np.zeros(12, np.uint64)
    lib.orc_stats_get(out.ctypes.data_as(C.POINTER(C.c_uint64)))
    if reset:
        lib.orc_stats_reset()
    keys = ["single_nodes", "single_prims", "packet_nodes", "packet_prims", "occl_nodes", "occl_prims", "single_rays", "packet_rays", "occl_rays"]
    return {k: int(out[i]) for i, k in enumerate(keys)}


def image_ppm(rgb, tonemap=True, kind="ref"):
    """Payload bytes of Image::savePpm (image.cpp:52-80) for an RGB32F image (H, W, 3) -> uint8 (H, W, 3)."""
    import tempfile
    lib = _Prefixed(_load(), "ref_") if kind == "ref" else _Prefixed(_load_oracle(), "orc_")
    rgb = _f32(rgb); h, w = rgb.shape[:2]
    out = np.zeros((h, w, 3), np.uint8)
    path = os.path.join(tempfile.gettempdir(), f"prt_ref_{os.getpid()}.ppm").encode()
    with quiet_stdout():
        r = lib.image_ppm(C.c_uint32(w), C.c_uint32(h), C.c_int(1 if tonemap else 0), _p(rgb, C.c_float), _p(out, C.c_uint8), path)
    assert r == 0, r
    return out


def image_exr(rgb, kind="ref"):
    """Pixels of Image::saveExr (image.cpp:82-139) read back: (H, W, 4) float32 holding half-precision values."""
    import tempfile
    lib = _Prefixed(_load(), "ref_") if kind == "ref" else _Prefixed(_load_oracle(), "orc_")
    rgb = _f32(rgb); h, w = rgb.shape[:2]
    out = np.zeros((h, w, 4), np.float32)
    path = os.path.join(tempfile.gettempdir(), f"prt_ref_{os.getpid()}.exr").encode()
    with quiet_stdout():
        r = lib.image_exr(C.c_uint32(w), C.c_uint32(h), _p(rgb, C.c_float), _p(out, C.c_float), path)
    assert r == 0, r
    return out


def half_bits(x):
    """tinyexr float_to_half_full (tinyexr.h:899-934) over an array, from the oracle restatement."""
    x = _f32(x); out = np.zeros(x.shape, np.uint16)
    _load_oracle().orc_half_bits(C.c_uint32(x.size), _p(x, C.c_float), _p(out, C.c_uint16))
    return out
