"""The C++ host API (prt_b200/host/prt_host.h: Scene/Mesh/Bvh/Camera/PathTracer/Image with the reference's names) over
the C ABI: compiles with g++, builds a scene in a host-only context on CPU, and — on a GPU — renders through
PathTracer::TraceBlock (whole frame and 16x16 tiles) bit-identically to the C-ABI render of the same scene."""
import os
import struct
import subprocess
import numpy as np
import pytest

import prt_b200
from prt_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "prt_b200", "host")
DEMO = os.path.join(HOST, "build", "demo")


def build_demo():
    os.makedirs(os.path.dirname(DEMO), exist_ok=True)
    src = os.path.join(HOST, "demo.cpp")
    if not os.path.exists(DEMO) or os.path.getmtime(DEMO) < max(os.path.getmtime(src), os.path.getmtime(os.path.join(HOST, "prt_host.h"))):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", src, "-o", DEMO, "-L" + os.path.join(ROOT, "prt_b200"),
                               "-lprt_cuda", "-Wl,-rpath," + os.path.join(ROOT, "prt_b200")])
    return DEMO


def write_scene(path, desc, recompute_normals=False):
    with open(path, "wb") as f:
        f.write(struct.pack("<I", len(desc["meshes"])))
        for m in desc["meshes"]:
            idx = np.ascontiguousarray(m["indices"], np.uint32); pos = np.ascontiguousarray(m["positions"], np.float32)
            has = 0 if m.get("normals") is None else (2 if recompute_normals else 1)
            f.write(struct.pack("<4I", idx.shape[0], pos.shape[0], len(m["materials"]), has))
            f.write(idx.tobytes()); f.write(pos.tobytes())
            if has:
                f.write(np.ascontiguousarray(m["normals"], np.float32).tobytes())
            f.write(np.ascontiguousarray(m["prim_material"], np.uint32).tobytes())
            for mm in m["materials"]:
                f.write(struct.pack("<6fI", *mm["diffuse"], *mm.get("emissive", (0, 0, 0)), int(mm.get("refl", 0))))
        light = desc.get("light")
        if light is not None and light["type"] == "directional":
            f.write(struct.pack("<I6f", 1, *[float(x) for x in light["dir"]], *[float(x) for x in light["intensity"]]))
        else:
            f.write(struct.pack("<I", 0))
        f.write(struct.pack("<6f", *desc["camera"]["pos"], *desc["camera"]["dir"]))


def test_cpp_host_api_builds_scene_without_gpu(tmp_path):
    demo = build_demo()
    desc = scenes.cornell_teapot(48, 32, teapot_refl=scenes.DIFFUSE, light="directional")
    write_scene(tmp_path / "s.bin", desc)
    out = subprocess.run([demo, str(tmp_path / "s.bin"), str(tmp_path / "o.raw"), "48", "32", "8", "3", "-1"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ctx = prt_b200.load_scene(desc, device=-1)
    assert f"radius {ctx.radius():.9g}" in out.stdout
    # the options API through the C++ mirror (Scene::buildFastTraversal / setFastTraversal)
    out = subprocess.run([demo, str(tmp_path / "s.bin"), str(tmp_path / "o.raw"), "48", "32", "8", "3", "-1"], capture_output=True, text=True,
                         env=dict(os.environ, PRT_DEMO_FAST="1"))
    assert out.returncode == 0 and "traversal fast" in out.stdout, out.stdout + out.stderr


def test_cpp_vertex_normals_match_reference_algorithm(tmp_path, oracle_mod):
    """Mesh::calculateVertexNormals in prt_host.h vs the oracle's restatement (mesh.cpp:108-149)."""
    src = tmp_path / "vn.cpp"
    src.write_text('#include "prt_host.h"\n#include <cstdio>\nint main(int, char** a){ using namespace prt; FILE* f=fopen(a[1],"rb"); uint32_t n[2]; fread(n,4,2,f);'
                   ' Mesh m; m.create(n[0],n[1],1,false); fread(m.getIndexBuffer(),12,n[0],f); fread(m.getPositionBuffer(),12,n[1],f); fclose(f);'
                   ' m.calculateVertexNormals(); f=fopen(a[2],"wb"); fwrite(m.getNormalBuffer(),12,n[1],f); fclose(f); return 0; }\n')
    exe = tmp_path / "vn"
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-I" + HOST, str(src), "-o", str(exe)])
    v, fcs = scenes.icosphere(3)
    pos = (v * (1 + 0.1 * np.sin(v[:, :1] * 9))).astype(np.float32); idx = fcs.astype(np.uint32)
    with open(tmp_path / "in.bin", "wb") as f:
        f.write(struct.pack("<2I", idx.shape[0], pos.shape[0])); f.write(idx.tobytes()); f.write(pos.tobytes())
    subprocess.check_call([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "out.bin")])
    got = np.fromfile(tmp_path / "out.bin", np.float32).reshape(-1, 3)
    want = oracle_mod.calc_vertex_normals(idx, pos, kind="oracle")
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_cpp_traceblock_matches_c_abi_render(tmp_path):
    demo = build_demo()
    W, H, spp, seed = 80, 48, 16, 5
    desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
    write_scene(tmp_path / "s.bin", desc)
    ctx = prt_b200.load_scene(desc, 0)
    want = ctx.render(spp, seed)
    for tile in (0, 16):
        out = subprocess.run([demo, str(tmp_path / "s.bin"), str(tmp_path / "o.raw"), str(W), str(H), str(spp), str(seed), "0", str(tile)],
                             capture_output=True, text=True)
        assert out.returncode == 0, out.stderr
        got = np.fromfile(tmp_path / "o.raw", np.float32).reshape(H, W, 3)
        assert np.array_equal(got, want), f"tile={tile}"
        assert "centre ray: t=" in out.stdout
    ctx.close()


@pytest.mark.gpu
def test_cpp_output_files_and_gbuffer(tmp_path, oracle_mod):
    """Image::savePpm / Image::saveExr / GbufferVisualizer of the C++ host API: the PPM file is the reference's header +
    the device-converted payload, the EXR file decodes (with the reference's own tinyexr, when oracle/_ref is present)
    to the half-rounded pixels, and the normal AOV equals the C-ABI call."""
    import ctypes as C
    demo = build_demo()
    W, H, spp, seed = 80, 48, 16, 5
    desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
    write_scene(tmp_path / "s.bin", desc)
    out = subprocess.run([demo, str(tmp_path / "s.bin"), str(tmp_path / "o.raw"), str(W), str(H), str(spp), str(seed), "0"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ctx = prt_b200.load_scene(desc, 0)
    img = np.fromfile(tmp_path / "o.raw", np.float32).reshape(H, W, 3)
    ppm = open(str(tmp_path / "o.raw") + ".ppm", "rb").read()
    hdr = f"P6\n{W} {H}\n255\n".encode()
    assert ppm[:len(hdr)] == hdr
    assert np.array_equal(np.frombuffer(ppm[len(hdr):], np.uint8).reshape(H, W, 3), oracle_mod.image_ppm(img, True, kind="oracle"))
    assert np.array_equal(np.fromfile(str(tmp_path / "o.raw") + ".normal.raw", np.float32).reshape(H, W, 3), ctx.gbuffer(2, seed))
    exr = open(str(tmp_path / "o.raw") + ".exr", "rb").read()
    assert exr[:4] == bytes([0x76, 0x2f, 0x31, 0x01])
    want = oracle_mod.image_exr(img, kind="oracle")
    # scanlines are stored raw: the last H*(8 + 6W) bytes are [y, size, B row, G row, R row] per line
    body = np.frombuffer(exr[-H * (8 + 6 * W):], np.uint8).reshape(H, 8 + 6 * W)
    rows = body[:, 8:].copy().view(np.float16).reshape(H, 3, W).astype(np.float32)
    assert np.array_equal(rows[:, 2], want[..., 0]) and np.array_equal(rows[:, 1], want[..., 1]) and np.array_equal(rows[:, 0], want[..., 2])
    if oracle_mod.available():
        lib = C.CDLL(oracle_mod.REF_SO)
        p = C.POINTER(C.c_float)(); w = C.c_int(); h = C.c_int(); err = C.c_char_p()
        assert lib.LoadEXR(C.byref(p), C.byref(w), C.byref(h), (str(tmp_path / "o.raw") + ".exr").encode(), C.byref(err)) == 0, err.value
        rgba = np.ctypeslib.as_array(p, shape=(H, W, 4)).copy()
        assert (w.value, h.value) == (W, H) and np.array_equal(rgba, want)
    ctx.close()


def _gpu_count():
    from prt_b200 import capi
    return capi.lib().prt_cuda_device_count()


@pytest.mark.gpu
@pytest.mark.skipif(_gpu_count() < 2, reason="needs 2 GPUs")
def test_cpp_multi_gpu_traceblock_matches_single_gpu(tmp_path):
    """the C++ demo renders C3's scene through PathTracer::TraceBlock on a multi-GPU Scene (no Python, no
    torch: bands per GPU + ncclAllGather + de-tile inside libprt_cuda.so) and the frame equals the 1-GPU frame bit for bit."""
    demo = build_demo()
    W, H, spp, seed = 200, 117, 16, 7
    desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.SPECULAR, light=None)     # the demo's scene format carries no sky image
    write_scene(tmp_path / "s.bin", desc)
    n = min(_gpu_count(), 4)
    one = subprocess.run([demo, str(tmp_path / "s.bin"), str(tmp_path / "one.raw"), str(W), str(H), str(spp), str(seed), "0"], capture_output=True, text=True)
    assert one.returncode == 0, one.stderr
    devs = ",".join(str(d) for d in range(n))
    many = subprocess.run([demo, str(tmp_path / "s.bin"), str(tmp_path / "many.raw"), str(W), str(H), str(spp), str(seed), devs], capture_output=True, text=True)
    assert many.returncode == 0, many.stderr
    assert f"gpus {n}" in many.stdout and "gpus 1" in one.stdout
    a = np.fromfile(tmp_path / "one.raw", np.float32).reshape(H, W, 3)
    b = np.fromfile(tmp_path / "many.raw", np.float32).reshape(H, W, 3)
    assert a.max() > 0 and np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"{int((a != b).any(axis=2).sum())} pixels differ"
    # the one-ray forms and Scene::getSurfaceProperties go through the same multi-GPU context
    line = lambda out, key: [l for l in out.stdout.splitlines() if l.startswith(key)]
    assert line(one, "centre ray") == line(many, "centre ray") and line(one, "centre surface") == line(many, "centre surface")
