"""Loader -> device path: Mesh::loadObj / Material::load / Texture::load of the host API
(prt_b200/host/prt_loader.h) against the reference's loader (tinyobjloader + stb_image) on a generated
OBJ + MTL + TGA scene: triangles and quads, v/vt/vn corner forms, negative indices, faces before any usemtl, an unknown
material, Kd-less map_Kd, RLE / uncompressed / bottom- and top-origin TGAs, an alpha-holed diffuse map, a grey map and
a 24-bit normal map used as map_bump.  Golden: tests/golden/loader.bin (the reference's dump, written where
/root/reference exists); live comparison when oracle/_ref is present.  Byte-exact."""
import os
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "prt_b200", "host")
GOLD = os.path.join(ROOT, "tests", "golden", "loader.bin")


def write_tga(path, img, rle=False, top_origin=False, grey=False):
    """img: (H, W) uint8 grey or (H, W, 3|4) uint8 RGB(A), row 0 = top."""
    h, w = img.shape[:2]
    comp = 1 if grey else img.shape[2]
    rows = img if top_origin else img[::-1]
    px = rows.reshape(h * w, comp)
    if comp >= 3:
        px = px[:, [2, 1, 0] + ([3] if comp == 4 else [])]
    hdr = bytearray(18)
    hdr[2] = (11 if grey else 10) if rle else (3 if grey else 2)
    hdr[12:14] = w.to_bytes(2, "little"); hdr[14:16] = h.to_bytes(2, "little")
    hdr[16] = 8 * comp; hdr[17] = (0x20 if top_origin else 0) | (8 if comp == 4 else 0)
    body = bytearray()
    if not rle:
        body += px.tobytes()
    else:
        i, n = 0, px.shape[0]
        while i < n:                                  # alternate run / raw packets
            j = i
            while j + 1 < n and j - i < 127 and (px[j + 1] == px[i]).all():
                j += 1
            if j > i:
                body.append(0x80 | (j - i)); body += px[i].tobytes(); i = j + 1
            else:
                k = i
                while k + 1 < n and k - i < 127 and not (px[k + 1] == px[k]).all():
                    k += 1
                body.append(k - i); body += px[i:k + 1].tobytes(); i = k + 1
    open(path, "wb").write(bytes(hdr) + bytes(body))


def make_scene(d):
    rng = np.random.RandomState(4)
    tex = rng.randint(0, 256, (8, 16, 4)).astype(np.uint8); tex[2:5, 3:9, 3] = 0; tex[:2] = tex[0, 0]      # alpha holes, constant rows (runs)
    write_tga(os.path.join(d, "leaf.tga"), tex, rle=True)
    write_tga(os.path.join(d, "wall.tga"), rng.randint(0, 256, (4, 4, 3)).astype(np.uint8), top_origin=True)
    write_tga(os.path.join(d, "grey.tga"), rng.randint(0, 256, (5, 7)).astype(np.uint8), grey=True)
    nrm = rng.randint(0, 256, (8, 8, 3)).astype(np.uint8); nrm[..., 2] |= 128
    write_tga(os.path.join(d, "nrm.tga"), nrm)
    open(os.path.join(d, "s.mtl"), "w").write(
        "# materials\nnewmtl wall\nKa 0.1 0.1 0.1\nKd 0.8 0.7 0.6\nKs 0 0 0\nKe 0 0 0\nmap_Kd wall.tga\nmap_bump -bm 0.5 nrm.tga\n\n"
        "newmtl leaf\nmap_Kd leaf.tga\nbump grey.tga\n\nnewmtl lamp\nKd 0.5 0.25 1\nKe 12.5 11 9.75\nillum 2\n")
    v = rng.uniform(-2, 2, (12, 3)).round(4)
    vt = rng.uniform(-1, 2, (9, 2)).round(3)
    L = ["# test scene", "mtllib s.mtl"]
    L += [f"v {a} {b} {c}" for a, b, c in v[:8]]
    L += [f"vt {a} {b}" for a, b in vt]
    L += ["vn 0 0 1", "f 1 2 3", "o first", "usemtl wall", "f 1/1 2/2 3/3 4/4", "f 2/5/1 3/6/1 5/7/1", "g grp", "usemtl leaf", "f 5//1 6//1 7//1 8//1"]
    L += [f"v {a} {b} {c}" for a, b, c in v[8:]]
    L += ["usemtl lamp", "f -1/-1 -2/-2 -3/-3", "f 9/1 10/2 11/3 12/4", "usemtl nosuch", "f 1 5 9", "s off", "f 3/9 7/8 11/7"]
    open(os.path.join(d, "s.obj"), "w").write("\n".join(L) + "\n")
    return os.path.join(d, "s.obj")


def write_png(path, img, ctype, palette=None, trns=None, level=6, split_idat=0, depth=8, interlace=False):
    """Non-interlaced or Adam7 PNG at any legal bit depth.  img holds SAMPLE VALUES (0 .. 2**depth - 1): (H, W) for colour types
    0 / 3, (H, W, 2 | 3 | 4) for 4 / 2 / 6.  Row y of every (sub-)image is stored with filter type y % 5 (none, sub, up, average,
    Paeth), so every un-filter path of the decoder runs."""
    import struct, zlib
    h, w = img.shape[:2]
    ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    fb = max(1, ch * depth // 8)

    def pack_rows(sub):                                   # (ph, pw, ch) sample values -> (ph, stride) bytes
        ph, pw = sub.shape[:2]
        v = sub.reshape(ph, pw * ch).astype(np.uint32)
        if depth == 8: return v.astype(np.uint8)
        if depth == 16: return np.stack([v >> 8, v & 255], axis=-1).reshape(ph, -1).astype(np.uint8)
        per = 8 // depth
        pad = (-v.shape[1]) % per
        v = np.pad(v, ((0, 0), (0, pad)))
        out = np.zeros((ph, v.shape[1] // per), np.uint32)
        for k in range(per): out |= v[:, k::per] << (8 - depth * (k + 1))
        return out.astype(np.uint8)

    def filtered(rows_u8):
        rows = rows_u8.astype(np.int32); out = bytearray()
        for y in range(rows.shape[0]):
            cur = rows[y]; up = rows[y - 1] if y else np.zeros_like(cur)
            a = np.concatenate([np.zeros(fb, np.int32), cur[:-fb]]) if cur.size > fb else np.zeros_like(cur)
            c = np.concatenate([np.zeros(fb, np.int32), up[:-fb]]) if cur.size > fb else np.zeros_like(cur)
            f = y % 5
            if f == 0: pred = 0
            elif f == 1: pred = a
            elif f == 2: pred = up
            elif f == 3: pred = (a + up) >> 1
            else:
                pp = a + up - c; pa, pb, pc = abs(pp - a), abs(pp - up), abs(pp - c)
                pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, up, c))
            out.append(f); out += ((cur - pred) & 255).astype(np.uint8).tobytes()
        return out

    im3 = img.reshape(h, w, ch)
    raw = bytearray()
    if not interlace:
        raw += filtered(pack_rows(im3))
    else:
        for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            sub = im3[y0::dy, x0::dx]
            if sub.shape[0] and sub.shape[1]: raw += filtered(pack_rows(sub))

    def chunk(t, body):
        return struct.pack(">I", len(body)) + t + body + struct.pack(">I", zlib.crc32(t + body) & 0xffffffff)
    z = zlib.compress(bytes(raw), level)
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 1 if interlace else 0))
    if palette is not None: out += chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if trns is not None: out += chunk(b"tRNS", bytes(trns))
    if split_idat: out += chunk(b"IDAT", z[:split_idat]) + chunk(b"IDAT", z[split_idat:])
    else: out += chunk(b"IDAT", z)
    open(path, "wb").write(out + chunk(b"IEND", b""))


def make_png_scene(d):
    """Six PNG flavours behind map_Kd / bump of three materials on two quads and a triangle."""
    rng = np.random.RandomState(11)
    rgba = rng.randint(0, 256, (9, 12, 4)).astype(np.uint8); rgba[3:6, 2:8, 3] = 0; rgba[:2] = rgba[0, 0]
    write_png(os.path.join(d, "leaf.png"), rgba, 6, split_idat=40)                                   # RGBA, alpha holes, two IDAT chunks
    write_png(os.path.join(d, "grey.png"), rng.randint(0, 256, (7, 5)).astype(np.uint8), 0, trns=b"\x00\x40", level=0)   # grey + colour key (dropped), stored blocks
    rgb = rng.randint(0, 256, (6, 6, 3)).astype(np.uint8); rgb[1, 1] = (9, 8, 7); rgb[4, 2] = (9, 8, 7)
    write_png(os.path.join(d, "wall.png"), rgb, 2, trns=b"\x00\x09\x00\x08\x00\x07", level=9)         # RGB + colour key -> alpha 0 at the key
    nrm = rng.randint(0, 256, (8, 8, 3)).astype(np.uint8); nrm[..., 2] |= 128
    write_png(os.path.join(d, "nrm.png"), nrm, 2, level=1)                                           # 3-component map_bump: normal map -> height
    pal = rng.randint(0, 256, (5, 3)).astype(np.uint8)
    write_png(os.path.join(d, "pal.png"), rng.randint(0, 5, (4, 9)).astype(np.uint8), 3, palette=pal, trns=b"\xff\x00\x80")   # palette + tRNS
    write_png(os.path.join(d, "ga.png"), rng.randint(0, 256, (5, 4, 2)).astype(np.uint8), 4)         # grey + alpha as a bump map
    open(os.path.join(d, "p.mtl"), "w").write(
        "newmtl wall\nKd 0.8 0.7 0.6\nmap_Kd wall.png\nmap_bump nrm.png\n\nnewmtl leaf\nKd 1 1 1\nmap_Kd leaf.png\nbump grey.png\n\n"
        "newmtl sign\nKd 0.5 0.5 0.5\nmap_Kd pal.png\nbump ga.png\n")
    L = ["mtllib p.mtl", "v 0 0 0", "v 1 0 0", "v 1 1 0", "v 0 1 0", "v 0 0 1", "v 1 0 1", "v 1 1 1", "v 0 1 1", "vt 0 0", "vt 1 0", "vt 1 1", "vt 0 1",
         "usemtl wall", "f 1/1 2/2 3/3 4/4", "usemtl leaf", "f 5/1 6/2 7/3 8/4", "usemtl sign", "f 1/1 6/2 8/4"]
    open(os.path.join(d, "p.obj"), "w").write("\n".join(L) + "\n")
    return os.path.join(d, "p.obj")


def build_dump():
    exe = os.path.join(HOST, "build", "loader_dump")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    src = os.path.join(HOST, "loader_dump.cpp")
    deps = [src, os.path.join(HOST, "prt_host.h"), os.path.join(HOST, "prt_loader.h")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(p) for p in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", src, "-o", exe, "-L" + os.path.join(ROOT, "prt_b200"),
                               "-lprt_cuda", "-Wl,-rpath," + os.path.join(ROOT, "prt_b200")])
    return exe


def test_loader_matches_reference_dump(tmp_path, oracle_mod):
    import prt_b200  # noqa: F401  (makes sure libprt_cuda.so is built)
    obj = make_scene(str(tmp_path))
    out = subprocess.run([build_dump(), obj, str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "scene.add: prt_cuda: add_mesh: primMaterial[0] = 4294967295" in out.stdout      # faces without a material: rejected, not UB
    ours = open(tmp_path / "ours.bin", "rb").read()
    if oracle_mod.available():
        import ctypes as C
        with oracle_mod.quiet_stdout():
            r = C.CDLL(oracle_mod.REF_SO).ref_load_obj_dump(obj.encode(), str(tmp_path / "ref.bin").encode())
        assert r == 0
        ref = open(tmp_path / "ref.bin", "rb").read()
        if os.environ.get("PRT_WRITE_GOLDEN"):
            open(GOLD, "wb").write(ref)
        assert ours == ref, f"loader dump differs from the live reference ({len(ours)} vs {len(ref)} bytes)"
    gold = open(GOLD, "rb").read()
    hdr = np.frombuffer(gold[:12], np.uint32)
    assert tuple(hdr) == (11, 12, 3)
    assert ours == gold, "loader dump differs from tests/golden/loader.bin (reference output)"


def test_png_textures_match_reference_dump(tmp_path, oracle_mod):
    """PNG ingestion (own inflate + un-filter + stb_image's component rules, prt_loader.h) against the reference's loader:
    RGBA, keyed grey, keyed RGB, RGB normal map as map_bump, palette with tRNS, grey+alpha; filters 0-4, stored / fixed /
    dynamic deflate blocks, split IDAT.  Golden: tests/golden/loader_png.bin (the reference's dump)."""
    import prt_b200  # noqa: F401
    gold_path = os.path.join(ROOT, "tests", "golden", "loader_png.bin")
    obj = make_png_scene(str(tmp_path))
    out = subprocess.run([build_dump(), obj, str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ours = open(tmp_path / "ours.bin", "rb").read()
    if oracle_mod.available():
        import ctypes as C
        with oracle_mod.quiet_stdout():
            r = C.CDLL(oracle_mod.REF_SO).ref_load_obj_dump(obj.encode(), str(tmp_path / "ref.bin").encode())
        assert r == 0
        ref = open(tmp_path / "ref.bin", "rb").read()
        if os.environ.get("PRT_WRITE_GOLDEN"):
            open(gold_path, "wb").write(ref)
        assert ours == ref, f"PNG loader dump differs from the live reference ({len(ours)} vs {len(ref)} bytes)"
    gold = open(gold_path, "rb").read()
    assert ours == gold, "PNG loader dump differs from tests/golden/loader_png.bin (reference output)"
    _, textures = parse_dump(ours)
    assert [(t["width"], t["height"], t["comp"]) for t in textures] == [(6, 6, 4), (8, 8, 1), (12, 9, 4), (5, 7, 1), (9, 4, 4), (4, 5, 4)]
    # a PNG whose header does not match its data is refused, not mis-decoded
    import struct, zlib
    bad = bytearray(open(os.path.join(str(tmp_path), "nrm.png"), "rb").read()); bad[24] = 16
    bad[29:33] = struct.pack(">I", zlib.crc32(bytes(bad[12:29])) & 0xffffffff)
    open(os.path.join(str(tmp_path), "nrm.png"), "wb").write(bytes(bad))
    out = subprocess.run([build_dump(), obj, str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 1 and "cannot decode" in out.stderr


EXR_GOLD = os.path.join(ROOT, "tests", "golden", "loader_exr.npz")


def _exr_cases():
    rng = np.random.RandomState(23)
    sky = np.abs(rng.standard_cauchy((40, 64, 4))).astype(np.float32); sky[5:9, 10:30] = 0.0; sky[20, :, :3] = 65504.0
    tiny = rng.rand(7, 9, 3).astype(np.float32) * 1e-6                      # half subnormals; < 16 x 16: tinyexr stores it uncompressed
    return [("sky_f32_rgba", sky, 0), ("sky_f16_rgba", sky, 1), ("rgb_f32", rng.randn(33, 17, 3).astype(np.float32), 0),
            ("tiny_f16_rgb", tiny, 1), ("one_channel", rng.rand(18, 20, 1).astype(np.float32), 0)]


def write_exr(path, img, names, compression, half=False):
    """Scan-line EXR with compression 0 (none), 1 (RLE) or 2 (ZIPS, one line per chunk) — the two tinyexr's SaveEXR never writes.
    img: (H, W, C) float32, channel c named names[c]; channels are stored in alphabetical order as the format requires."""
    import struct, zlib
    h, w, _ = img.shape
    order = sorted(range(len(names)), key=lambda c: names[c])
    def attr(name, typ, val):
        return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<I", len(val)) + val
    chl = b"".join(names[c].encode() + b"\0" + struct.pack("<IB3xII", 1 if half else 2, 0, 1, 1) for c in order) + b"\0"
    box = struct.pack("<4i", 0, 0, w - 1, h - 1)
    hdr = struct.pack("<II", 20000630, 2) + attr("channels", "chlist", chl) + attr("compression", "compression", bytes([compression])) + \
        attr("dataWindow", "box2i", box) + attr("displayWindow", "box2i", box) + attr("lineOrder", "lineOrder", b"\0") + \
        attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) + attr("screenWindowCenter", "v2f", struct.pack("<2f", 0, 0)) + \
        attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0"
    chunks = []
    for y in range(h):
        line = b"".join((img[y, :, c].astype(np.float16) if half else img[y, :, c]).tobytes() for c in order)
        if compression:
            a = np.frombuffer(line, np.uint8)
            t = np.concatenate([a[0::2], a[1::2]]).astype(np.int32)                       # interleave halves, then delta + 128
            t[1:] = (t[1:] - t[:-1] + 128) & 255
            t = t.astype(np.uint8).tobytes()
            if compression == 2:
                z = zlib.compress(t, 6)
            else:
                z = bytearray(); i = 0
                while i < len(t):
                    j = i
                    while j + 1 < len(t) and t[j + 1] == t[i] and j - i < 126: j += 1
                    if j - i >= 2: z += bytes([j - i, t[i]]); i = j + 1
                    else:
                        k = i
                        while k + 1 < len(t) and k - i < 126 and not (k + 2 < len(t) and t[k] == t[k + 1] == t[k + 2]): k += 1
                        z += struct.pack("b", -(k - i + 1)) + t[i:k + 1]; i = k + 1
                z = bytes(z)
            data = z if len(z) < len(line) else line
        else:
            data = line
        chunks.append(struct.pack("<iI", y, len(data)) + data)
    off = len(hdr) + 8 * h
    table = b""
    for c in chunks:
        table += struct.pack("<Q", off); off += len(c)
    open(path, "wb").write(hdr + table + b"".join(chunks))


def test_exr_rle_and_zips_against_tinyexr(tmp_path, oracle_mod):
    """RLE and ZIPS scan-line files (written here, tinyexr's SaveEXR only writes NONE / ZIP): our decoder against tinyexr's
    LoadEXR where the reference library is present, and against the source pixels (FLOAT is lossless, HALF is the numpy cast)."""
    import ctypes as C
    import prt_b200  # noqa: F401
    exe = build_dump()
    rng = np.random.RandomState(5)
    img = rng.randn(11, 23, 4).astype(np.float32); img[3:6] = 0.25; img[:, 5:9, 1] = -3.0
    for comp in (1, 2):
        for half in (False, True):
            for names in (["R", "G", "B", "A"], ["B", "G", "R"], ["Y"]):
                src = img[..., :len(names)]
                path = str(tmp_path / f"c{comp}{int(half)}{len(names)}.exr")
                write_exr(path, src, names, comp, half)
                out = subprocess.run([exe, "--exr", path, str(tmp_path / "o.bin")], capture_output=True, text=True)
                assert out.returncode == 0, out.stderr
                got = np.frombuffer(open(tmp_path / "o.bin", "rb").read()[8:], np.float32).reshape(11, 23, 4)
                v = src.astype(np.float16).astype(np.float32) if half else src
                exp = np.ones((11, 23, 4), np.float32)
                if len(names) == 1: exp[:] = v
                else:
                    for c, nm in enumerate(names): exp[..., "RGBA".index(nm)] = v[..., c]
                assert np.array_equal(got.view(np.uint32), exp.view(np.uint32)), (comp, half, names)
                if oracle_mod.available():
                    lib = C.CDLL(oracle_mod.REF_SO)
                    o = C.POINTER(C.c_float)(); w = C.c_int(); h = C.c_int(); err = C.c_char_p()
                    assert lib.LoadEXR(C.byref(o), C.byref(w), C.byref(h), path.encode(), C.byref(err)) == 0, err.value
                    ref = np.ctypeslib.as_array(o, shape=(11, 23, 4))
                    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), (comp, half, names, "tinyexr")


PIZ_GOLD = os.path.join(ROOT, "tests", "golden", "loader_exr_piz.npz")


def test_exr_piz_matches_tinyexr(tmp_path, oracle_mod):
    """PIZ-compressed scan-line files (wavelet + Huffman; written once with OpenCV, committed in tests/golden/loader_exr_piz.npz
    together with tinyexr's decode): HALF and FLOAT, 3 and 4 channels, sizes that leave partial 32-line blocks and odd wavelet
    levels, a constant image and a > 14-bit value range.  Live comparison with the reference's LoadEXR where it is present."""
    import ctypes as C
    import prt_b200  # noqa: F401
    exe = build_dump()
    if os.environ.get("PRT_WRITE_GOLDEN"):
        os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
        import cv2
        rng = np.random.RandomState(9)
        yy, xx = np.mgrid[0:70, 0:53]
        smooth = np.stack([np.sin(xx / 9.0) + 1.2, np.cos(yy / 7.0) + 1.5, (xx + yy) / 60.0], -1).astype(np.float32)
        cases = {"half_rgb": (smooth, cv2.IMWRITE_EXR_TYPE_HALF), "float_rgb": (smooth * 1000.0 + rng.rand(70, 53, 3).astype(np.float32), cv2.IMWRITE_EXR_TYPE_FLOAT),
                 "half_rgba": (np.dstack([smooth[:33, :17], np.ones((33, 17), np.float32)]), cv2.IMWRITE_EXR_TYPE_HALF),
                 "noise_half": (rng.standard_cauchy((37, 41, 3)).astype(np.float32), cv2.IMWRITE_EXR_TYPE_HALF),
                 "const": (np.full((40, 40, 3), 0.25, np.float32), cv2.IMWRITE_EXR_TYPE_FLOAT), "tiny": (smooth[:3, :5], cv2.IMWRITE_EXR_TYPE_HALF)}
        lib = C.CDLL(oracle_mod.REF_SO)
        save = {}
        for name, (img, typ) in cases.items():
            path = str(tmp_path / (name + ".exr"))
            assert cv2.imwrite(path, img, [cv2.IMWRITE_EXR_COMPRESSION, cv2.IMWRITE_EXR_COMPRESSION_PIZ, cv2.IMWRITE_EXR_TYPE, typ])
            o = C.POINTER(C.c_float)(); w = C.c_int(); h = C.c_int(); err = C.c_char_p()
            assert lib.LoadEXR(C.byref(o), C.byref(w), C.byref(h), path.encode(), C.byref(err)) == 0, err.value
            save["file_" + name] = np.frombuffer(open(path, "rb").read(), np.uint8)
            save["rgba_" + name] = np.ctypeslib.as_array(o, shape=(h.value, w.value, 4)).copy()
        np.savez_compressed(PIZ_GOLD, **save)
    g = np.load(PIZ_GOLD)
    names = [k[5:] for k in g.files if k.startswith("file_")]
    assert len(names) == 6
    for name in names:
        b = g["file_" + name].tobytes()
        assert b[b.find(b"compression\0compression\0") + 28] == 4                 # really PIZ
        path = str(tmp_path / (name + "_in.exr"))
        open(path, "wb").write(b)
        out = subprocess.run([exe, "--exr", path, str(tmp_path / "o.bin")], capture_output=True, text=True)
        assert out.returncode == 0, name + ": " + out.stderr
        gold = g["rgba_" + name]
        raw = open(tmp_path / "o.bin", "rb").read()
        assert tuple(np.frombuffer(raw[:8], np.uint32)) == (gold.shape[1], gold.shape[0])
        got = np.frombuffer(raw[8:], np.uint32).reshape(gold.shape)
        assert np.array_equal(got, gold.view(np.uint32)), name
        if oracle_mod.available():
            lib = C.CDLL(oracle_mod.REF_SO)
            o = C.POINTER(C.c_float)(); w = C.c_int(); h = C.c_int(); err = C.c_char_p()
            assert lib.LoadEXR(C.byref(o), C.byref(w), C.byref(h), path.encode(), C.byref(err)) == 0
            assert np.array_equal(got, np.ctypeslib.as_array(o, shape=gold.shape).view(np.uint32)), name + " (live tinyexr)"


def test_exr_matches_tinyexr(tmp_path, oracle_mod):
    """Texture::loadExr / Scene::setInfiniteAreaLight(path) (prt_loader.h decodeExr: header, offset table, ZIP blocks through
    the own inflate + predictor + de-interleave, HALF / FLOAT channels, tinyexr's channel rules) against the reference's
    tinyexr on files its SaveEXR wrote.  Golden: tests/golden/loader_exr.npz (files + the reference's decode)."""
    import ctypes as C
    import prt_b200  # noqa: F401
    exe = build_dump()
    files, want = {}, {}
    if oracle_mod.available():
        lib = C.CDLL(oracle_mod.REF_SO)
        for name, img, fp16 in _exr_cases():
            path = str(tmp_path / (name + ".exr")).encode()
            err = C.c_char_p()
            hgt, wid, comps = img.shape
            assert lib.SaveEXR(np.ascontiguousarray(img).ctypes.data_as(C.c_void_p), wid, hgt, comps, fp16, path, C.byref(err)) == 0
            out = C.POINTER(C.c_float)(); w = C.c_int(); h = C.c_int()
            assert lib.LoadEXR(C.byref(out), C.byref(w), C.byref(h), path, C.byref(err)) == 0 and (w.value, h.value) == (wid, hgt)
            want[name] = np.ctypeslib.as_array(out, shape=(hgt, wid, 4)).copy()
            files[name] = np.frombuffer(open(path, "rb").read(), np.uint8)
        if os.environ.get("PRT_WRITE_GOLDEN"):
            np.savez_compressed(EXR_GOLD, **{"file_" + k: v for k, v in files.items()}, **{"rgba_" + k: v for k, v in want.items()})
    g = np.load(EXR_GOLD)
    for name, _, _ in _exr_cases():
        for src in ([files[name]] if name in files else []) + [g["file_" + name]]:
            path = str(tmp_path / (name + "_in.exr"))
            open(path, "wb").write(src.tobytes())
            out = subprocess.run([exe, "--exr", path, str(tmp_path / "o.bin")], capture_output=True, text=True)
            assert out.returncode == 0, out.stderr
            b = open(tmp_path / "o.bin", "rb").read()
            ref = want.get(name, g["rgba_" + name]); gold = g["rgba_" + name]
            assert tuple(np.frombuffer(b[:8], np.uint32)) == (ref.shape[1], ref.shape[0])
            got = np.frombuffer(b[8:], np.uint32).reshape(ref.shape)
            assert np.array_equal(got, ref.view(np.uint32)), name
            assert np.array_equal(got, gold.view(np.uint32)), name + " (golden)"
    # refused, not mis-decoded: a tiled-flag file and a truncated one
    bad = bytearray(g["file_rgb_f32"].tobytes()); bad[5] |= 0x02
    open(tmp_path / "bad.exr", "wb").write(bytes(bad))
    out = subprocess.run([exe, "--exr", str(tmp_path / "bad.exr"), str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 1 and "cannot decode" in out.stderr
    open(tmp_path / "cut.exr", "wb").write(g["file_sky_f32_rgba"].tobytes()[:-100])
    out = subprocess.run([exe, "--exr", str(tmp_path / "cut.exr"), str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 1


def test_png_depths_and_interlace_match_reference_dump(tmp_path, oracle_mod):
    """The rest of what stb_image's PNG reader accepts: 1 / 2 / 4-bit grey and palette, 16-bit samples (reduced to the high byte,
    colour key compared at 16 bits), Adam7 interlace with odd sizes.  Golden: tests/golden/loader_png2.bin (the reference's dump)."""
    import prt_b200  # noqa: F401
    d = str(tmp_path)
    rng = np.random.RandomState(31)
    pal = rng.randint(0, 256, (16, 3)).astype(np.uint8)
    rgb16 = rng.randint(0, 65536, (5, 7, 3)); rgb16[2, 3] = (0x1234, 0x00ff, 0xff00); rgb16[4, 1] = (0x1234, 0x00ff, 0xff01)
    # (name, image of sample values, colour type, kwargs) in (map_Kd, bump) pairs.  Diffuse maps are never 1-component files: the reference's
    # Texture::isAlphaTestRequired (texture.cpp:338-350) reads texels[4 * i + 3] whatever the component count — a heap over-read for those.
    files = [
        ("p1", rng.randint(0, 2, (7, 10)), 3, dict(depth=1, palette=pal[:2])), ("g1", rng.randint(0, 2, (9, 13)), 0, dict(depth=1)),
        ("p2", rng.randint(0, 4, (9, 9)), 3, dict(depth=2, palette=pal[:4], trns=b"\x10\xff\x00", interlace=True)), ("g2", rng.randint(0, 4, (6, 7)), 0, dict(depth=2)),
        ("p4", rng.randint(0, 16, (3, 11)), 3, dict(depth=4, palette=pal)), ("g4", rng.randint(0, 16, (5, 9)), 0, dict(depth=4, interlace=True)),
        ("rgb16k", rgb16, 2, dict(depth=16, trns=b"\x12\x34\x00\xff\xff\x00")), ("g16", rng.randint(0, 65536, (6, 5)), 0, dict(depth=16)),
        ("rgba16", rng.randint(0, 65536, (4, 6, 4)), 6, dict(depth=16, interlace=True)), ("rgb16kb", rgb16, 2, dict(depth=16, trns=b"\x12\x34\x00\xff\xff\x00", interlace=True)),
        ("rgba8i", rng.randint(0, 256, (11, 17, 4)), 6, dict(interlace=True)), ("rgb8i", rng.randint(0, 256, (1, 1, 3)), 2, dict(interlace=True)),
        ("ga8i", rng.randint(0, 256, (2, 3, 2)), 4, dict(interlace=True)), ("pal8i", rng.randint(0, 16, (10, 6)), 3, dict(palette=pal, interlace=True)),
        ("ga16", rng.randint(0, 65536, (7, 3, 2)), 4, dict(depth=16)), ("rgb16b", rng.randint(0, 65536, (3, 4, 3)), 2, dict(depth=16))]
    mtl, obj = [], ["mtllib q.mtl", "v 0 0 0", "v 1 0 0", "v 0 1 0", "vt 0 0", "vt 1 0", "vt 0 1"]
    for k in range(0, len(files), 2):
        for name, img, ct, kw in files[k:k + 2]:
            write_png(os.path.join(d, name + ".png"), np.asarray(img), ct, **kw)
        mtl.append(f"newmtl m{k}\nKd 0.5 0.5 0.5\nmap_Kd {files[k][0]}.png\nbump {files[k + 1][0]}.png\n")
        obj += [f"usemtl m{k}", "f 1/1 2/2 3/3"]
    open(os.path.join(d, "q.mtl"), "w").write("\n".join(mtl))
    open(os.path.join(d, "q.obj"), "w").write("\n".join(obj) + "\n")
    gold_path = os.path.join(ROOT, "tests", "golden", "loader_png2.bin")
    out = subprocess.run([build_dump(), os.path.join(d, "q.obj"), str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ours = open(tmp_path / "ours.bin", "rb").read()
    if oracle_mod.available():
        import ctypes as C
        with oracle_mod.quiet_stdout():
            r = C.CDLL(oracle_mod.REF_SO).ref_load_obj_dump(os.path.join(d, "q.obj").encode(), str(tmp_path / "ref.bin").encode())
        assert r == 0
        ref = open(tmp_path / "ref.bin", "rb").read()
        if os.environ.get("PRT_WRITE_GOLDEN"):
            open(gold_path, "wb").write(ref)
        assert ours == ref, f"PNG loader dump differs from the live reference ({len(ours)} vs {len(ref)} bytes)"
    assert ours == open(gold_path, "rb").read(), "PNG loader dump differs from tests/golden/loader_png2.bin (reference output)"
    _, textures = parse_dump(ours)
    assert [t["comp"] for t in textures] == [4, 1, 4, 1, 4, 1, 4, 1, 4, 4, 4, 1, 4, 1, 4, 1]
    # [9]: a keyed RGB file as map_bump counts 4 components and stays RGBA; [11], [13], [15]: 3-component bump maps (RGB, palette, 16-bit RGB) become heights
    key = textures[6]["texels"].reshape(5, 7, 4)
    assert key[2, 3, 3] == 0 and key[4, 1, 3] == 255                        # the colour key is compared on all 16 bits


JPG_GOLD = os.path.join(ROOT, "tests", "golden", "loader_jpg.npz")


def _jpeg_files():
    """name -> JPEG bytes: written with Pillow where it is installed (and refreshed into the golden with PRT_WRITE_GOLDEN), else the
    committed copies.  4:4:4 / 4:2:2 / 4:2:0 chroma, grey, odd sizes down to 1 x 1, optimised Huffman tables, restart intervals."""
    try:
        from PIL import Image
    except ImportError:
        g = np.load(JPG_GOLD)
        return {k[5:]: g[k].tobytes() for k in g.files if k.startswith("file_")}
    import io
    rng = np.random.RandomState(41)
    def smooth(h, w, c):
        yy, xx = np.mgrid[0:h, 0:w]
        img = np.stack([127 + 100 * np.sin(xx / (3.0 + k) + k) * np.cos(yy / (4.0 + k)) for k in range(c)], -1) + rng.randn(h, w, c) * 12
        return np.clip(img, 0, 255).astype(np.uint8)
    out = {}
    cases = [("c444", smooth(24, 40, 3), dict(quality=90, subsampling=0)), ("c422", smooth(23, 37, 3), dict(quality=75, subsampling=1)),
             ("c420", smooth(35, 29, 3), dict(quality=60, subsampling=2)), ("c420opt", smooth(17, 50, 3), dict(quality=95, subsampling=2, optimize=True)),
             ("grey", smooth(19, 21, 1)[..., 0], dict(quality=85)), ("one", smooth(1, 1, 3), dict(quality=80, subsampling=2)),
             ("noise420", rng.randint(0, 256, (16, 16, 3)).astype(np.uint8), dict(quality=30, subsampling=2)),
             ("rst", smooth(40, 48, 3), dict(quality=70, subsampling=2, restart_marker_blocks=2)), ("thin", smooth(9, 3, 3), dict(quality=88, subsampling=1)),
             ("p444", smooth(26, 33, 3), dict(quality=85, subsampling=0, progressive=True)), ("p420", smooth(45, 38, 3), dict(quality=65, subsampling=2, progressive=True)),
             ("p422opt", smooth(21, 27, 3), dict(quality=92, subsampling=1, progressive=True, optimize=True)), ("pgrey", smooth(30, 18, 1)[..., 0], dict(quality=75, progressive=True)),
             ("pnoise", rng.randint(0, 256, (24, 24, 3)).astype(np.uint8), dict(quality=50, subsampling=2, progressive=True)),
             ("prst", smooth(33, 41, 3), dict(quality=80, subsampling=2, progressive=True, restart_marker_blocks=3))]
    cases += [("cmyk", rng.randint(0, 256, (13, 19, 4)).astype(np.uint8), dict(quality=90, mode="CMYK")),            # Adobe marker, transform 0, 4 components
              ("cmyksub", np.dstack([smooth(20, 24, 3), smooth(20, 24, 1)]), dict(quality=70, mode="CMYK", progressive=True)),
              ("rgbkeep", smooth(15, 22, 3), dict(quality=85, keep_rgb=True))]                                           # stored as R, G, B
    for name, img, kw in cases:
        b = io.BytesIO()
        mode = kw.pop("mode", None)
        im = Image.fromarray(img, mode) if mode else Image.fromarray(img)
        try:
            im.save(b, "JPEG", **kw)
        except TypeError:
            kw.pop("restart_marker_blocks", None); kw.pop("keep_rgb", None); im.save(b, "JPEG", **kw)
        out[name] = b.getvalue()
    return out


def test_jpeg_textures_match_reference_dump(tmp_path, oracle_mod):
    """Sequential and progressive JPEG ingestion (prt_loader.h JpegDecoder: Huffman, spectral selection / successive approximation,
    stb_image's integer IDCT, its chroma up-sampling filters and fixed-point YCbCr -> RGB) against the reference's loader, byte for byte.  Golden: tests/golden/loader_jpg.npz (files + the reference's dump)."""
    import prt_b200  # noqa: F401
    d = str(tmp_path)
    files = _jpeg_files()
    names = sorted(files)
    for n in names:
        open(os.path.join(d, n + ".jpg"), "wb").write(files[n])
    colour = [n for n in names if n not in ("grey", "pgrey")]
    mtl, obj = [], ["mtllib j.mtl", "v 0 0 0", "v 1 0 0", "v 0 1 0", "vt 0 0", "vt 1 0", "vt 0 1"]
    for k, n in enumerate(colour):                       # grey only as a bump map (1-component diffuse maps: see the PNG test)
        bump = ("grey", "pgrey")[k // 2 % 2] if k % 2 == 0 else colour[(k + 1) % len(colour)]      # 3-component bump maps become heights
        mtl.append(f"newmtl m{k}\nKd 0.5 0.5 0.5\nmap_Kd {n}.jpg\nbump {bump}.jpg\n")
        obj += [f"usemtl m{k}", "f 1/1 2/2 3/3"]
    open(os.path.join(d, "j.mtl"), "w").write("\n".join(mtl))
    open(os.path.join(d, "j.obj"), "w").write("\n".join(obj) + "\n")
    out = subprocess.run([build_dump(), os.path.join(d, "j.obj"), str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ours = open(tmp_path / "ours.bin", "rb").read()
    if oracle_mod.available():
        import ctypes as C
        with oracle_mod.quiet_stdout():
            r = C.CDLL(oracle_mod.REF_SO).ref_load_obj_dump(os.path.join(d, "j.obj").encode(), str(tmp_path / "ref.bin").encode())
        assert r == 0
        ref = open(tmp_path / "ref.bin", "rb").read()
        if os.environ.get("PRT_WRITE_GOLDEN"):
            np.savez_compressed(JPG_GOLD, dump=np.frombuffer(ref, np.uint8), **{"file_" + k: np.frombuffer(v, np.uint8) for k, v in files.items()})
        if ours != ref:
            (_, to), (_, tr) = parse_dump(ours), parse_dump(ref)
            bad = [(i, t["width"], t["height"], t["comp"], int((t["texels"] != u["texels"]).sum()) if t["texels"].shape == u["texels"].shape else -1) for i, (t, u) in enumerate(zip(to, tr))
                   if t["comp"] != u["comp"] or not np.array_equal(t["texels"], u["texels"])]
            assert False, f"JPEG loader dump differs from the live reference: textures {bad}"
    g = np.load(JPG_GOLD)
    if all(np.array_equal(np.frombuffer(files[n], np.uint8), g["file_" + n]) for n in names if "file_" + n in g.files) and len(names) == sum(k.startswith("file_") for k in g.files):
        assert ours == g["dump"].tobytes(), "JPEG loader dump differs from tests/golden/loader_jpg.npz (reference output)"
    # an arithmetic-coded file is refused
    prog = bytearray(files["c444"]); i = prog.find(b"\xff\xc0"); prog[i + 1] = 0xc9
    open(os.path.join(d, "c444.jpg"), "wb").write(bytes(prog))
    out = subprocess.run([build_dump(), os.path.join(d, "j.obj"), str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 1 and "cannot decode" in out.stderr


def test_crlf_tabs_and_trailing_blanks_change_nothing(tmp_path):
    """The same scene with CRLF line ends, tabs between face corners and trailing blanks loads to the same buffers (as it does in the
    reference: verified against the live loader when this test was written)."""
    import prt_b200  # noqa: F401
    obj = make_scene(str(tmp_path))
    for name in ("s.obj", "s.mtl"):
        b = open(tmp_path / name, "rb").read().replace(b"\n", b"\r\n").replace(b"f 1 2 3", b"f\t1  2\t3 ").replace(b"usemtl wall", b"usemtl wall  ")
        open(tmp_path / name, "wb").write(b)
    out = subprocess.run([build_dump(), obj, str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert open(tmp_path / "ours.bin", "rb").read() == open(GOLD, "rb").read()


@pytest.mark.skipif(not os.path.exists("/root/reference/data/teapot/teapot.obj"), reason="the reference's data directory is only in the build container")
def test_real_teapot_obj_matches_reference_loader(tmp_path, oracle_mod):
    """The one OBJ the reference ships (data/teapot/teapot.obj, 15 704 triangles, 7 850 vertices, main.cpp:36): our loader's buffers
    equal the reference loader's byte for byte, and the positions are those bundled in prt_b200/data/teapot.npz before its transform."""
    import ctypes as C
    import prt_b200  # noqa: F401
    obj = "/root/reference/data/teapot/teapot.obj"
    out = subprocess.run([build_dump(), obj, str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0 and "loaded 15704 triangles, 7850 vertices" in out.stdout, out.stdout + out.stderr
    if oracle_mod.available():
        with oracle_mod.quiet_stdout():
            assert C.CDLL(oracle_mod.REF_SO).ref_load_obj_dump(obj.encode(), str(tmp_path / "ref.bin").encode()) == 0
        assert open(tmp_path / "ours.bin", "rb").read() == open(tmp_path / "ref.bin", "rb").read()
    mesh, _ = parse_dump(open(tmp_path / "ours.bin", "rb").read())
    bundled = np.load(os.path.join(ROOT, "prt_b200", "data", "teapot.npz"))
    assert np.array_equal(mesh["indices"], bundled["indices"])


def test_numbers_are_read_like_tinyobj_not_strtod(tmp_path, oracle_mod):
    """tinyobj's own decimal reader (tiny_obj_loader.h:866-1000) rounds differently from strtod on float ties: the first twelve
    coordinates below come out one float apart under strtod.  Plus the token forms it accepts or zeroes.  Golden: loader_num.bin."""
    import prt_b200  # noqa: F401
    ties = ["-66.5789875e5", "-10.3264260e7", "-90.0579232e7", "-0.5530217e8", "8.7553060e7", "5.028113400e7", "2.2976151e7", "-12.4969492e7",
            "83.3516768e7", "-14.3350665e6", "-82.0288280e6", "-14.9236965e6"]
    odd = [".5", "-.25", "+3", "1.5abc", "1e", "nan", "7.", "1e+2", "2E-3", "0.1234567890123456789", "123456789012345678", "1e400", "-0", "1.e1", "e5"]
    toks = ties + odd
    while len(toks) % 3: toks.append("0.25")
    L = ["mtllib m.mtl", "usemtl a"] + [f"v {toks[i]} {toks[i + 1]} {toks[i + 2]}" for i in range(0, len(toks), 3)]
    nvert = len(toks) // 3
    L += [f"vt {toks[(2 * i) % len(toks)]} {toks[(2 * i + 1) % len(toks)]}" for i in range(nvert)]
    L += [f"f {i + 1}/{i + 1} {(i + 1) % nvert + 1}/{(i + 1) % nvert + 1} {(i + 2) % nvert + 1}/{(i + 2) % nvert + 1}" for i in range(nvert)]
    open(tmp_path / "m.mtl", "w").write("newmtl a\nKd 0.1234567890123 5.028113400e-1 .75\nKe 8.7553060e-7 1e-2 3\n")
    open(tmp_path / "n.obj", "w").write("\n".join(L) + "\n")
    gold_path = os.path.join(ROOT, "tests", "golden", "loader_num.bin")
    out = subprocess.run([build_dump(), str(tmp_path / "n.obj"), str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ours = open(tmp_path / "ours.bin", "rb").read()
    if oracle_mod.available():
        import ctypes as C
        with oracle_mod.quiet_stdout():
            r = C.CDLL(oracle_mod.REF_SO).ref_load_obj_dump(str(tmp_path / "n.obj").encode(), str(tmp_path / "ref.bin").encode())
        assert r == 0
        ref = open(tmp_path / "ref.bin", "rb").read()
        if os.environ.get("PRT_WRITE_GOLDEN"):
            open(gold_path, "wb").write(ref)
        assert ours == ref, "number parsing differs from the live reference"
    assert ours == open(gold_path, "rb").read(), "number parsing differs from tests/golden/loader_num.bin (reference output)"
    mesh, _ = parse_dump(ours)
    strtod = np.array([np.float32(float(t)) for t in ties], np.float32)
    got = mesh["positions"].reshape(-1)[:len(ties)]
    assert int((got != strtod).sum()) >= 10                   # the reference really is off the correctly rounded float here


def test_polygons_with_more_than_four_corners_match_reference(tmp_path, oracle_mod):
    """tinyobj's ear clipping (tiny_obj_loader.h:1536-1820) restated: convex, star-shaped, collinear-corner and self-crossing polygons of
    5-10 corners in random planes, with vt per corner.  Golden: tests/golden/loader_ngon.bin (the reference's dump)."""
    import prt_b200  # noqa: F401
    rng = np.random.RandomState(77)
    V, VT, F = [], [], []
    for k in range(60):
        n = rng.randint(5, 11)
        ang = np.sort(rng.uniform(0, 2 * np.pi, n))
        kind = k % 4
        rad = np.ones(n) if kind == 0 else rng.uniform(0.3, 1.0, n)          # convex / star-shaped
        p2 = np.stack([rad * np.cos(ang), rad * np.sin(ang)], 1)
        if kind == 2: p2[1] = 0.5 * (p2[0] + p2[2])                           # a collinear corner
        if kind == 3: p2[[1, 3]] = p2[[3, 1]]                                 # self-crossing: the walk may give up
        q, _ = np.linalg.qr(rng.randn(3, 3))
        p3 = (np.c_[p2, np.zeros(n)] @ q.T) * rng.uniform(0.5, 3) + rng.uniform(-2, 2, 3)
        if k % 5 == 0: p3 = np.c_[p2, np.full(n, 0.25)][:, rng.permutation(3)]  # axis-aligned planes: ties in the axis choice
        base = len(V)
        V += [tuple(np.float32(x) for x in row) for row in p3]
        VT += [tuple(np.float32(x) for x in row) for row in rng.uniform(0, 1, (n, 2))]
        order = range(n) if k % 2 == 0 else range(n - 1, -1, -1)            # both windings
        F.append(" ".join(f"{base + i + 1}/{base + i + 1}" for i in order))
    L = ["mtllib m.mtl", "usemtl a"] + [f"v {x:.9g} {y:.9g} {z:.9g}" for x, y, z in V] + [f"vt {u:.9g} {v:.9g}" for u, v in VT] + ["f " + f for f in F]
    open(tmp_path / "m.mtl", "w").write("newmtl a\nKd 1 0 0\n")
    open(tmp_path / "n.obj", "w").write("\n".join(L) + "\n")
    gold_path = os.path.join(ROOT, "tests", "golden", "loader_ngon.bin")
    out = subprocess.run([build_dump(), str(tmp_path / "n.obj"), str(tmp_path / "ours.bin")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    ours = open(tmp_path / "ours.bin", "rb").read()
    if oracle_mod.available():
        import ctypes as C
        with oracle_mod.quiet_stdout():
            r = C.CDLL(oracle_mod.REF_SO).ref_load_obj_dump(str(tmp_path / "n.obj").encode(), str(tmp_path / "ref.bin").encode())
        assert r == 0
        ref = open(tmp_path / "ref.bin", "rb").read()
        if os.environ.get("PRT_WRITE_GOLDEN"):
            open(gold_path, "wb").write(ref)
        assert ours == ref, f"n-gon dump differs from the live reference ({len(ours)} vs {len(ref)} bytes)"
    assert ours == open(gold_path, "rb").read(), "n-gon dump differs from tests/golden/loader_ngon.bin (reference output)"
    mesh, _ = parse_dump(ours)
    assert 150 < len(mesh["indices"]) < 60 * 8


def test_loader_rejects_what_it_does_not_restate(tmp_path):
    import prt_b200  # noqa: F401
    open(tmp_path / "m.mtl", "w").write("newmtl a\nKd 1 0 0\n")
    open(tmp_path / "ok.obj", "w").write("mtllib m.mtl\nusemtl a\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0.5\nf 1 2 3 4\n")
    out = subprocess.run([build_dump(), str(tmp_path / "ok.obj"), str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 0 and "loaded 2 triangles, 4 vertices, 1 materials" in out.stdout and "scene.add: ok" in out.stdout
    out = subprocess.run([build_dump(), str(tmp_path / "missing.obj"), str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert out.returncode == 1 and "Failed to load" in out.stderr


def parse_dump(b):
    h = np.frombuffer(b[:12], np.uint32); nprim, nvert, nmat = (int(x) for x in h); o = 12
    idx = np.frombuffer(b[o:o + 12 * nprim], np.uint32).reshape(-1, 3); o += 12 * nprim
    pos = np.frombuffer(b[o:o + 12 * nvert], np.float32).reshape(-1, 3); o += 12 * nvert
    uv = np.frombuffer(b[o:o + 8 * nvert], np.float32).reshape(-1, 2); o += 8 * nvert
    pm = np.frombuffer(b[o:o + 4 * nprim], np.uint32); o += 4 * nprim
    mats, textures = [], []
    for _ in range(nmat):
        v = np.frombuffer(b[o:o + 24], np.float32); u = np.frombuffer(b[o + 24:o + 32], np.uint32); o += 32
        ids = []
        for _k in range(2):
            t = np.frombuffer(b[o:o + 12], np.uint32); o += 12
            n = int(t[0]) * int(t[1]) * int(t[2])
            if n:
                textures.append(dict(width=int(t[0]), height=int(t[1]), comp=int(t[2]), texels=np.frombuffer(b[o:o + n], np.uint8).copy()))
                ids.append(len(textures) - 1)
            else:
                ids.append(-1)
            o += n
        mats.append(dict(diffuse=tuple(v[:3]), emissive=tuple(v[3:6]), refl=int(u[0]), diffuse_tex=ids[0], bump_tex=ids[1]))
    return dict(indices=idx, positions=pos, normals=None, texcoords=uv, prim_material=pm, materials=mats), textures


@pytest.mark.gpu
def test_obj_on_disk_renders_like_the_same_buffers_through_the_c_abi(tmp_path):
    """Disk -> device: a textured OBJ scene (alpha-holed leaves, bumped wall, emissive lamp) loaded and rendered by the C++
    host API equals the render of the very same buffers handed to the C ABI from Python, bit for bit."""
    import prt_b200
    rng = np.random.RandomState(8)
    d = str(tmp_path)
    tex = rng.randint(0, 256, (32, 32, 4)).astype(np.uint8); tex[..., 3] = np.where(rng.rand(32, 32) < 0.3, 0, 255)
    write_tga(os.path.join(d, "leaf.tga"), tex, rle=True)
    write_tga(os.path.join(d, "wall.tga"), rng.randint(60, 256, (16, 16, 3)).astype(np.uint8))
    nrm = rng.randint(96, 160, (16, 16, 3)).astype(np.uint8); nrm[..., 2] = 230
    write_tga(os.path.join(d, "nrm.tga"), nrm)
    open(os.path.join(d, "s.mtl"), "w").write("newmtl wall\nKd 0.8 0.8 0.8\nmap_Kd wall.tga\nmap_bump nrm.tga\nnewmtl leaf\nKd 0.4 0.9 0.3\nmap_Kd leaf.tga\n"
                                              "newmtl lamp\nKd 0 0 0\nKe 8 8 6\n")
    L = ["mtllib s.mtl"]
    quads = [  # floor, back wall, a leaf card in front of it, a lamp quad overhead
        ("wall", [(-2, 0, -2), (2, 0, -2), (2, 0, 2), (-2, 0, 2)]), ("wall", [(-2, 0, -2), (-2, 3, -2), (2, 3, -2), (2, 0, -2)]),
        ("leaf", [(-1, 0.2, -0.5), (1, 0.2, -0.5), (1, 2.2, -0.6), (-1, 2.2, -0.6)]), ("lamp", [(-0.5, 2.9, -0.5), (0.5, 2.9, -0.5), (0.5, 2.9, 0.5), (-0.5, 2.9, 0.5)])]
    L += ["vt 0 0", "vt 2 0", "vt 2 2", "vt 0 2"]
    nv = 0
    for name, q in quads:
        L += [f"v {x} {y} {z}" for x, y, z in q] + [f"usemtl {name}", "f " + " ".join(f"{nv + k + 1}/{k + 1}" for k in range(4))]
        nv += 4
    open(os.path.join(d, "s.obj"), "w").write("\n".join(L) + "\n")
    exe = os.path.join(HOST, "build", "obj_render")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", os.path.join(HOST, "obj_render.cpp"), "-o", exe,
                           "-L" + os.path.join(ROOT, "prt_b200"), "-lprt_cuda", "-Wl,-rpath," + os.path.join(ROOT, "prt_b200")])
    W, H, spp, seed = 96, 64, 16, 3
    cam = dict(pos=(0.0, 1.5, 4.5), dir=(0.0, -0.1, -1.0)); light = dict(dir=(0.3, 0.8, 0.5), intensity=(6.0, 6.0, 5.0))
    args = [str(x) for x in (*cam["pos"], *cam["dir"], *light["dir"], *light["intensity"])]
    out = subprocess.run([exe, os.path.join(d, "s.obj"), os.path.join(d, "o.raw"), str(W), str(H), str(spp), str(seed), "0", *args], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    got = np.fromfile(os.path.join(d, "o.raw"), np.float32).reshape(H, W, 3)
    out = subprocess.run([build_dump(), os.path.join(d, "s.obj"), os.path.join(d, "dump.bin")], capture_output=True, text=True)
    assert out.returncode == 0 and "scene.add: ok" in out.stdout, out.stdout + out.stderr
    mesh, textures = parse_dump(open(os.path.join(d, "dump.bin"), "rb").read())
    desc = dict(meshes=[mesh], textures=textures, light=dict(type="directional", **light), camera=dict(width=W, height=H, **cam))
    ctx = prt_b200.load_scene(desc, 0)
    want = ctx.render(spp, seed)
    assert ctx.alpha_flags[0][1] == 1 and got.max() > 0
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    ctx.close()


@pytest.mark.gpu
def test_png_exr_scene_on_disk_renders_like_c_abi(tmp_path):
    """Disk -> device with PNG textures and an .exr sky (Scene::setInfiniteAreaLight(path)): the C++ host render equals the render
    of the same decoded buffers handed to the C ABI from Python, bit for bit."""
    import prt_b200
    rng = np.random.RandomState(12)
    d = str(tmp_path)
    tex = rng.randint(0, 256, (32, 32, 4)).astype(np.uint8); tex[..., 3] = np.where(rng.rand(32, 32) < 0.3, 0, 255)
    write_png(os.path.join(d, "leaf.png"), tex, 6)
    write_png(os.path.join(d, "wall.png"), rng.randint(60, 256, (16, 16, 3)).astype(np.uint8), 2)
    nrm = rng.randint(96, 160, (16, 16, 3)).astype(np.uint8); nrm[..., 2] = 230
    write_png(os.path.join(d, "nrm.png"), nrm, 2)
    sky = np.zeros((16, 32, 4), np.float32); sky[..., :3] = 0.4 + rng.rand(16, 32, 3) * 0.2; sky[2:5, 6:10, :3] = 40.0; sky[..., 3] = 1.0
    write_exr(os.path.join(d, "sky.exr"), sky, ["R", "G", "B", "A"], 2)
    open(os.path.join(d, "s.mtl"), "w").write("newmtl wall\nKd 0.8 0.8 0.8\nmap_Kd wall.png\nmap_bump nrm.png\nnewmtl leaf\nKd 0.4 0.9 0.3\nmap_Kd leaf.png\n")
    L = ["mtllib s.mtl", "vt 0 0", "vt 2 0", "vt 2 2", "vt 0 2"]
    quads = [("wall", [(-2, 0, -2), (2, 0, -2), (2, 0, 2), (-2, 0, 2)]), ("wall", [(-2, 0, -2), (-2, 3, -2), (2, 3, -2), (2, 0, -2)]),
             ("leaf", [(-1, 0.2, -0.5), (1, 0.2, -0.5), (1, 2.2, -0.6), (-1, 2.2, -0.6)])]
    nv = 0
    for name, q in quads:
        L += [f"v {x} {y} {z}" for x, y, z in q] + [f"usemtl {name}", "f " + " ".join(f"{nv + k + 1}/{k + 1}" for k in range(4))]
        nv += 4
    open(os.path.join(d, "s.obj"), "w").write("\n".join(L) + "\n")
    exe = os.path.join(HOST, "build", "obj_render")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", os.path.join(HOST, "obj_render.cpp"), "-o", exe,
                           "-L" + os.path.join(ROOT, "prt_b200"), "-lprt_cuda", "-Wl,-rpath," + os.path.join(ROOT, "prt_b200")])
    W, H, spp, seed = 96, 64, 16, 4
    cam = dict(pos=(0.0, 1.5, 4.5), dir=(0.0, -0.1, -1.0))
    args = [str(x) for x in (*cam["pos"], *cam["dir"], 0, 1, 0, 1, 1, 1)]
    out = subprocess.run([exe, os.path.join(d, "s.obj"), os.path.join(d, "o.raw"), str(W), str(H), str(spp), str(seed), "0", *args, os.path.join(d, "sky.exr")],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    got = np.fromfile(os.path.join(d, "o.raw"), np.float32).reshape(H, W, 3)
    out = subprocess.run([build_dump(), os.path.join(d, "s.obj"), os.path.join(d, "dump.bin")], capture_output=True, text=True)
    assert out.returncode == 0 and "scene.add: ok" in out.stdout, out.stdout + out.stderr
    mesh, textures = parse_dump(open(os.path.join(d, "dump.bin"), "rb").read())
    desc = dict(meshes=[mesh], textures=textures, light=dict(type="infinite", image=sky), camera=dict(width=W, height=H, **cam))
    ctx = prt_b200.load_scene(desc, 0)
    want = ctx.render(spp, seed)
    assert got.max() > 0 and np.array_equal(got.view(np.uint32), want.view(np.uint32))
    ctx.close()
