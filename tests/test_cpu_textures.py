"""CPU: cuda_pathtracer_b200/textures.py (the image-file half of loadTexture / loadTextureL / loadTextureHDR,
reference src/use_cuda.h:108-263) against the reference's decoder, stb_image, compiled where it lies (oracle/stb_check.cpp ->
oracle/_ref/libstbref.so, `ref` tests: build container only), and against small files written by the test itself
(run everywhere)."""
import ctypes as C
import glob
import os
import struct

import numpy as np
import pytest

from cuda_pathtracer_b200 import textures

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
STB = os.path.join(ROOT, "oracle", "_ref", "libstbref.so")
needs_ref = pytest.mark.skipif(not (os.path.exists(STB) and os.path.isdir(REF)), reason="oracle/_ref/libstbref.so + /root/reference (build container)")


def stb():
    L = C.CDLL(STB)
    L.ref_stbi_load.restype = C.POINTER(C.c_ubyte)
    L.ref_stbi_load.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    L.ref_stbi_loadf.restype = C.POINTER(C.c_float)
    L.ref_stbi_loadf.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int, C.c_float]
    L.ref_stbi_free.argtypes = [C.c_void_p]
    return L


def stb_load(path, req=0):
    L = stb(); w, h, c = C.c_int(), C.c_int(), C.c_int()
    p = L.ref_stbi_load(path.encode(), C.byref(w), C.byref(h), C.byref(c), req)
    assert p, path
    n = req or c.value
    a = np.ctypeslib.as_array(p, shape=(h.value, w.value, n)).copy()
    L.ref_stbi_free(p)
    return a, c.value


def stb_loadf(path, req=0):
    L = stb(); w, h, c = C.c_int(), C.c_int(), C.c_int()
    p = L.ref_stbi_loadf(path.encode(), C.byref(w), C.byref(h), C.byref(c), req, 1.0)
    assert p, path
    n = req or c.value
    a = np.ctypeslib.as_array(p, shape=(h.value, w.value, n)).copy()
    L.ref_stbi_free(p)
    return a, c.value


def ref_load_texture(path):
    """The loop of loadTexture (use_cuda.h:124-137) on stb's bytes."""
    d, c = stb_load(path)
    assert c >= 3
    out = np.empty(d.shape[:2] + (4,), np.float32)
    out[..., :3] = d[::-1, :, :3].astype(np.float32) * np.float32(1.0 / 255.0)
    out[..., 3] = 1
    return out


PNGS = sorted(p for p in glob.glob(os.path.join(REF, "*.png")) + glob.glob(os.path.join(REF, "textures", "*.png"))
              if os.path.getsize(p) < 6 << 20)[:14]


@needs_ref
@pytest.mark.ref
def test_png_textures_equal_the_reference_loader_bitwise():
    checked = 0
    for p in PNGS:
        d, c = stb_load(p)
        if c < 3:
            with pytest.raises(ValueError):
                textures.load_texture(p)
            continue
        assert np.array_equal(textures.load_texture(p).view(np.uint32), ref_load_texture(p).view(np.uint32)), p
        checked += 1
    assert checked >= 5


@needs_ref
@pytest.mark.ref
def test_blue_noise_tile_equals_loadTextureL_bitwise():
    p = os.path.join(REF, "bluenoise.png")
    d, c = stb_loadf(p, 1)                                  # loadTextureL: stbi_loadf(.., 1), gamma 1, rows not flipped
    mine = textures.load_texture_l(p)
    assert mine.shape == d.shape[:2] and np.array_equal(mine.view(np.uint32), d[..., 0].view(np.uint32))
    q = os.path.join(REF, "vokselia_spawn.png")             # a colour file through the one-channel request: stbi__compute_y
    d, c = stb_loadf(q, 1)
    assert np.array_equal(textures.load_texture_l(q).view(np.uint32), d[..., 0].view(np.uint32))


@needs_ref
@pytest.mark.ref
def test_jpeg_sky_dome_within_decoder_tolerance():
    """skydome.jpg through loadTextureHDR's path (stbi_loadf, gamma 1, rows flipped, a = 1).  stb_image and libjpeg differ in
    IDCT rounding and chroma upsampling: bounded, not exact."""
    p = os.path.join(REF, "skydome.jpg")
    d, c = stb_loadf(p, 0)
    assert c == 3
    mine = textures.load_texture_hdr(p)
    assert mine.shape == d.shape[:2] + (4,) and np.all(mine[..., 3] == 1)
    diff = np.abs(mine[..., :3] - d[::-1])
    assert diff.mean() < 1.5 / 255 and np.percentile(diff, 99.9) < 12 / 255


def write_hdr(path, rgbe, rle):
    h, w = rgbe.shape[:2]
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w))
        if not rle:
            f.write(rgbe.tobytes()); return
        for y in range(h):
            f.write(bytes([2, 2, w >> 8, w & 255]))
            for k in range(4):
                row = rgbe[y, :, k]; x = 0
                while x < w:
                    run = 1
                    while x + run < w and run < 127 and row[x + run] == row[x]:
                        run += 1
                    if run >= 3:
                        f.write(bytes([128 + run, int(row[x])])); x += run
                    else:
                        n = min(w - x, 16)
                        f.write(bytes([n]) + row[x:x + n].tobytes()); x += n


@pytest.mark.parametrize("rle", [False, True])
def test_rgbe_reader_known_values(tmp_path, rle):
    rng = np.random.default_rng(5)
    rgbe = rng.integers(0, 256, (9, 40, 4), dtype=np.uint8)
    rgbe[:, 5:25, 0] = 77                                    # runs for the RLE writer
    rgbe[2, :, 3] = 0                                        # exponent 0 => black
    rgbe[..., 3] = np.where(rgbe[..., 3] == 0, 0, 120 + rgbe[..., 3] % 20)
    p = str(tmp_path / "t.hdr"); write_hdr(p, rgbe, rle)
    rgb = textures.read_rgbe(p)
    want = rgbe[..., :3].astype(np.float64) * np.exp2(rgbe[..., 3].astype(np.float64) - 136)[..., None]
    want[rgbe[..., 3] == 0] = 0
    assert np.array_equal(rgb, want.astype(np.float32))
    out = textures.load_texture_hdr(p)
    assert np.array_equal(out[..., :3], rgb[::-1]) and np.all(out[..., 3] == 1)
    if os.path.exists(STB):                                  # and the same file through stb_image
        d, c = stb_loadf(p, 0)
        assert c == 3 and np.array_equal(d.view(np.uint32), rgb.view(np.uint32))
        d1, _ = stb_loadf(p, 1)
        assert np.array_equal(d1[..., 0].view(np.uint32), textures.load_texture_l(p).view(np.uint32))


def test_conventions_on_a_file_written_here(tmp_path):
    """Rows flipped for loadTexture / loadTextureHDR, not for loadTextureL; rgb * (1.0f/255.0f) vs / 255.0f; alpha ignored."""
    from PIL import Image
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (6, 10, 4), dtype=np.uint8)
    p = str(tmp_path / "t.png"); Image.fromarray(a, "RGBA").save(p)
    t = textures.load_texture(p)
    assert t.dtype == np.float32 and t.shape == (6, 10, 4) and np.all(t[..., 3] == 1)
    assert np.array_equal(t[..., :3], a[::-1, :, :3].astype(np.float32) * np.float32(1.0 / 255.0))
    g = textures.load_texture_l(p)
    y = ((a[..., 0].astype(np.uint32) * 77 + a[..., 1].astype(np.uint32) * 150 + a[..., 2].astype(np.uint32) * 29) >> 8).astype(np.float32)
    assert np.array_equal(g, y / np.float32(255.0))
    with pytest.raises(ValueError):
        textures.load_texture_hdr(p)                          # four channels: the reference asserts nrChannels == 3
    q = str(tmp_path / "g.png"); Image.fromarray(a[..., 0], "L").save(q)
    with pytest.raises(ValueError):
        textures.load_texture(q)                              # one channel: the reference asserts nrChannels >= 3
    assert np.array_equal(textures.load_texture_l(q), a[..., 0].astype(np.float32) / np.float32(255.0))


def write_textured_quad(tmp_path, with_normal_map=True):
    """A 4 x 4 quad in the xy plane with uvs, an MTL naming a map_Kd (and norm) PNG, the PNGs themselves."""
    from PIL import Image
    rng = np.random.default_rng(11)
    kd = rng.integers(0, 256, (16, 16, 3), dtype=np.uint8)
    Image.fromarray(kd, "RGB").save(str(tmp_path / "kd.png"))
    nm = np.full((8, 8, 3), (128, 128, 255), np.uint8); nm[::2, :, 0] = 160
    Image.fromarray(nm, "RGB").save(str(tmp_path / "nm.png"))
    (tmp_path / "quad.mtl").write_text("newmtl painted\nKd 0.8 0.8 0.8\nKs 0 0 0\nd 1\nNs 0\nNi 1\nmap_Kd kd.png\n" + ("norm nm.png\n" if with_normal_map else "") +
                                       "newmtl plain\nKd 0.5 0.4 0.3\nKs 0 0 0\nd 1\nNs 0\nNi 1\n")
    (tmp_path / "quad.obj").write_text("mtllib quad.mtl\nv -2 -2 0\nv 2 -2 0\nv 2 2 0\nv -2 2 0\nv -6 -2 0\nv -2.5 -2 0\nv -2.5 2 0\n"
                                       "vt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nvn 0 0 -1\n"
                                       "usemtl painted\nf 1/1/1 2/2/1 3/3/1\nf 1/1/1 3/3/1 4/4/1\nusemtl plain\nf 5/1/1 6/2/1 7/3/1\n")
    return str(tmp_path / "quad.obj")


def test_mtl_texture_names_become_texture_arrays(tmp_path):
    """Scene::addModel's map_Kd / norm handling (scene.h:218-243) through the Python mirror: the MTL's file names come back
    per material and load_material_textures turns them into the arrays Pathtracer.Init(textures=, normal_maps=) uploads."""
    from cuda_pathtracer_b200.scene import Scene
    obj = write_textured_quad(tmp_path)
    s = Scene()
    m = s.addModel(obj, useMtl=True)
    s.addObject(m)
    s.finalize()
    mats = s.arrays()["materials"]
    painted = [i for i in range(len(mats)) if s.textureName(i)]
    assert len(painted) == 1 and s.textureName(painted[0]) == "kd.png" and s.textureName(painted[0], True) == "nm.png"
    assert mats[painted[0]]["hasTexture"] == 1 and mats[painted[0]]["hasNormalMap"] == 1
    tex, nrm = textures.load_material_textures(s, str(tmp_path))
    assert list(tex) == painted and list(nrm) == painted
    assert np.array_equal(tex[painted[0]], textures.load_texture(str(tmp_path / "kd.png"))) and tex[painted[0]].shape == (16, 16, 4)
    assert nrm[painted[0]].shape == (8, 8, 4)
    with pytest.raises(FileNotFoundError):
        textures.load_material_textures(s, str(tmp_path / "elsewhere"))     # the reference exits when a texture is missing (use_cuda.h:114-117)
