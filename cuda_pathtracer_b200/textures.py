"""Image files -> the float arrays pt_texture_create / pt_set_blue_noise / pt_set_skydome take: the host half of the
reference's loadTexture / loadTextureL / loadTextureHDR (src/use_cuda.h:108-263).  The reference decodes with stb_image
(vendored third-party, src/stb_image.h v2.22) and then applies its own conventions; this module decodes with Pillow (PNG,
JPEG) or its own Radiance-RGBE reader (.hdr) and applies the same conventions, so that a scene's MTL textures, the
blue-noise tile and the sky dome give the product the arrays the reference's texture objects hold:

  loadTexture     8-bit RGB(A) * (1.0f/255.0f), rows flipped, a = 1           -> (h, w, 4) float32, linear filter, wrap
  loadTextureL    stbi_loadf(.., 1) with gamma 1: grey = (77 r + 150 g + 29 b) >> 8 on the 8-bit data, then / 255.0f,
                  rows NOT flipped                                           -> (h, w) float32, point filter, wrap
  loadTextureHDR  stbi_loadf(.., 0) with gamma 1: .hdr = mantissa * 2^(e-136); 8-bit files / 255.0f; rows flipped, a = 1

PNG is lossless, so these arrays equal the reference's bit for bit (tests/test_cpu_textures.py compares them with stb_image
itself, compiled where it lies); JPEG decoders differ in their IDCT / chroma upsampling, there the test bounds the difference.
Device work (texture objects, filtering) is pt_texture_create's; nothing here touches the GPU."""
import os

import numpy as np


def _decode_8bit(path):
    """(h, w, c) uint8 with the channel count stb_image reports for the file (1, 2, 3 or 4)."""
    from PIL import Image
    im = Image.open(path)
    im.load()
    if im.mode == "P":                                   # stb expands palettes to RGB (RGBA with a tRNS chunk)
        im = im.convert("RGBA" if "transparency" in im.info else "RGB")
    elif im.mode in ("1",):
        im = im.convert("L")
    elif im.mode in ("I;16", "I;16B", "I;16L", "I"):      # 16-bit grey: stb keeps the high byte
        a = np.asarray(im).astype(np.uint32)
        return (a >> 8).astype(np.uint8)[..., None]
    elif im.mode == "CMYK":
        im = im.convert("RGB")
    elif im.mode not in ("L", "LA", "RGB", "RGBA"):
        im = im.convert("RGBA" if "A" in im.mode else "RGB")
    a = np.asarray(im)
    if a.dtype != np.uint8:                              # 16-bit RGB(A)
        a = (a.astype(np.uint32) >> 8).astype(np.uint8)
    return a[..., None] if a.ndim == 2 else a


def _grey(a):
    """stbi__compute_y on 8-bit data: what stb does when one channel is requested from a colour file."""
    c = a.shape[2]
    if c <= 2:
        return a[..., 0]
    r, g, b = (a[..., k].astype(np.uint32) for k in range(3))
    return ((r * 77 + g * 150 + b * 29) >> 8).astype(np.uint8)


def load_texture(path):
    """loadTexture (use_cuda.h:108-164): (h, w, 4) float32, rows flipped, rgb = byte * (1.0f / 255.0f), a = 1."""
    a = _decode_8bit(path)
    if a.shape[2] < 3:
        raise ValueError("%s: %d channel(s); the reference asserts nrChannels >= 3 (use_cuda.h:120)" % (path, a.shape[2]))
    h, w = a.shape[:2]
    out = np.empty((h, w, 4), np.float32)
    out[..., :3] = a[::-1, :, :3].astype(np.float32) * np.float32(1.0 / 255.0)
    out[..., 3] = 1.0
    return out


def load_texture_l(path):
    """loadTextureL (use_cuda.h:166-200): one channel, (h, w) float32 = grey / 255.0f, rows as stored."""
    if path.lower().endswith(".hdr"):                     # stbi__hdr_convert with req_comp 1: (r + g + b) * 2^(e-136) / 3
        rgbe = _rgbe_bytes(path)
        e = rgbe[..., 3].astype(np.int32)
        s = rgbe[..., :3].astype(np.int32).sum(axis=2).astype(np.float32)
        out = s * np.ldexp(np.float32(1.0), e - 136).astype(np.float32) / np.float32(3)
        out[e == 0] = 0.0
        return out.astype(np.float32)
    return _grey(_decode_8bit(path)).astype(np.float32) / np.float32(255.0)


def load_texture_hdr(path):
    """loadTextureHDR (use_cuda.h:202-263): (h, w, 4) float32 radiance, rows flipped, a = 1; three channels required."""
    if path.lower().endswith(".hdr"):
        rgb = read_rgbe(path)
    else:
        a = _decode_8bit(path)
        if a.shape[2] != 3:
            raise ValueError("%s: %d channels; the reference asserts nrChannels == 3 (use_cuda.h:214)" % (path, a.shape[2]))
        rgb = a.astype(np.float32) / np.float32(255.0)    # stbi__ldr_to_hdr with gamma 1, scale 1
    h, w = rgb.shape[:2]
    out = np.empty((h, w, 4), np.float32)
    out[..., :3] = rgb[::-1]
    out[..., 3] = 1.0
    return out


def read_rgbe(path):
    """Radiance .hdr (RGBE), flat or new-style run-length encoded scanlines, -Y +X orientation: (h, w, 3) float32 with
    stb_image's conversion (stbi__hdr_convert): rgb = mantissa * 2^(e - 136), all zero when e == 0."""
    rgbe = _rgbe_bytes(path)
    e = rgbe[..., 3].astype(np.int32)
    scale = np.ldexp(np.float32(1.0), e - 136).astype(np.float32)
    out = rgbe[..., :3].astype(np.float32) * scale[..., None]
    out[e == 0] = 0.0
    return out


def _rgbe_bytes(path):
    """(h, w, 4) uint8 mantissas + shared exponent of a Radiance .hdr file."""
    with open(path, "rb") as f:
        data = f.read()
    pos = 0

    def line():
        nonlocal pos
        end = data.index(b"\n", pos)
        s = data[pos:end]
        pos = end + 1
        return s
    magic = line()
    if magic not in (b"#?RADIANCE", b"#?RGBE"):
        raise ValueError("%s: not a Radiance HDR file" % path)
    fmt_ok = False
    while True:
        s = line()
        if not s:
            break
        if s == b"FORMAT=32-bit_rle_rgbe":
            fmt_ok = True
    if not fmt_ok:
        raise ValueError("%s: unsupported HDR format" % path)
    res = line().split()
    if len(res) != 4 or res[0] != b"-Y" or res[2] != b"+X":
        raise ValueError("%s: unsupported HDR orientation" % path)
    h, w = int(res[1]), int(res[3])
    buf = np.frombuffer(data, np.uint8)
    rgbe = np.empty((h, w, 4), np.uint8)
    if w < 8 or w >= 32768:
        rgbe[:] = buf[pos:pos + h * w * 4].reshape(h, w, 4)
    else:
        for y in range(h):
            if not (buf[pos] == 2 and buf[pos + 1] == 2 and not (buf[pos + 2] & 0x80)):
                if y != 0:
                    raise ValueError("%s: mixed flat / RLE scanlines" % path)
                rgbe[:] = buf[pos:pos + h * w * 4].reshape(h, w, 4)      # flat file
                break
            if (int(buf[pos + 2]) << 8 | int(buf[pos + 3])) != w:
                raise ValueError("%s: corrupt scanline width" % path)
            pos += 4
            for k in range(4):
                x = 0
                while x < w:
                    count = int(buf[pos]); pos += 1
                    if count > 128:
                        count -= 128
                        rgbe[y, x:x + count, k] = buf[pos]; pos += 1
                    else:
                        rgbe[y, x:x + count, k] = buf[pos:pos + count]; pos += count
                    x += count
    return rgbe


def load_material_textures(scene, base_dir="."):
    """What Scene::addModel does with an MTL's map_Kd / norm entries (scene.h:218-243): returns
    ({material id: image}, {material id: normal map}) for Pathtracer.Init(textures=..., normal_maps=...).  `scene` is a
    cuda_pathtracer_b200.scene.Scene; file names are resolved against base_dir (the reference resolves them against the
    working directory)."""
    tex, nrm = {}, {}
    n = len(scene.arrays()["materials"])
    cache = {}
    for mid in range(n):
        for normal_map, dst in ((False, tex), (True, nrm)):
            name = scene.textureName(mid, normal_map)
            if not name:
                continue
            p = name if os.path.isabs(name) else os.path.join(base_dir, name)
            if p not in cache:
                cache[p] = load_texture(p)
            dst[mid] = cache[p]
    return tex, nrm
