"""TEST / MEASUREMENT INFRASTRUCTURE.  Driver for oracle/_ref/refgpu_* -- the reference's own device path
(src/pathtracer.h + src/kernels.h, unmodified, compiled for sm_100a; see oracle/refgpu_main.cu).  A scene is described
once (SceneSpec) and realised twice: as the text file the reference harness parses into the reference's `Scene`, and
through this repo's host scene code, so both renderers see the same OBJ files, materials, instances and camera."""
import os
import struct
import subprocess
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def binary_path(kind, w, h, depth):
    return os.path.join(HERE, "_ref", "refgpu_%s_%dx%dx%d" % (kind, w, h, depth))


def write_png_gray(path, img_u8):
    """8-bit greyscale PNG (what stbi_loadf(.., 1) reads for bluenoise.png, use_cuda.h:166-200)."""
    h, w = img_u8.shape
    raw = b"".join(b"\x00" + img_u8[y].tobytes() for y in range(h))
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 0, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b""))


def write_hdr_constant(path, w, h, value=1.0):
    """Radiance .hdr with flat (non-RLE) RGBE pixels of one grey value (cave.hdr for loadTextureHDR, use_cuda.h:202-263)."""
    m, e = np.frexp(value)                       # value = m * 2^e, m in [0.5, 1)
    px = bytes([int(m * 256), int(m * 256), int(m * 256), int(e + 128)])
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w) + px * (w * h))


class SceneSpec:
    def __init__(self):
        self.materials, self.models, self.objects, self.spheres, self.planes = [], [], [], [], []
        self.camera = None
        self.interactive_depth = 1

    def material(self, diffuse, emission=(0, 0, 0), reflect=0.0, glossy=0.0, transmit=0.0, ior=1.0, absorption=(0, 0, 0)):
        self.materials.append(dict(diffuse=diffuse, emission=emission, reflect=reflect, glossy=glossy, transmit=transmit, ior=ior, absorption=absorption))
        return len(self.materials) - 1

    def model(self, obj_path, material, scale=1.0, use_mtl=False):
        self.models.append((obj_path, scale, material, use_mtl)); return len(self.models) - 1

    def object(self, model, material=-1, position=(0, 0, 0), rotation=(0, 0, 0), scale=(1, 1, 1)):
        self.objects.append((model, material, position, rotation, scale))

    def write(self, path, clear_frames, converge_frames, nee, cache, out, timed_frames=0):
        L = []
        for m in self.materials:
            L.append("material " + " ".join(repr(float(x)) for x in (*m["diffuse"], *m["emission"], m["reflect"], m["glossy"], m["transmit"], m["ior"], *m["absorption"])))
        for p, s, mat, use_mtl in self.models:
            L.append("model %s %r 0 0 0 0 0 0 %d %d" % (p, float(s), mat, int(use_mtl)))
        for mo, mat, pos, rot, sc in self.objects:
            L.append("object %d %d " % (mo, mat) + " ".join(repr(float(x)) for x in (*pos, *rot, *sc)))
        for pos, r, mat in self.spheres:
            L.append("sphere %r %r %r %r %d" % (*map(float, pos), float(r), mat))
        for n, d, mat in self.planes:
            L.append("plane %r %r %r %r %d" % (*map(float, n), float(d), mat))
        eye, dr, d, focal, ap = self.camera
        L.append("camera " + " ".join(repr(float(x)) for x in (*eye, *dr, d, focal, ap)))
        L.append("interactive_depth %d" % self.interactive_depth)
        L.append("render %d %d %d %d %d" % (clear_frames, converge_frames, int(nee), int(cache), timed_frames))
        L.append("out %s" % out)
        open(path, "w").write("\n".join(L) + "\n")

    def build_ours(self):
        """The same scene through this repo's host scene code (pts_* C++).  Returns (Scene, camera factory)."""
        from cuda_pathtracer_b200.scene import Scene, material
        s = Scene()
        s.interactive_depth = self.interactive_depth
        for m in self.materials:
            s.addMaterial(material(m["diffuse"], emission=m["emission"], reflect=m["reflect"], glossy=m["glossy"], transmit=m["transmit"],
                                   refractive_index=m["ior"], absorption=m["absorption"]))
        ids = [s.addModel(p, sc, (0, 0, 0), (0, 0, 0), mat, use_mtl) for p, sc, mat, use_mtl in self.models]
        for mo, mat, pos, rot, sc in self.objects:
            s.addObject(ids[mo], material=0xFFFFFFFF if mat < 0 else mat, position=pos, rotation=rot, scale=sc)
        for pos, r, mat in self.spheres:
            s.addSphere(pos, r, mat)
        for n, d, mat in self.planes:
            s.addPlane(n, d, mat)
        return s.finalize()

    def camera_ours(self, w, h):
        from cuda_pathtracer_b200.scene import make_camera
        eye, dr, d, focal, ap = self.camera
        return make_camera(eye, dr, d, focal, ap, w, h)


def run_reference(binary, spec, workdir, w, h, clear_frames, converge_frames, nee, cache, blue_u8, sky_value=1.0, timed_frames=0, timeout=600):
    """Runs the reference device build in `workdir` (it loads ./bluenoise.png and ./cave.hdr).  Returns
    (luminance (h,w,4), albedo (h,w,4), ms_per_sample or None, stdout)."""
    write_png_gray(os.path.join(workdir, "bluenoise.png"), blue_u8)
    write_hdr_constant(os.path.join(workdir, "cave.hdr"), 64, 32, sky_value)
    scene_txt = os.path.join(workdir, "scene.txt")
    out = os.path.join(workdir, "ref_out")
    spec.write(scene_txt, clear_frames, converge_frames, nee, cache, out, timed_frames)
    r = subprocess.run([binary, scene_txt], cwd=workdir, capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError("reference device build failed (%d): %s" % (r.returncode, (r.stdout + r.stderr)[-2000:]))
    data = np.fromfile(out + ".bin", np.float32).reshape(2, h, w, 4)
    ms = None
    for line in r.stdout.splitlines():
        if line.startswith("REFGPU timed_frames"):
            ms = float(line.split()[-1])
    return data[0], data[1], ms, r.stdout


def write_obj(path, tri):
    with open(path, "w") as f:
        v = tri.reshape(-1, 3)
        f.write("".join("v %r %r %r\n" % (float(a), float(b), float(c)) for a, b, c in v))
        f.write("".join("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3) for i in range(len(tri))))


def sibenik_spec(workdir):
    """The bench workload (cuda_pathtracer_b200.scenes.sibenik_scene: getSibenikScene, sceneBuilder.h:119-218, with the
    seeded proxies) as a SceneSpec over OBJ files written into `workdir`: same materials (same ids), instances, spheres."""
    from cuda_pathtracer_b200 import scenes
    sp = SceneSpec()
    glass_emit = dict(transmit=1.0, ior=1.1, glossy=0.02, absorption=(0.1, 0.5, 0.8))
    sp.material((0.4, 0.4, 0.4))
    cube_w = sp.material((1, 1, 1), emission=(10, 10, 10), **glass_emit)
    for e in ((3, 0, 0), (0, 3, 0), (0, 0, 3)):
        sp.material((1, 1, 1), emission=e, **glass_emit)
    sib_mat = sp.material((0.2, 0.2, 0.2))
    sp.material((1, 1, 1), reflect=0.6, glossy=0.08)
    lucy_mat = sp.material((0.98, 0.745, 0.02), reflect=0.7, glossy=0.08)
    sp.material((1, 1, 1), transmit=1.0, ior=1.544, absorption=(0.01, 0.4, 0.4))
    white_glass = sp.material((1, 1, 1), transmit=1.0, ior=1.5)
    mirror = sp.material((1, 1, 1), reflect=1.0, ior=1.4)
    paths = {}
    for name, soup in (("cathedral", scenes.cathedral_soup(1)), ("statue", scenes.statue_soup(4)), ("cube", scenes.cube_soup())):
        paths[name] = os.path.join(workdir, name + ".obj")
        write_obj(paths[name], soup)
    sib = sp.model(paths["cathedral"], sib_mat)
    lucy = sp.model(paths["statue"], lucy_mat)
    cube = sp.model(paths["cube"], cube_w)
    sp.object(sib, position=(0, 12, 0))
    sp.object(lucy)
    sp.object(cube, material=cube_w, position=(0, 3, 0))
    sp.spheres.append(((-2, -1, -3), 2, white_glass))
    sp.spheres.append(((-2, -1, 3), 2, mirror))
    return sp


def outside_spec(workdir):
    """BASELINE.json configs[1] (getOutsideScene, sceneBuilder.h:15-117; cuda_pathtracer_b200.scenes.outside_scene): 10 glass cubes
    on a circle (handler positions at t = 0, sceneBuilder.h:89-100) over the checker plane, as a SceneSpec."""
    from cuda_pathtracer_b200 import scenes
    sp = SceneSpec()
    sp.interactive_depth = 5
    white = sp.material((0.4, 0.4, 0.4))
    cube_mat = sp.material((1, 1, 1), transmit=1.0, ior=1.1, glossy=0.02, absorption=(0.1, 0.5, 0.8))
    path = os.path.join(workdir, "cube.obj")
    write_obj(path, scenes.cube_soup())
    cube = sp.model(path, cube_mat)
    f = np.float32(0)
    for _ in range(10):
        sp.object(cube, position=(float(np.float32(10) * np.sin(f, dtype=np.float32)), 0.0, float(np.float32(10) * np.cos(f, dtype=np.float32))), rotation=(float(f), 0.0, 0.0))
        f = np.float32(f + np.float32(2 * 0.3141592))
    sp.planes.append(((0, -1, 0), -3.0, white))
    return sp


def write_obj_uv(path, tri, uv, mtllib, usemtl):
    """(n, 9) triangles + (n, 6) uvs -> OBJ with v / vt / f v/vt lines, one material."""
    with open(path, "w") as f:
        f.write("mtllib %s\n" % mtllib)
        f.write("".join("v %r %r %r\n" % (float(a), float(b), float(c)) for a, b, c in tri.reshape(-1, 3)))
        f.write("".join("vt %r %r\n" % (float(a), float(b)) for a, b in uv.reshape(-1, 2)))
        f.write("usemtl %s\n" % usemtl)
        f.write("".join("f %d/%d %d/%d %d/%d\n" % (3 * i + 1, 3 * i + 1, 3 * i + 2, 3 * i + 2, 3 * i + 3, 3 * i + 3) for i in range(len(tri))))


def minecraft_spec(workdir, n=256):
    """BASELINE.json configs[4] (getMinecraftScene, sceneBuilder.h:220-239) with the voxel-terrain proxy: an OBJ with uvs, an MTL
    whose map_Kd names the atlas PNG written next to it (both renderers decode the same file), the bokeh camera of
    scenes.minecraft_camera."""
    from PIL import Image
    from cuda_pathtracer_b200 import scenes
    tri, uv = scenes.voxel_terrain_soup(n, 3)
    atlas = np.clip(np.round(scenes.voxel_atlas(256)[..., :3] * 255.0), 0, 255).astype(np.uint8)
    Image.fromarray(atlas, "RGB").save(os.path.join(workdir, "atlas.png"))
    open(os.path.join(workdir, "terrain.mtl"), "w").write("newmtl terrain\nKd 1 1 1\nKs 0 0 0\nd 1\nNs 0\nNi 1\nmap_Kd atlas.png\n")
    path = os.path.join(workdir, "terrain.obj")
    write_obj_uv(path, tri, uv, "terrain.mtl", "terrain")
    sp = SceneSpec()
    white = sp.material((0.4, 0.4, 0.4))
    m = sp.model(path, white, use_mtl=True)
    sp.object(m)
    return sp
