"""Preset scenes (reference src/sceneBuilder.h:15-261) and the seeded procedural proxies that stand in for
the meshes missing from the reference mount (sibenik.obj, lucy.obj, 2Mtris.obj, vokselia_spawn.obj; see
SURVEY.md 8(d)).  When ``$PT_ASSET_DIR`` holds the real OBJ file it is loaded instead.  Proxies are pure
functions of their seed, so the CPU oracle and the GPU path always see identical arrays.
"""
import math
import os

import numpy as np

from .scene import Scene, make_camera, material

ASSET_DIR = os.environ.get("PT_ASSET_DIR", "")


def _asset(name):
    p = os.path.join(ASSET_DIR, name) if ASSET_DIR else ""
    return p if p and os.path.exists(p) else None


def xorshift_uniform(n, seed):
    """n floats in [0,1) from xorshift32 (13,17,5), the reference's generator (use_cuda.h:71-78), vectorised by skipping."""
    out = np.empty(n, np.float32)
    s = np.uint32(seed)
    state = np.empty(n, np.uint32)
    x = int(s)
    for i in range(n):
        x ^= (x << 13) & 0xFFFFFFFF
        x ^= x >> 17
        x ^= (x << 5) & 0xFFFFFFFF
        state[i] = x
    out[:] = state.astype(np.float32) * np.float32(2.3283064365387e-10)
    return out


def _grid(origin, du, dv, nu, nv):
    """nu x nv quads spanning origin + s*du + t*dv -> (nu*nv*2, 9) float32 triangles (CCW about cross(du,dv))."""
    o, du, dv = (np.asarray(a, np.float64) for a in (origin, du, dv))
    s = np.linspace(0, 1, nu + 1)
    t = np.linspace(0, 1, nv + 1)
    P = o[None, None, :] + s[:, None, None] * du[None, None, :] + t[None, :, None] * dv[None, None, :]
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    t1 = np.concatenate([a, b, c], axis=-1).reshape(-1, 9)
    t2 = np.concatenate([a, c, d], axis=-1).reshape(-1, 9)
    return np.concatenate([t1, t2]).astype(np.float32)


def _param_surface(fn, nu, nv, flip=False):
    u = np.linspace(0, 1, nu + 1)
    v = np.linspace(0, 1, nv + 1)
    U, V = np.meshgrid(u, v, indexing="ij")
    P = fn(U, V)
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    if flip:
        b, d = d, b
    t1 = np.concatenate([a, b, c], axis=-1).reshape(-1, 9)
    t2 = np.concatenate([a, c, d], axis=-1).reshape(-1, 9)
    return np.concatenate([t1, t2]).astype(np.float32)


def cube_soup():
    """The 12 triangles of a 2x2x2 cube centred at the origin (same extent as the reference's cube.obj)."""
    f = []
    for axis in range(3):
        for sgn in (-1.0, 1.0):
            o = np.zeros(3)
            o[axis] = sgn
            a1, a2 = (axis + 1) % 3, (axis + 2) % 3
            du, dv = np.zeros(3), np.zeros(3)
            du[a1], dv[a2] = 2.0, 2.0
            o[a1], o[a2] = -1.0, -1.0
            if sgn < 0:
                du, dv = dv, du
            f.append(_grid(o, du, dv, 1, 1))
    return np.concatenate(f)


def cathedral_soup(seed=1):
    """Proxy for sibenik.obj (~80k triangles): a closed 40 x 20 x 15 nave, 2 x 8 columns, a ribbed barrel vault.
    Model space has the floor at y = -15 so that the reference's `sibenikObj.position.y = 12` puts it at y = -3."""
    y0 = -15.0
    parts = [
        _grid((-20, y0, -7.5), (0, 0, 15), (40, 0, 0), 60, 160),            # floor (normal +y)
        _grid((-20, y0, -7.5), (40, 0, 0), (0, 20, 0), 160, 80),            # wall z = -7.5 (normal +z)
        _grid((-20, y0, 7.5), (0, 20, 0), (40, 0, 0), 80, 160),             # wall z = +7.5 (normal -z)
        _grid((-20, y0, -7.5), (0, 20, 0), (0, 0, 15), 80, 60),             # wall x = -20 (normal +x)
        _grid((20, y0, -7.5), (0, 0, 15), (0, 20, 0), 60, 80),              # wall x = +20 (normal -x)
    ]

    def vault(U, V):                                                        # barrel vault closing the top (normal down)
        ang = np.pi * V
        return np.stack([-20 + 40 * U, y0 + 20 + 4.0 * np.sin(ang), -7.5 * np.cos(ang)], -1)
    parts.append(_param_surface(vault, 160, 48))
    for i in range(8):                                                      # columns
        for z in (-4.5, 4.5):
            cx = -17.5 + 5.0 * i

            def col(U, V, cx=cx, z=z):
                a = 2 * np.pi * U
                r = 0.55 + 0.12 * np.cos(8 * a) * (V > 0.08) * (V < 0.92) + 0.25 * ((V <= 0.08) | (V >= 0.92))
                return np.stack([cx + r * np.cos(a), y0 + 17.0 * V, z + r * np.sin(a)], -1)
            parts.append(_param_surface(col, 32, 24, flip=True))
    for i in range(9):                                                      # vault ribs
        cx = -20 + 5.0 * i

        def rib(U, V, cx=cx):
            ang = np.pi * U
            a = 2 * np.pi * V
            R = 1.0 + 0.18 * np.cos(a)
            return np.stack([cx + 0.18 * np.sin(a), y0 + 20 + (4.0 * R - 0.2) * np.sin(ang) * 0.97, -7.3 * R * np.cos(ang) * 0.97], -1)
        parts.append(_param_surface(rib, 48, 10))
    tri = np.concatenate(parts)
    jit = (xorshift_uniform(4096, seed) - np.float32(0.5)) * np.float32(2e-3)
    k = (np.abs(tri * np.float32(977.0)).astype(np.int64).sum(axis=1) % 4096)
    tri = tri + jit[k][:, None]                                             # tiny per-triangle offset: breaks exact coplanarity
    return np.ascontiguousarray(tri, np.float32)


def _fbm(nx, ny, seed, octaves=5):
    rng = np.random.default_rng(seed)
    h = np.zeros((nx, ny))
    amp = 1.0
    for o in range(octaves):
        n = 2 ** (o + 2)
        g = rng.standard_normal((n + 1, n + 1))
        x = np.linspace(0, n, nx, endpoint=False)
        y = np.linspace(0, n, ny, endpoint=False)
        xi, yi = x.astype(int), y.astype(int)
        fx, fy = x - xi, y - yi
        fx, fy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
        a = g[np.ix_(xi, yi)]
        b = g[np.ix_(xi + 1, yi)]
        c = g[np.ix_(xi, yi + 1)]
        d = g[np.ix_(xi + 1, yi + 1)]
        h += amp * ((a * (1 - fx)[:, None] + b * fx[:, None]) * (1 - fy)[None, :] + (c * (1 - fx)[:, None] + d * fx[:, None]) * fy[None, :])
        amp *= 0.5
    return h


def statue_soup(seed=4, n=100):
    """Proxy for lucy.obj: a displaced blob, 2*n*n triangles, standing on the floor near (3, -3, 4)."""
    disp = _fbm(n + 1, n + 1, seed, 4)

    def blob(U, V):
        th, ph = np.pi * V, 2 * np.pi * U
        r = 0.9 + 0.18 * disp * np.sin(th)
        return np.stack([3 + r * np.sin(th) * np.cos(ph), -1.2 + 1.9 * r * np.cos(th), 4 + r * np.sin(th) * np.sin(ph)], -1)
    return _param_surface(blob, n, n)


def heightfield_soup(nx=1004, ny=1005, seed=2):
    """Proxy for 2Mtris.obj: nx x ny displaced quads = 2,018,040 triangles for the defaults (reference: 2,018,232)."""
    h = _fbm(nx + 1, ny + 1, seed, 5)
    X, Y = np.meshgrid(np.linspace(-100, 100, nx + 1), np.linspace(-100, 100, ny + 1), indexing="ij")
    P = np.stack([X, Y, 12.0 * h], -1)
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    t1 = np.concatenate([a, b, c], axis=-1).reshape(-1, 9)
    t2 = np.concatenate([a, c, d], axis=-1).reshape(-1, 9)
    return np.ascontiguousarray(np.concatenate([t1, t2]) * 0.2, np.float32)     # addModel scale 0.2 (sceneBuilder.h:249)


# ---- presets ------------------------------------------------------------------------------------------------
def outside_scene(t=0.0):
    """getOutsideScene, sceneBuilder.h:15-117: 10 glass cubes on a circle over a checker plane."""
    s = Scene()
    s.interactive_depth, s.interactive_samples = 5, 3
    white = s.addMaterial(material((0.4, 0.4, 0.4)))
    cube_mat = s.addMaterial(material((1, 1, 1), transmit=1.0, refractive_index=1.1, glossy=0.02, absorption=(0.1, 0.5, 0.8)))
    cube = s.addModelSoup(cube_soup(), material=cube_mat)
    f = np.float32(0)
    for _ in range(10):                                                     # handler, sceneBuilder.h:89-100
        a = np.float32(f + np.float32(t) / np.float32(10.0))
        s.addObject(cube, position=(float(np.float32(10) * np.sin(a, dtype=np.float32)), 0.0, float(np.float32(10) * np.cos(a, dtype=np.float32))),
                    rotation=(float(f), 0.0, 0.0))
        f = np.float32(f + np.float32(2 * 0.3141592))
    s.addPlane((0, -1, 0), -3, white)
    return s.finalize()


def sibenik_scene():
    """getSibenikScene, sceneBuilder.h:119-218, with the cathedral / statue proxies for the missing meshes."""
    s = Scene()
    s.addMaterial(material((0.4, 0.4, 0.4)))
    glass_emit = dict(transmit=1.0, refractive_index=1.1, glossy=0.02, absorption=(0.1, 0.5, 0.8))
    cube_w = s.addMaterial(material((1, 1, 1), emission=(10, 10, 10), **glass_emit))
    for e in ((3, 0, 0), (0, 3, 0), (0, 0, 3)):
        s.addMaterial(material((1, 1, 1), emission=e, **glass_emit))
    sib_mat = s.addMaterial(material((0.2, 0.2, 0.2)))
    s.addMaterial(material((1, 1, 1), reflect=0.6, glossy=0.08))
    lucy_mat = s.addMaterial(material((0.98, 0.745, 0.02), reflect=0.7, glossy=0.08))
    s.addMaterial(material((1, 1, 1), transmit=1.0, refractive_index=1.544, absorption=(0.01, 0.4, 0.4)))
    white_glass = s.addMaterial(material((1, 1, 1), transmit=1.0, refractive_index=1.5))
    mirror = s.addMaterial(material((1, 1, 1), reflect=1.0, refractive_index=1.4))
    p = _asset("sibenik.obj")
    sib = s.addModel(p, 1.0, (0, 0, 0), (0, 0, 0), sib_mat, False) if p else s.addModelSoup(cathedral_soup(1), material=sib_mat)
    s.addObject(sib, position=(0, 12, 0))
    p = _asset("lucy.obj")
    lucy = s.addModel(p, 0.005, (-3.1415926 / 2, 0, 3.1415926 / 2), (3, 0, 4.0), lucy_mat) if p else s.addModelSoup(statue_soup(4), material=lucy_mat)
    s.addObject(lucy)
    cube = s.addModelSoup(cube_soup(), material=cube_w)
    s.addObject(cube, material=cube_w, position=(0, 3, 0))
    s.addSphere((-2, -1, -3), 2, white_glass)
    s.addSphere((-2, -1, 3), 2, mirror)
    return s.finalize()


def two_million_scene(nx=1004, ny=1005):
    """get2MillionScene, sceneBuilder.h:241-261, with the heightfield proxy."""
    s = Scene()
    white = s.addMaterial(material((0.4, 0.4, 0.4)))
    p = _asset("2Mtris.obj")
    m = s.addModel(p, 0.2, (0, 0, 0), (0, 0, 0), white, False) if p else s.addModelSoup(heightfield_soup(nx, ny, 2), material=white)
    s.addObject(m, rotation=(-3.1415926535 / 2, 0, 0))
    return s.finalize()


def _quads(a, b, c, d):
    """Quads (n,3)x4 -> (2n, 9) triangles (a,b,c),(a,c,d) and (2n, 6) unit-square uvs."""
    t = np.concatenate([np.concatenate([a, b, c], -1), np.concatenate([a, c, d], -1)])
    n = len(a)
    uv1 = np.tile(np.array([0, 0, 1, 0, 1, 1], np.float32), (n, 1))
    uv2 = np.tile(np.array([0, 0, 1, 1, 0, 1], np.float32), (n, 1))
    return t.astype(np.float32), np.concatenate([uv1, uv2])


def voxel_terrain_soup(n=256, seed=3):
    """Proxy for vokselia_spawn.obj / lost_empire.obj: an n x n field of voxel columns
    (integer heights from seeded fBm) emitted as axis-aligned quads -- one top face per column, one wall per
    height step between neighbours -- with uvs into a 16 x 16-tile atlas (tile chosen by face kind and height).
    Returns (triangles (m, 9), uvs (m, 6)); n = 256 gives about 225k triangles."""
    h = np.floor(6.0 * _fbm(n, n, seed, 4) + 8.0).astype(np.int64)
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    x0, z0, y = ii.ravel().astype(np.float64), jj.ravel().astype(np.float64), h.ravel().astype(np.float64)
    P = lambda x, yy, z: np.stack([x, yy, z], -1)
    tris, uvs, tiles = [], [], []
    t, u = _quads(P(x0, y, z0), P(x0, y, z0 + 1), P(x0 + 1, y, z0 + 1), P(x0 + 1, y, z0))        # tops (normal +y)
    tris.append(t); uvs.append(u); tiles.append(np.tile(np.where(y > 9, 2, np.where(y > 6, 0, 1)), 2))
    for axis in (0, 1):                                                                          # walls between neighbours
        ha = h[:-1, :] if axis == 0 else h[:, :-1]
        hb = h[1:, :] if axis == 0 else h[:, 1:]
        ia, ja = (ii[:-1, :], jj[:-1, :]) if axis == 0 else (ii[:, :-1], jj[:, :-1])
        m = ha != hb
        lo, hi = np.minimum(ha, hb)[m].astype(np.float64), np.maximum(ha, hb)[m].astype(np.float64)
        a_hi = (ha > hb)[m]                                                                      # wall faces the lower column
        i0, j0 = ia[m].astype(np.float64), ja[m].astype(np.float64)
        if axis == 0:
            xw = i0 + 1
            A, B, C, D = P(xw, lo, j0), P(xw, lo, j0 + 1), P(xw, hi, j0 + 1), P(xw, hi, j0)      # normal -x ... flipped below
        else:
            zw = j0 + 1
            A, B, C, D = P(i0 + 1, lo, zw), P(i0, lo, zw), P(i0, hi, zw), P(i0 + 1, hi, zw)
        flip = a_hi if axis == 0 else a_hi
        A2, B2, C2, D2 = np.where(flip[:, None], B, A), np.where(flip[:, None], A, B), np.where(flip[:, None], D, C), np.where(flip[:, None], C, D)
        t, u = _quads(A2, B2, C2, D2)
        tris.append(t); uvs.append(u); tiles.append(np.tile(np.where(hi > 9, 3, 4), 2))
    tri = np.concatenate(tris); uv = np.concatenate(uvs); tile = np.concatenate(tiles)
    tx, ty = (tile % 16).astype(np.float32), (tile // 16).astype(np.float32)
    uv = uv.reshape(-1, 3, 2)
    uv = np.stack([(tx[:, None] + 0.05 + 0.9 * uv[..., 0]) / 16.0, (ty[:, None] + 0.05 + 0.9 * uv[..., 1]) / 16.0], -1).reshape(-1, 6)
    tri = tri.reshape(-1, 3, 3) - np.array([n / 2.0, 8.0, n / 2.0])
    return np.ascontiguousarray(tri.reshape(-1, 9), np.float32), np.ascontiguousarray(uv, np.float32)


def voxel_atlas(size=256, seed=5):
    """Deterministic stand-in for vokselia_spawn.png: 16 x 16 tiles of noisy colour, float4 RGBA, a = 1
    (the layout loadTexture produces, use_cuda.h:108-164)."""
    rng = np.random.default_rng(seed)
    base = np.array([[0.25, 0.6, 0.2], [0.5, 0.35, 0.2], [0.9, 0.9, 0.95], [0.5, 0.5, 0.55], [0.45, 0.3, 0.18]], np.float32)
    img = np.ones((size, size, 4), np.float32)
    ts = size // 16
    for ty in range(16):
        for tx in range(16):
            c = base[(ty * 16 + tx) % len(base)]
            noise = 0.75 + 0.25 * rng.random((ts, ts, 1), dtype=np.float32)
            img[ty * ts:(ty + 1) * ts, tx * ts:(tx + 1) * ts, :3] = c[None, None, :] * noise
    return img


def minecraft_scene(n=256, copies=1):
    """getMinecraftScene, sceneBuilder.h:220-239, with the voxel-terrain proxy.  The diffuse material is textured:
    set material `scene.textured_material`'s `texture` handle (pt_texture_create / ora_add_texture of voxel_atlas())
    and hasTexture before uploading.  The reference instantiates the model 20 times along x (sceneBuilder.h:228)."""
    s = Scene()
    tex_mat = s.addMaterial(material((1, 1, 1)))
    p = _asset("vokselia_spawn.obj")
    if p:
        m = s.addModel(p, 1.0, (0, 0, 0), (0, 0, 0), tex_mat, True)
    else:
        tri, uv = voxel_terrain_soup(n, 3)
        m = s.addModelSoup(tri, uvs=uv, material=tex_mat)
    for c in range(copies):
        s.addObject(m, position=(float(c * n), 0.0, 0.0))
    s.finalize()
    s.textured_material = tex_mat
    return s


def minecraft_camera(width, height, n=256):
    """Bokeh camera of config C5: aperture 0.05, focalLength = distance to the terrain centre."""
    eye = np.array([0.35 * n, 14.0, -0.35 * n])
    d = np.array([0.0, 0.0, 0.0]) - eye
    dist = float(np.linalg.norm(d))
    return make_camera(tuple(eye), d / dist, 1.5, dist, 0.05, width, height)


def teapot_like_scene(n=40):
    """Small single-instance diffuse scene for quick tests when teapot.obj is not at hand."""
    s = Scene()
    m = s.addMaterial(material((1, 1, 1)))
    s.addObject(s.addModelSoup(statue_soup(4, n), material=m))
    return s.finalize()


def camera_sweep(scene_name, width, height, k=16):
    """K cameras per scene: an ellipse inside the nave for sibenik, an orbit otherwise."""
    cams = []
    for i in range(k):
        a = 2 * math.pi * i / k
        if scene_name == "sibenik":
            eye = (14.0 * math.cos(a), 1.5 + 1.0 * math.sin(2 * a), 4.5 * math.sin(a))
            target = (0.0, 2.0, 0.0)
        elif scene_name == "2mtris":
            eye = (21.0 * math.cos(a), 8.0, 21.0 * math.sin(a))      # the heightfield spans +-20 in x/z after the 0.2 scale
            target = (0.0, -1.0, 0.0)
        else:
            eye = (18.0 * math.cos(a), 6.0, 18.0 * math.sin(a))
            target = (0.0, 0.0, 0.0)
        d = np.array(target) - np.array(eye)
        d = d / np.linalg.norm(d)
        cams.append(make_camera(eye, d, 1.5, 5.0, 0.01, width, height))     # initialCamera() intrinsics, stateLoader.h:30-33
    return cams


SCENES = {"outside": outside_scene, "sibenik": sibenik_scene, "2mtris": two_million_scene, "minecraft": minecraft_scene}
