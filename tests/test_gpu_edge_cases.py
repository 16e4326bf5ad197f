"""-m gpu: edge cases of the path against the CPU oracle -- ragged image sizes (tiles cut by the border), primitive-only
scenes (no triangles, no instances), rays that all miss (empty extension / shadow queues after the first bounce),
the maximum depth (64 bounces: 61 of them in k_tail), 3-rank tile partition of a ragged image, a 1x1 image."""
import numpy as np
import pytest

from cuda_pathtracer_b200 import abi, api, scenes
from cuda_pathtracer_b200.scene import Scene, make_camera, material
from oracle.oracle_py import OracleLib

pytestmark = pytest.mark.gpu


def rel_rmse(a, b):
    return float(np.sqrt(np.mean((a.astype(np.float64) - b) ** 2)) / max(1e-12, np.sqrt(np.mean(b.astype(np.float64) ** 2))))


def both(arrays, cfg, cam, frames, nee=True, guiding=False):
    pt = api.Pathtracer(arrays, cfg).Init()
    ora = OracleLib(cfg)
    desc, keep = abi.make_scene_desc(arrays)
    ora.upload_scene(desc, keep)
    pt.HNEE, pt.HCACHE = nee, guiding
    ora.set_toggles(nee, guiding)
    pt.Render(cam, shouldClear=True); ora.render(cam, True)
    for _ in range(frames):
        pt.Render(cam); ora.render(cam, False)
    pt.Finish()
    out = pt.read_image().copy(), ora.read_image().copy(), pt.stats(), ora.stats()[0]
    pt.close(); ora.close()
    return out


def primitive_scene():
    s = Scene()
    w = s.addMaterial(material((0.6, 0.6, 0.6)))
    g = s.addMaterial(material((1, 1, 1), transmit=1.0, refractive_index=1.5))
    m = s.addMaterial(material((1, 1, 1), reflect=1.0))
    s.addSphere((0, 0, 0), 1.0, g); s.addSphere((2.5, 0, 0), 1.0, w); s.addSphere((-2.5, 0, 0), 1.0, m)
    s.addPlane((0, -1, 0), -1.0, w)
    return s.finalize().arrays()


@pytest.mark.parametrize("w,h", [(37, 23), (1, 1), (33, 65)])
def test_primitive_only_scene_ragged_sizes_max_depth(w, h):
    """No triangles, no instances, no TLAS; image sizes that are not multiples of the 32x32 tile; MAX_RAY_DEPTH 64."""
    cfg = abi.default_config(w, h, max_ray_depth=64, interactive_depth=1)
    cam = make_camera((0, 1, -6), (0, -0.1, 1), 1.5, 5.0, 0.01, w, h)
    g, r, st, ost = both(primitive_scene(), cfg, cam, 3)
    assert g.shape == (h, w, 4) and np.all(g[..., 3] == 4.0) and np.isfinite(g).all()
    assert rel_rmse(g[..., :3], r[..., :3]) <= 1e-3
    assert st.primary_rays == ost.primary_rays == 4 * w * h and st.shadow_rays == 0


def test_all_rays_miss_empty_queues():
    """The camera looks away from the scene: every primary ray ends on the sky, all later queues are empty."""
    a = scenes.teapot_like_scene(20).arrays()
    cfg = abi.default_config(64, 48, max_ray_depth=8, interactive_depth=1)
    cam = make_camera((0, 50, 0), (0, 1, 0), 1.5, 5.0, 0.01, 64, 48)
    g, r, st, ost = both(a, cfg, cam, 2)
    assert st.extend_rays == st.primary_rays == 3 * 64 * 48 and st.shadow_rays == 0
    assert np.array_equal(g, r)                      # constant sky: exact


def test_ragged_image_three_ranks_sum_to_the_single_rank_image():
    a = scenes.sibenik_scene().arrays()
    cam = scenes.camera_sweep("sibenik", 70, 45, 16)[6]
    imgs = []
    for rank, world in [(0, 1), (0, 3), (1, 3), (2, 3)]:
        cfg = abi.default_config(70, 45, max_ray_depth=8, interactive_depth=1, rank=rank, world_size=world)
        pt = api.Pathtracer(a, cfg).Init()
        pt.HNEE, pt.HCACHE = True, False
        pt.Render(cam, shouldClear=True)
        for _ in range(3):
            pt.Render(cam)
        pt.Finish()
        imgs.append(pt.read_image().copy())
        pt.close()
    assert np.array_equal(imgs[0], imgs[1] + imgs[2] + imgs[3])
    assert np.all(imgs[0][..., 3] == 4.0)


def test_deep_paths_through_glass_in_the_tail_kernel():
    """outside scene (glass cubes, checker plane): many paths survive past bounce 3, so k_tail carries real load;
    depth 24, NEE on (no emitters => device NEE off), guiding off."""
    from util import load_golden_scene
    a = load_golden_scene("outside")
    cfg = abi.default_config(160, 120, max_ray_depth=24, interactive_depth=1)
    g, r, st, ost = both(a, cfg, a["camera_160x120"], 4)
    # libm ulps (expf in Beer absorption, sinf/cosf in the glossy lobe) compound over ~10 refraction events per path
    assert rel_rmse(g[..., :3], r[..., :3]) <= 5e-3
    assert st.extend_rays > 1.5 * st.primary_rays and abs(int(st.extend_rays) - int(ost.extend_rays)) <= 1e-3 * ost.extend_rays
    # ... and the scheduling is not the source: wavefront-only, one sample in flight gives the same bits
    cfg2 = abi.default_config(160, 120, max_ray_depth=24, interactive_depth=1, flags=1 | 2)
    pt = api.Pathtracer(a, cfg2).Init()
    pt.HNEE, pt.HCACHE = True, False
    pt.Render(a["camera_160x120"], shouldClear=True)
    for _ in range(4):
        pt.Render(a["camera_160x120"])
    pt.Finish()
    assert np.array_equal(pt.read_image(), g) and pt.stats().extend_rays == st.extend_rays
    pt.close()
