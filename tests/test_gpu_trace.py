"""-m gpu: stage-level parity of the CUDA traversal (pt_debug_trace, through the C ABI) against the CPU
oracle and the committed reference vectors.  Bar: hit ids bit-exact, t within 1e-5 relative (it is in
fact compared bit-for-bit), occlusion booleans bit-exact."""
import numpy as np
import pytest

from cuda_pathtracer_b200 import abi, api
from oracle.oracle_py import OracleLib
from util import classify_hit_mismatches, load_golden_scene

pytestmark = pytest.mark.gpu
T_RTOL = 1e-5


@pytest.fixture(scope="module", params=["teapot", "outside"])
def setup(request):
    a = load_golden_scene(request.param)
    cfg = abi.default_config(160, 120)
    pt = api.Pathtracer(a, cfg).Init()
    ora = OracleLib(cfg)
    desc, keep = abi.make_scene_desc(a)
    ora.upload_scene(desc, keep)
    yield a, pt, ora
    pt.close()
    ora.close()


def _check(h_gpu, h_ref, any_hit):
    n, ties, real = classify_hit_mismatches(h_gpu, h_ref)
    assert real == 0, "%d real hit-id mismatches (%d exact-t ties)" % (real, ties)
    assert n == 0, "%d exact-t ties resolved differently from the reference" % n
    if not any_hit:
        hit = h_ref["primitive_id"] != 0xFFFFFFFF
        assert np.array_equal(h_gpu["t"][hit].view(np.uint32), h_ref["t"][hit].view(np.uint32))
        np.testing.assert_allclose(h_gpu["t"][hit], h_ref["t"][hit], rtol=T_RTOL)


def test_primary_hits_match_reference_vectors(setup):
    a, pt, ora = setup
    _check(pt.debug_trace(a["rays_160x120"], False), a["hits_160x120"], False)


def test_primary_occlusion_matches_reference_vectors(setup):
    a, pt, ora = setup
    _check(pt.debug_trace(a["rays_160x120"], True), a["anyhits_160x120"], True)


def test_primary_rays_match_oracle(setup):
    a, pt, ora = setup
    cam = a["camera_160x120"]
    r_gpu = pt.debug_primary_rays(cam, 3)
    r_ora = ora.primary_rays(cam, 3)
    assert np.array_equal(r_gpu["pixeli"], r_ora["pixeli"])
    # sinf/cosf of the lens sample differ by ulps between glibc and the CUDA math library
    np.testing.assert_allclose(r_gpu["origin"], r_ora["origin"], rtol=0, atol=2e-5)
    np.testing.assert_allclose(r_gpu["direction"], r_ora["direction"], rtol=0, atol=2e-6)


def test_secondary_rays_closest_and_any(setup):
    """Incoherent rays: hemisphere bounces off the primary hit points, plus short shadow-like segments."""
    a, pt, ora = setup
    rays = a["rays_160x120"]
    hits = a["hits_160x120"]
    rng = np.random.default_rng(7)
    sel = np.nonzero(hits["primitive_id"] != 0xFFFFFFFF)[0]
    sec = np.zeros(len(sel) * 4, abi.RAY)
    p = rays["origin"][sel] + rays["direction"][sel] * hits["t"][sel][:, None]
    for k in range(4):
        d = rng.normal(size=(len(sel), 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
        s = sec[k * len(sel):(k + 1) * len(sel)]
        s["origin"] = p + d * np.float32(1e-3)
        s["direction"] = d
        s["length"] = np.float32(9999999) if k < 2 else rng.uniform(0.05, 3.0, len(sel)).astype(np.float32)
        s["pixeli"] = sel
    for any_hit in (False, True):
        _check(pt.debug_trace(sec, any_hit), ora.trace(sec, any_hit), any_hit)
