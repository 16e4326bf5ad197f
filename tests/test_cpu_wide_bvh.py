"""not gpu: the BVH2 -> 8-wide conversion of the host converter (csrc/wide_bvh.cpp through pt_host_wide_bvh), structurally.
The reference has no wide tree (its traversal walks the BVH2 of createBVHBinned), so there is nothing of the reference to compare
the bytes with; what a correct tree must satisfy regardless of how it is cut:
  * every triangle of the model appears in exactly one leaf, leaves hold <= 3 triangles, a node <= 8 children and <= 24 triangles;
  * the decoded (quantised) box of every child contains everything below it -- the traversal never culls a triangle it should test;
  * child / triangle indices form one breadth-first numbering without gaps;
  * the SAH-optimal collapse never costs more than round 1's largest-area rule, and is what pt_upload_scene uses by default.
The device converter is compared byte for byte with this one in tests/test_gpu_bvh_build.py; hit parity in tests/test_gpu_trace.py."""
import numpy as np
import pytest

from cuda_pathtracer_b200 import api, scenes
from cuda_pathtracer_b200.scene import Scene, material


def decode(node):
    """One 80-byte node (20 uint32) -> origin, scale, imask, child_base, tri_base, meta[8], qlo[3][8], qhi[3][8] (wide_bvh.h)."""
    b = node.view(np.uint8)
    p = node[:3].view(np.float32).astype(np.float64)
    e = b[12:15].astype(np.int64) - 127
    return p, np.ldexp(1.0, e), int(b[15]), int(node[4]), int(node[5]), b[24:32], b[32:56].reshape(3, 8).astype(np.float64), b[56:80].reshape(3, 8).astype(np.float64)


def check_tree(nodes, tris, verts, tri_start, tri_count):
    ids = tris[:, 3]
    assert np.array_equal(np.sort(ids), np.arange(tri_start, tri_start + tri_count, dtype=np.uint32)), "every triangle exactly once"
    v = np.stack([verts["v0"], verts["v1"], verts["v2"]], 1).astype(np.float64)          # [n, 3 vertices, 3]
    v0 = tris[:, 0:3].view(np.float32)
    assert np.array_equal(v0, verts["v0"][ids]), "records carry their triangle's first vertex"
    seen_nodes, seen_tris = np.zeros(len(nodes), bool), np.zeros(len(tris), bool)
    seen_nodes[0] = True
    lo_of, hi_of = np.full((len(nodes), 3), np.inf), np.full((len(nodes), 3), -np.inf)      # bounds of everything below a node
    child_box = {}
    order = []
    queue = [0]
    while queue:
        n = queue.pop(0); order.append(n)
        p, scale, imask, child_base, tri_base, meta, qlo, qhi = decode(nodes[n])
        n_internal, n_tri = 0, 0
        for s in range(8):
            m = int(meta[s])
            if m == 0:
                assert not (imask >> s) & 1
                continue
            lo, hi = p + qlo[:, s] * scale, p + qhi[:, s] * scale
            if (imask >> s) & 1:
                assert m == (0x20 | (24 + s))
                c = child_base + n_internal; n_internal += 1
                assert c < len(nodes) and not seen_nodes[c]
                seen_nodes[c] = True; child_box[c] = (lo, hi); queue.append(c)
            else:
                cnt = {1: 1, 3: 2, 7: 3}[m >> 5]
                off = m & 31
                assert off == n_tri, "leaf children take consecutive records"
                r = np.arange(tri_base + off, tri_base + off + cnt)
                assert r[-1] < len(tris) and not seen_tris[r].any()
                seen_tris[r] = True; n_tri += cnt
                tv = v[ids[r]].reshape(-1, 3)
                assert np.all(tv >= lo - 0.0) and np.all(tv <= hi + 0.0), "leaf box contains its triangles"
                lo_of[n] = np.minimum(lo_of[n], tv.min(0)); hi_of[n] = np.maximum(hi_of[n], tv.max(0))
        assert n_internal + bin(imask).count("1") * 0 <= 8 and n_tri <= 24
    assert seen_nodes.all() and seen_tris.all(), "no node or record left out"
    for n in reversed(order):                          # children come after their parent in breadth-first order
        p, scale, imask, child_base, tri_base, meta, qlo, qhi = decode(nodes[n])
        k = 0
        for s in range(8):
            if (imask >> s) & 1:
                c = child_base + k; k += 1
                lo_of[n] = np.minimum(lo_of[n], lo_of[c]); hi_of[n] = np.maximum(hi_of[n], hi_of[c])
        if n in child_box:
            lo, hi = child_box[n]
            assert np.all(lo_of[n] >= lo) and np.all(hi_of[n] <= hi), "an inner child's box contains its whole subtree"
    return len(nodes)


def soups():
    rng = np.random.default_rng(7)
    yield "statue", scenes.statue_soup(6, 40)
    yield "cube", scenes.cube_soup()
    yield "random-soup", (rng.uniform(-5, 5, (3000, 1, 3)) + rng.normal(0, 0.2, (3000, 3, 3))).astype(np.float32).reshape(-1, 9)
    yield "identical-x40", np.tile(np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32), (40, 1))       # a reference leaf of 40 triangles: the cut-leaf path
    yield "tiny-5", scenes.statue_soup(6, 40)[:5]


@pytest.mark.parametrize("name,soup", list(soups()), ids=[n for n, _ in soups()])
def test_wide_bvh_is_a_complete_conservative_tree(name, soup):
    s = Scene()
    s.addMaterial(material())
    s.addModelSoup(np.ascontiguousarray(soup, np.float32).reshape(-1, 9))
    s.addObject(0)
    a = s.finalize().arrays()
    start, count = int(a["models"][0]["triangleStart"]), int(a["models"][0]["nrTriangles"])
    out = {}
    for optimal in (False, True):
        nodes, tris, cost, depth = api.host_wide_bvh(a["bvh"][0], a["vertices"], optimal)
        out[optimal] = (check_tree(nodes, tris, a["vertices"], start, count), cost, depth)
    print(name, "greedy (nodes, cost, depth)", out[False], "optimal", out[True])
    assert out[True][1] <= out[False][1] * (1 + 1e-6), "the optimal cover never costs more"


def test_optimal_collapse_pays_on_the_bench_scene():
    a = scenes.sibenik_scene().arrays()
    m = int(np.argmax(a["models"]["nrTriangles"]))
    g = api.host_wide_bvh(a["bvh"][m], a["vertices"], False)
    o = api.host_wide_bvh(a["bvh"][m], a["vertices"], True)
    print("cathedral shell: greedy %d nodes cost %.2f, optimal %d nodes cost %.2f" % (len(g[0]), g[2], len(o[0]), o[2]))
    assert len(o[0]) < 0.7 * len(g[0]) and o[2] < 0.97 * g[2]
    assert np.array_equal(np.sort(o[1][:, 3]), np.sort(g[1][:, 3]))
