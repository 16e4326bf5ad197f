"""Shared helpers for the tests: golden scenes, oracle/product construction, comparisons."""
import os

import numpy as np

from cuda_pathtracer_b200 import abi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden_scene(name):
    z = np.load(os.path.join(GOLDEN, name + "_scene.npz"))
    a = {k: z[k] for k in z.files}
    bvh = []
    i = 0
    while "bvh%d" % i in a:
        bvh.append(a.pop("bvh%d" % i))
        i += 1
    a["bvh"] = bvh
    return a


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def classify_hit_mismatches(h_gpu, h_ref):
    """Returns (n_mismatch_ids, n_exact_t_ties, n_real): an id mismatch whose t agrees bit-for-bit is an
    exact-t tie between two triangles (first-found-wins in the reference, lowest-id-wins here)."""
    bad = (h_gpu["primitive_id"] != h_ref["primitive_id"]) | (h_gpu["primitive_type"] != h_ref["primitive_type"]) | \
          (h_gpu["instance_id"] != h_ref["instance_id"])
    idx = np.nonzero(bad)[0]
    ties = int(np.sum(h_gpu["t"][idx].view(np.uint32) == h_ref["t"][idx].view(np.uint32)))
    return len(idx), ties, len(idx) - ties


def read_pfm(path):
    """Colour PFM ("PF", bottom row first) -> float32 (H, W, 3), top row first."""
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(x) for x in f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return np.ascontiguousarray(data[::-1]).astype(np.float32)


def write_obj(path, tri):
    """(n, 9) float32 triangle soup -> a minimal OBJ (v / f lines only), vertices written with repr precision."""
    with open(path, "w") as f:
        for t in tri.reshape(-1, 3):
            f.write("v %s %s %s\n" % (repr(float(t[0])), repr(float(t[1])), repr(float(t[2]))))
        for i in range(len(tri)):
            f.write("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3))


DROPIN = os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "dropin_160x120")
