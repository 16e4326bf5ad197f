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
