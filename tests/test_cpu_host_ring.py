"""not gpu: the host side of distributed.HostRing (the process-shared framebuffer block of pt_host_ring_*) with a stand-in for the
Pathtracer: file layout, zero fill, clean-up, and the refusal paths (a rank that cannot map the block, a /dev/shm that is too small).
The device side runs in tests/test_gpu_peer_exchange.py (one rank) and tools/multi_gpu_check.py (2-8 ranks)."""
import os
import types

import numpy as np
import pytest

from cuda_pathtracer_b200 import abi
from cuda_pathtracer_b200.distributed import HostRing


class FakePt:
    def __init__(self, w=40, h=24, fail=False):
        self.cfg = types.SimpleNamespace(width=w, height=h, rank=0, world_size=1)
        self.fail, self.attached, self.detached = fail, None, False

    def host_ring_attach(self, block, slots):
        if self.fail:
            raise RuntimeError("cannot map")
        self.attached = (block, slots)

    def host_ring_detach(self):
        self.detached = True
        self.attached = None


def test_block_layout_and_cleanup():
    pt = FakePt()
    ring = HostRing(pt, slots=3)
    block, slots = pt.attached
    assert slots == 3 and block.dtype == np.uint8 and block.nbytes == abi.HOST_RING_HEADER_BYTES + 3 * 24 * 40 * 16
    assert os.path.getsize(ring.path) == block.nbytes and not block.any()          # header and images start at zero
    block[5] = 7                                                                     # writable, shared mapping
    with open(ring.path, "rb") as f:
        assert f.read(6)[5] == 7
    path = ring.path
    del block
    ring.close()
    assert pt.detached and not os.path.exists(path)
    ring.close()                                                                     # idempotent


def test_a_rank_that_cannot_map_refuses_for_everyone():
    pt = FakePt(fail=True)
    before = set(os.listdir("/dev/shm"))
    with pytest.raises(RuntimeError, match="cannot map"):
        HostRing(pt, slots=2)
    assert set(os.listdir("/dev/shm")) == before                                     # nothing left behind


def test_too_small_shared_memory_is_refused_before_anything_is_created(monkeypatch):
    real = os.statvfs

    def tiny(path):
        st = real(path)
        return types.SimpleNamespace(f_bavail=1, f_frsize=st.f_frsize)
    monkeypatch.setattr(os, "statvfs", tiny)
    pt = FakePt()
    before = set(os.listdir("/dev/shm"))
    with pytest.raises(RuntimeError, match="not enough room"):
        HostRing(pt, slots=4)
    assert pt.attached is None and set(os.listdir("/dev/shm")) == before
