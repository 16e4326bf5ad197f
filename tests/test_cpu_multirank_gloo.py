"""world_size-2 gloo test of the multi-rank host logic (no GPU): the tile partition deals every pixel to
exactly one rank, and the all-reduce hook sums the per-rank radiance-cache additions in place so that the
propagate step sees what a single rank would have accumulated."""
import ctypes as C
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cuda_pathtracer_b200 import abi, api
from cuda_pathtracer_b200.distributed import gather_image, make_allreduce_hook

W, H, T = 200, 120, 257


def _propagate(rc, add):          # kernels.h:884-905 in numpy, float32
    rc = rc.copy()
    for t in range(8):
        cnt = add[8 + t]
        m = cnt >= np.float32(0.001)
        old = rc[:, t]
        inc = np.divide(add[t], cnt, out=np.zeros_like(cnt), where=m)
        new = np.maximum(np.float32(0.1), np.minimum(np.float32(0.95) * old + np.float32(1 - np.float32(0.95)) * inc, np.float32(2.0)))
        rc[:, t] = np.where(m, new, old)
    return rc


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    L = api.load_library()
    L.pt_tile_num_slots.restype = C.c_uint32
    cfg = abi.default_config(W, H, rank=rank, world_size=world)
    own = np.zeros((H, W, 4), np.float32)
    x, y = C.c_uint32(), C.c_uint32()
    for slot in range(L.pt_tile_num_slots(C.byref(cfg))):
        if L.pt_tile_slot_to_pixel(C.byref(cfg), slot, C.byref(x), C.byref(y)) == 1:
            own[y.value, x.value] += (1.0, rank, 0.0, 1.0)
    full = gather_image(own).numpy()
    ok_cover = bool(np.all(full[..., 0] == 1.0) and np.all(full[..., 3] == 1.0))
    # additions: every rank contributes the samples of ITS pixels; the hook must leave the global sum everywhere
    rng = np.random.default_rng(100 + rank)
    add = (rng.random((16, T)) * (rng.random((16, T)) < 0.3)).astype(np.float32)
    add[8:] = np.round(add[8:] * 4)
    mine = add.copy()
    hook = make_allreduce_hook("cpu")
    assert hook(add.ctypes.data, add.size, 0) == 0
    q.put((rank, ok_cover, mine, add))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_partition_and_allreduce(built):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=180) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert all(r[1] for r in res), "tile partition does not cover every pixel exactly once"
    total = res[0][2] + res[1][2]
    for r in res:
        assert np.array_equal(r[3], total), "all-reduce hook did not leave the global sum on rank %d" % r[0]
    rc = np.full((T, 8), 0.1, np.float32)
    assert np.array_equal(_propagate(rc, res[0][3]), _propagate(rc, res[1][3]))      # replicas stay identical


# ---- radiance-cache exchange over peer memory: host protocol (distributed.connect_peers) and the owner-range algorithm ----
class _FakePt:
    """Stands in for api.Pathtracer on a machine without a GPU: records what connect_peers does with the handles."""

    def __init__(self, rank, fail):
        self.rank, self.fail, self.connected, self.disconnected = rank, fail, None, False

    def peer_export(self):
        return bytes([self.rank]) * 64

    def peer_connect(self, handles):
        if self.fail:
            raise RuntimeError("cannot map the block of rank 0")
        self.connected = list(handles)

    def peer_disconnect(self):
        self.disconnected = True


def _owner_range_exchange(rc, adds, world):
    """What k_peer_reduce_propagate does (pt_kernels.cuh), in numpy: the owner of a contiguous triangle range sums every
    rank's additions for it in rank order, propagates once (kernels.h:884-905) and hands the rows to every rank."""
    n = rc.shape[0]
    out = rc.copy()
    for owner in range(world):
        lo, hi = n * owner // world, n * (owner + 1) // world          # pt_api.cu: T * rank / world
        total = np.zeros((16, hi - lo), np.float32)
        for r in range(world):
            total += adds[r][:, lo:hi]
        out[lo:hi] = _propagate(rc[lo:hi], total)
    return out


def _peer_worker(rank, world, port, q):
    from cuda_pathtracer_b200.distributed import connect_peers
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    good = _FakePt(rank, False)
    ok1 = connect_peers(good)
    bad = _FakePt(rank, rank == 1)                   # one rank cannot map its peers: every rank must fall back
    ok2 = connect_peers(bad)
    q.put((rank, ok1, good.connected, good.disconnected, ok2, bad.disconnected))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_peer_handle_exchange_and_fallback(built):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_peer_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=180) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, ok1, handles, disc1, ok2, disc2 in res:
        assert ok1 is True and handles == [bytes([0]) * 64, bytes([1]) * 64] and not disc1      # every rank got all handles, rank order
        assert ok2 is False and disc2                                                           # one failure => all ranks disconnect


@pytest.mark.parametrize("world", [2, 3, 8])
def test_owner_range_exchange_equals_allreduce_then_propagate(world):
    rng = np.random.default_rng(world)
    rc = (0.1 + 1.9 * rng.random((T, 8))).astype(np.float32)
    adds = []
    for r in range(world):
        a = (rng.random((16, T)) * (rng.random((16, T)) < 0.3)).astype(np.float32)
        a[8:] = np.round(a[8:] * 4)
        a[:8] *= a[8:] > 0                           # energy only where something was counted
        adds.append(a)
    total = np.zeros((16, T), np.float32)
    for a in adds:                                   # the same rank order as the owner kernels use
        total += a
    assert np.array_equal(_owner_range_exchange(rc, adds, world), _propagate(rc, total))
