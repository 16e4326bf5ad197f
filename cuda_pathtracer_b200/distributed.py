"""Multi-GPU plumbing: one process per GPU, image tiles dealt round-robin to ranks (pt_config.rank /
world_size), scene replicated, and ONE exchange per guiding-training frame: the radiance-cache additions
(additionCache || additionCacheCount, 16 floats per triangle) are summed over the ranks between the local
update and the propagate step (reference kernels: src/kernels.h:863-905).  Two transports:
connect_peers() -- the library's own fused reduce + propagate + broadcast kernel over NVLink peer memory (default on
one box); make_allreduce_hook() -- an in-place SUM all-reduce through torch.distributed (NCCL on GPUs, gloo in the
CPU tests) followed by the local propagate kernel."""
import ctypes as C

import numpy as np


class _DevBuf:
    """__cuda_array_interface__ view of a raw device pointer handed out by the C ABI."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def wrap_float_buffer(ptr, n, device):
    """torch view (no copy) of n floats at `ptr`: device memory for cuda devices, host memory for 'cpu'."""
    import torch
    if str(device).startswith("cuda"):
        return torch.as_tensor(_DevBuf(ptr, n), device=device)
    buf = (C.c_float * n).from_address(int(ptr))
    return torch.from_numpy(np.frombuffer(buf, dtype=np.float32, count=n))


def make_allreduce_hook(device, stream=None, group=None):
    """Returns fn(ptr, count, stream_handle) -> 0 for Pathtracer.set_allreduce_hook: sums the additions buffer
    over all ranks in place, ordered on the path tracer's stream."""
    import torch
    import torch.distributed as dist

    streams = {}

    def hook(ptr, count, stream_handle):
        t = wrap_float_buffer(ptr, count, device)
        st = stream
        if st is None and stream_handle and str(device).startswith("cuda"):
            # the library passes the stream the sample runs on (two samples alternate between two streams)
            st = streams.get(stream_handle)
            if st is None:
                st = streams[stream_handle] = torch.cuda.ExternalStream(stream_handle, device=device)
        if st is not None:
            with torch.cuda.stream(st):
                dist.all_reduce(t, group=group)
        else:
            dist.all_reduce(t, group=group)
        return 0
    return hook


def connect_peers(pt, device=None, group=None):
    """Switches the radiance-cache exchange of `pt` to the fused reduce + propagate kernel over NVLink peer memory
    (pt_peer_export / pt_peer_connect): the cudaIpc handles travel through one all_gather, after that the data path
    never touches NCCL.  Every rank of the group must call it after Init().  Returns True when every rank mapped every
    peer; otherwise every rank stays on (or falls back to) the all-reduce hook and False is returned."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    handles = [None] * world
    dist.all_gather_object(handles, pt.peer_export(), group=group)
    ok = True
    try:
        pt.peer_connect(handles)
    except Exception:                                    # noqa: BLE001 -- reported through the return value on every rank
        ok = False
    oks = [None] * world
    dist.all_gather_object(oks, ok, group=group)
    if not all(oks):
        pt.peer_disconnect()
        return False
    return True


def disconnect_peers(pt, group=None):
    """Unmaps the peers' blocks; collective (a barrier on both sides keeps a rank from freeing a block in use)."""
    import torch.distributed as dist
    dist.barrier(group=group)
    pt.peer_disconnect()
    dist.barrier(group=group)


def gather_image(local_image, group=None):
    """Every rank holds a full-size float4 image with zeros outside its tiles: the sum is the framebuffer."""
    import torch
    import torch.distributed as dist
    t = local_image if isinstance(local_image, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(local_image))
    dist.all_reduce(t, group=group)
    return t
