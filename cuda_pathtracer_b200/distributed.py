"""Multi-GPU plumbing: one process per GPU, image tiles dealt round-robin to ranks (pt_config.rank /
world_size), scene replicated, and ONE exchange per guiding-training frame: the radiance-cache additions
(additionCache || additionCacheCount, 16 floats per triangle) are summed over the ranks between the local
update and the propagate step (reference kernels: src/kernels.h:863-905).  Two transports:
connect_peers() -- the library's own fused reduce + propagate + broadcast kernel over NVLink peer memory (default on
one box); make_allreduce_hook() -- an in-place SUM all-reduce through torch.distributed (NCCL on GPUs, gloo in the
CPU tests) followed by the local propagate kernel."""
import ctypes as C
import os

import numpy as np

from . import abi


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


class HostRing:
    """The process-shared host framebuffer ring of pt_host_ring_* for one job: rank 0 creates a file in /dev/shm (zero-filled:
    header + `slots` images of H*W float4), every rank maps it and attaches it to its Pathtracer.  After that each rank's
    deliver_image_async() writes its own tiles through its own PCIe link and the consumer (rank 0) reads whole frames in place.
    Works without torch.distributed for a single rank."""

    def __init__(self, pt, slots=4, group=None, name=None):
        import mmap
        self.pt, self.slots = pt, int(slots)
        nbytes = abi.HOST_RING_HEADER_BYTES + self.slots * pt.cfg.height * pt.cfg.width * 16
        world = pt.cfg.world_size
        dist = None
        if world > 1:
            import torch.distributed as dist
        self.owner = pt.cfg.rank == 0
        names = [name or ("/dev/shm/ptb200_ring_%d_%d" % (os.getpid(), id(self) & 0xffff)), True]
        if self.owner:                                 # a container may cap /dev/shm (64 MB by default): refuse on every rank rather than fault on a page
            try:
                st = os.statvfs(os.path.dirname(names[0]))
                names[1] = st.f_bavail * st.f_frsize >= nbytes + (8 << 20)
            except OSError:
                names[1] = False
        if world > 1:
            dist.broadcast_object_list(names, src=0, group=group)
        self.path = names[0]
        if not names[1]:
            self.pt = None
            raise RuntimeError("HostRing: not enough room in %s for %d bytes of shared framebuffers" % (os.path.dirname(self.path), nbytes))
        if self.owner:
            with open(self.path, "wb") as f:
                f.truncate(nbytes)                     # sparse file of zeros: header and images start at 0
        if world > 1:
            dist.barrier(group=group)
        self._f = open(self.path, "r+b")
        self._map = mmap.mmap(self._f.fileno(), nbytes)
        self.block = np.frombuffer(self._map, dtype=np.uint8)
        self._group, self._dist = group, dist
        ok, why = True, ""
        try:
            pt.host_ring_attach(self.block, self.slots)
        except Exception as ex:                        # noqa: BLE001 -- agreed on by all ranks below
            ok, why = False, str(ex)
        oks = [(ok, why)]
        if world > 1:
            oks = [None] * world
            dist.all_gather_object(oks, (ok, why), group=group)
        if not all(o for o, _ in oks):                 # one rank could not map the block: nobody uses it
            if ok:
                pt.host_ring_detach()
            self.block = None
            self._map.close(); self._f.close()
            if world > 1:
                dist.barrier(group=group)
            if self.owner:
                os.unlink(self.path)
            self.pt = None
            raise RuntimeError("HostRing: " + "; ".join(w for o, w in oks if not o))

    def close(self):
        if self.pt is None:
            return
        if self._dist is not None:
            self._dist.barrier(group=self._group)
        self.pt.host_ring_detach()
        self.block = None
        try:
            self._map.close()
        except BufferError:                           # a consumer still holds a view: the mapping goes with the process
            pass
        self._f.close()
        if self._dist is not None:
            self._dist.barrier(group=self._group)
        if self.owner:
            try:
                os.unlink(self.path)
            except OSError:
                pass
        self.pt = None
