"""Host-side mirror of the reference's operator surface over the C ABI (include/ptb200.h).

``Pathtracer`` has the shape of the reference's ``class Pathtracer : public Application``
(reference src/application.h:7-20, src/pathtracer.h:46-310): ``Init()``, ``Render(camera, currentTime,
frameTime, shouldClear)``, ``Finish()``, plus the two host toggles HNEE / HCACHE (src/globals.h:19-20).
There is no CPU fallback: loading fails loudly when libptb200.so is missing, and pt_create fails
loudly without a CUDA device.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class PtError(RuntimeError):
    pass


def lib_path():
    # PTB200_LIB: developer knob for kernel tuning (tools/tune.sh builds variants of the same library)
    return os.environ.get("PTB200_LIB") or os.path.join(_HERE, "libptb200.so")


def load_library():
    """Load libptb200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing."""
    global _LIB
    if _LIB is not None:
        return _LIB
    p = lib_path()
    if not os.path.exists(p):
        raise PtError("libptb200.so not found at %s -- run __graft_entry__.build(); there is no fallback path" % p)
    L = C.CDLL(p)
    L.pt_version.restype = C.c_char_p
    L.pt_last_error.restype = C.c_char_p
    L.pt_last_error.argtypes = [C.c_void_p]
    L.pt_cuda_stream.restype = C.c_void_p
    L.pt_cuda_stream.argtypes = [C.c_void_p]
    L.pt_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(abi.Config)]
    L.pt_destroy.argtypes = [C.c_void_p]
    L.pt_destroy.restype = None
    L.pt_upload_scene.argtypes = [C.c_void_p, C.POINTER(abi.SceneDesc)]
    L.pt_build_bvh.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32)]
    L.pt_build_bvh_last_error.restype = C.c_char_p
    L.pt_read_image_async.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
    L.pt_read_tiles_async.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
    L.pt_wait_read.argtypes = [C.c_void_p, C.c_uint64]
    L.pt_debug_read_wide_bvh.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p, C.POINTER(C.c_uint32)]
    L.pt_update_instances.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
    L.pt_update_instances_device_tlas.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    L.pt_debug_read_tlas.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    L.pt_set_toggles.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.pt_set_skydome.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    L.pt_set_sky_constant.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float]
    L.pt_set_blue_noise.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    L.pt_texture_create.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]
    L.pt_set_allreduce_hook.argtypes = [C.c_void_p, abi.ALLREDUCE_HOOK, C.c_void_p]
    L.pt_peer_export.argtypes = [C.c_void_p, C.c_void_p]
    L.pt_peer_connect.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
    L.pt_peer_disconnect.argtypes = [C.c_void_p]
    L.pt_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.pt_finish.argtypes = [C.c_void_p]
    L.pt_join.argtypes = [C.c_void_p]
    L.pt_read_tiles.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.pt_tile_num_slots.argtypes = [C.POINTER(abi.Config)]
    L.pt_tile_num_slots.restype = C.c_uint32
    L.pt_measure_gather_bandwidth.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.POINTER(C.c_double)]
    L.pt_postprocess.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p]
    L.pt_write_pfm.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    L.pt_set_guiding_frames.argtypes = [C.c_void_p, C.c_uint32]
    L.pt_read_image.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.pt_read_radiance_cache.argtypes = [C.c_void_p, C.c_void_p]
    L.pt_get_stats.argtypes = [C.c_void_p, C.POINTER(abi.Stats)]
    L.pt_reset_stats.argtypes = [C.c_void_p]
    L.pt_set_profiling.argtypes = [C.c_void_p, C.c_int]
    L.pt_get_profile.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    L.pt_get_rand_state.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.pt_debug_primary_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    L.pt_debug_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
    L.pt_set_sky_sampling.argtypes = [C.c_void_p, C.c_int]
    L.pt_set_ray_sorting.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32]
    L.pt_set_counting.argtypes = [C.c_void_p, C.c_int]
    L.pt_get_tally.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    L.pt_debug_read_counters.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]
    L.pt_gather_image_async.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
    L.pt_trace_static.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
    L.pt_host_ring_attach.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]
    L.pt_host_ring_detach.argtypes = [C.c_void_p]
    L.pt_deliver_image_async.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
    L.pt_host_ring_wait.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.POINTER(C.c_float))]
    L.pt_host_ring_release.argtypes = [C.c_void_p, C.c_uint64]
    L.pt_host_wide_bvh.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_double), C.POINTER(C.c_uint32)]
    L.pt_debug_check_report.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
    L.pt_debug_trace_resident.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.POINTER(C.c_float), C.c_void_p]
    _LIB = L
    return L


class Pathtracer:
    """Drop-in for the reference's Pathtracer application object.

    scene: dict of numpy arrays (abi.make_scene_desc keys) -- the host vectors of the reference's
    ``Scene`` (src/scene.h:124-133) -- or a ``cuda_pathtracer_b200.scene.Scene``.
    """

    def __init__(self, scene, config=None):
        self.L = load_library()
        self.cfg = config if config is not None else abi.default_config()
        self.scene = scene
        self.h = None
        self.HNEE = True      # globals.h:19, set in Init (pathtracer.h:213)
        self.HCACHE = True    # globals.h:20, set in Init (pathtracer.h:214)
        self._keep = []
        self._hook = None

    # -- error plumbing --
    def _ck(self, r, what):
        if r < 0:
            msg = self.L.pt_last_error(self.h)
            raise PtError("%s failed (%d): %s" % (what, r, (msg or b"").decode()))
        return r

    def _arrays(self):
        return self.scene.arrays() if hasattr(self.scene, "arrays") else self.scene

    # -- Application interface --
    def Init(self, textures=None, normal_maps=None):
        """textures / normal_maps: {material id: float4 RGBA image (h, w, 4)} -- what Scene::addModel does with the
        MTL's map_Kd / norm entries (scene.h:218-243): the image becomes a texture object whose handle is stored in
        the material before the scene is uploaded."""
        h = C.c_void_p()
        r = self.L.pt_create(C.byref(h), C.byref(self.cfg))
        if r < 0:
            raise PtError("pt_create failed (%d): %s" % (r, (self.L.pt_last_error(None) or b"").decode()))
        self.h = h
        arrays = self._arrays()
        if textures or normal_maps:
            arrays = dict(arrays)
            mats = arrays["materials"].copy()
            for mid, img in (textures or {}).items():
                mats[mid]["texture"] = self.create_texture(img); mats[mid]["hasTexture"] = 1
            for mid, img in (normal_maps or {}).items():
                mats[mid]["normal_texture"] = self.create_texture(img); mats[mid]["hasNormalMap"] = 1
            arrays["materials"] = mats
        desc, keep = abi.make_scene_desc(arrays)
        self._keep.append(keep)
        self.num_lights = self._ck(self.L.pt_upload_scene(self.h, C.byref(desc)), "pt_upload_scene")
        self.num_triangles = desc.num_triangles
        self.HNEE = True
        self.HCACHE = True
        return self

    def Render(self, camera, currentTime=0.0, frameTime=0.0, shouldClear=False):
        """camera: numpy record of abi.CAMERA (lt/u/v prepared like Camera::recalculate, types.h:590-600)."""
        if shouldClear:
            # pathtracer.h:242-246: toggles and instances/TLAS are re-synced only on a clearing frame
            self._ck(self.L.pt_set_toggles(self.h, int(self.HNEE), int(self.HCACHE)), "pt_set_toggles")
            a = self._arrays()
            if len(a["instances"]):
                inst = np.ascontiguousarray(a["instances"])
                tlas = np.ascontiguousarray(a["topBvh"])
                if getattr(self, "DEVICE_TLAS", False):      # top-level BVH rebuilt on the device from the instance matrices
                    self._ck(self.L.pt_update_instances_device_tlas(self.h, abi.ptr(inst), len(inst)), "pt_update_instances_device_tlas")
                else:
                    self._ck(self.L.pt_update_instances(self.h, abi.ptr(inst), len(inst), abi.ptr(tlas), len(tlas)), "pt_update_instances")
        cam = np.ascontiguousarray(camera)
        self._ck(self.L.pt_render(self.h, abi.ptr(cam), int(bool(shouldClear))), "pt_render")

    def Finish(self):
        self._ck(self.L.pt_finish(self.h), "pt_finish")

    # -- headless replacements for the GL textures --
    def read_image(self, which=0, out=None):
        if out is None:
            out = np.empty((self.cfg.height, self.cfg.width, 4), np.float32)
        self._ck(self.L.pt_read_image(self.h, abi.ptr(out), which), "pt_read_image")
        return out

    def read_radiance_cache(self):
        out = np.zeros(self.num_triangles, abi.RADIANCE_CACHE)
        self._ck(self.L.pt_read_radiance_cache(self.h, abi.ptr(out)), "pt_read_radiance_cache")
        return out

    # -- environment --
    def set_skydome(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.float32)
        self._ck(self.L.pt_set_skydome(self.h, abi.ptr(rgba), rgba.shape[1], rgba.shape[0]), "pt_set_skydome")

    def set_sky_constant(self, r, g, b):
        self._ck(self.L.pt_set_sky_constant(self.h, r, g, b), "pt_set_sky_constant")

    def set_blue_noise(self, lum):
        lum = np.ascontiguousarray(lum, np.float32)
        self._ck(self.L.pt_set_blue_noise(self.h, abi.ptr(lum), lum.shape[1], lum.shape[0]), "pt_set_blue_noise")

    def create_texture(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.float32)
        out = C.c_uint64()
        self._ck(self.L.pt_texture_create(self.h, abi.ptr(rgba), rgba.shape[1], rgba.shape[0], C.byref(out)), "pt_texture_create")
        return out.value

    def set_sky_sampling(self, on):
        """Opt-in skydome importance sampling through the CDF of pathtracer.h:85-122 (the reference builds it, no kernel samples it)."""
        self._ck(self.L.pt_set_sky_sampling(self.h, int(bool(on))), "pt_set_sky_sampling")

    def set_ray_sorting(self, on, bits=0, bounces=0):
        """Opt-in counting sort of the extension / shadow queues by (origin cell, direction octant) between shade and trace
        (the reference's disabled experiment, src/raySorting.h); images are bit-identical."""
        self._ck(self.L.pt_set_ray_sorting(self.h, int(bool(on)), int(bits), int(bounces)), "pt_set_ray_sorting")

    def set_counting(self, on):
        self._ck(self.L.pt_set_counting(self.h, int(bool(on))), "pt_set_counting")

    def tally(self):
        """{'extend': (node steps, triangle tests, top-level steps), 'connect': (...)} since set_counting(True)."""
        out = (C.c_uint64 * 8)()
        self._ck(self.L.pt_get_tally(self.h, out), "pt_get_tally")
        return {"extend": tuple(int(x) for x in out[0:3]), "connect": tuple(int(x) for x in out[3:6]), "longest_ray_steps": (int(out[6]), int(out[7]))}

    def debug_read_counters(self, n):
        e, s = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        self._ck(self.L.pt_debug_read_counters(self.h, abi.ptr(e), abi.ptr(s), n), "pt_debug_read_counters")
        return e, s

    def gather_image_async(self, out, which=0):
        """Collective over the connected peers: the N-rank framebuffer assembled over NVLink in rank 0's HBM and shipped to
        `out` ((H, W, 4) pinned float32 on rank 0, None elsewhere) with one copy.  Returns a ticket for wait_read()."""
        t = C.c_uint64()
        self._ck(self.L.pt_gather_image_async(self.h, abi.ptr(out) if out is not None else None, int(which), C.byref(t)), "pt_gather_image_async")
        return t.value

    # -- one framebuffer in shared host memory, every rank through its own PCIe link (pt_host_ring_*) --
    def host_ring_attach(self, block, slots):
        """block: writable uint8 buffer over process-shared host memory (distributed.HostRing creates one per job),
        abi.HOST_RING_HEADER_BYTES + slots * H * W * 16 bytes, zero-filled by its creator."""
        self._ck(self.L.pt_host_ring_attach(self.h, abi.ptr(block), block.nbytes, int(slots)), "pt_host_ring_attach")
        self._ring_block = block                       # keep the mapping alive while the GPU writes into it

    def host_ring_detach(self):
        self._ck(self.L.pt_host_ring_detach(self.h), "pt_host_ring_detach")
        self._ring_block = None

    def deliver_image_async(self, which=0):
        """Every rank: this rank's tiles of the accumulated image into the ring slot of the next epoch; returns the epoch."""
        e = C.c_uint64()
        self._ck(self.L.pt_deliver_image_async(self.h, int(which), C.byref(e)), "pt_deliver_image_async")
        return e.value

    def host_ring_wait(self, epoch):
        """Consumer: the (H, W, 4) framebuffer of `epoch` as a view into the shared block, once all ranks have delivered."""
        p = C.POINTER(C.c_float)()
        self._ck(self.L.pt_host_ring_wait(self.h, int(epoch), C.byref(p)), "pt_host_ring_wait")
        return np.ctypeslib.as_array(p, shape=(self.cfg.height, self.cfg.width, 4))

    def host_ring_release(self, epoch):
        self._ck(self.L.pt_host_ring_release(self.h, int(epoch)), "pt_host_ring_release")

    def set_guiding_frames(self, frames):
        """Frames during which the radiance cache trains (reference pathtracer.h:292 hard-codes 100)."""
        self._ck(self.L.pt_set_guiding_frames(self.h, int(frames)), "pt_set_guiding_frames")

    def write_pfm(self, path, which=0):
        """Headless render-to-PFM: the display value rgb / sample count (reference main.cpp:58-62)."""
        self._ck(self.L.pt_write_pfm(self.h, os.fsencode(path), int(which)), "pt_write_pfm")

    def postprocess(self, blur=True, lum=None, albedo=None, nr_samples=None):
        """The reference's display pass (main.cpp:316-376) headless: demodulated separable Gaussian, remodulate,
        gamma 2, vignette.  lum/albedo: float4 images (e.g. a gathered multi-GPU framebuffer) or None for the
        context's own accumulators.  Returns the RGBA float image the fragment shader writes."""
        w, h = self.cfg.width, self.cfg.height
        out = np.empty((h, w, 4), np.float32)
        if lum is not None:
            lum = np.ascontiguousarray(lum, np.float32); albedo = np.ascontiguousarray(albedo, np.float32)
        if nr_samples is None:
            nr_samples = float(self.rand_state()[0])
        self._ck(self.L.pt_postprocess(self.h, abi.ptr(lum), abi.ptr(albedo), float(nr_samples), int(blur), abi.ptr(out)), "pt_postprocess")
        return out

    def measure_gather_bandwidth(self, array_bytes=32 << 20, access_bytes=32, repeats=5):
        """GB/s of random 16/32-byte reads over an array (default 32 MB: L2-resident) -- the ceiling for BVH gathers."""
        out = C.c_double()
        self._ck(self.L.pt_measure_gather_bandwidth(self.h, int(array_bytes), int(access_bytes), int(repeats), C.byref(out)), "pt_measure_gather_bandwidth")
        return out.value

    def num_slots(self):
        return int(self.L.pt_tile_num_slots(C.byref(self.cfg)))

    def update_instances_device_tlas(self, instances):
        """Scene::update's instance upload with BuildTopLevelBVH (scene.h:81-118) run on the device."""
        inst = np.ascontiguousarray(instances)
        self._ck(self.L.pt_update_instances_device_tlas(self.h, abi.ptr(inst), len(inst)), "pt_update_instances_device_tlas")

    def debug_read_tlas(self):
        out = np.zeros(len(self._arrays()["topBvh"]), abi.TLAS_NODE)
        self._ck(self.L.pt_debug_read_tlas(self.h, abi.ptr(out), len(out)), "pt_debug_read_tlas")
        return out

    def read_image_async(self, out, which=0, tiles=False):
        """Streaming read-back into `out` (pinned host memory: (H, W, 4) float32, or (num_slots, 4) with tiles=True): the
        image as it is after every Render call made so far, copied while later samples render.  Returns a ticket for
        wait_read(); Finish() / join() cover all outstanding copies."""
        t = C.c_uint64()
        fn = self.L.pt_read_tiles_async if tiles else self.L.pt_read_image_async
        self._ck(fn(self.h, abi.ptr(out), int(which), C.byref(t)), "pt_read_image_async")
        return t.value

    def wait_read(self, ticket):
        self._ck(self.L.pt_wait_read(self.h, C.c_uint64(ticket)), "pt_wait_read")

    def debug_read_wide_bvh(self):
        """(nodes [n, 20] uint32, triangle records [m, 12] uint32) of the 8-wide BVH as it sits on the device."""
        nn, nt = C.c_uint32(), C.c_uint32()
        self._ck(self.L.pt_debug_read_wide_bvh(self.h, None, C.byref(nn), None, C.byref(nt)), "pt_debug_read_wide_bvh")
        nodes, tris = np.zeros((nn.value, 20), np.uint32), np.zeros((nt.value, 12), np.uint32)
        self._ck(self.L.pt_debug_read_wide_bvh(self.h, abi.ptr(nodes), None, abi.ptr(tris), None), "pt_debug_read_wide_bvh")
        return nodes, tris

    def read_tiles(self, which=0, out=None):
        """This rank's pixels only, (num_slots, 4) float32 in slot order (see tile_pixels()): with N ranks each
        moves 1/N of the framebuffer."""
        if out is None:
            out = np.empty((self.num_slots(), 4), np.float32)
        self._ck(self.L.pt_read_tiles(self.h, abi.ptr(out), int(which)), "pt_read_tiles")
        return out

    def join(self):
        """Asynchronous: orders every Render enqueued so far before later work on stream() (renders run on
        private streams, two samples in flight)."""
        self._ck(self.L.pt_join(self.h), "pt_join")

    def set_allreduce_hook(self, fn):
        """fn(dev_ptr:int, count:int, stream:int) -> 0 on success; called once per guiding-training frame."""
        if fn is None:
            self._hook = abi.ALLREDUCE_HOOK(0)
        else:
            self._hook = abi.ALLREDUCE_HOOK(lambda user, p, n, s: int(fn(p or 0, n, s or 0)))
        self._ck(self.L.pt_set_allreduce_hook(self.h, self._hook, None), "pt_set_allreduce_hook")

    def peer_export(self):
        """cudaIpc handles (2 x 64 bytes: radiance-cache block, framebuffer-gather block) of this rank; gather them over all
        ranks for peer_connect."""
        buf = C.create_string_buffer(128)
        self._ck(self.L.pt_peer_export(self.h, buf), "pt_peer_export")
        return buf.raw

    def peer_connect(self, handles):
        """handles: the 128-byte handle blobs of all ranks in rank order.  Training frames then exchange the radiance-cache
        additions with the fused reduce + propagate kernel over NVLink peer memory instead of hook + propagate."""
        blob = b"".join(handles)
        self._ck(self.L.pt_peer_connect(self.h, blob, len(blob) // 128), "pt_peer_connect")

    def peer_disconnect(self):
        self._ck(self.L.pt_peer_disconnect(self.h), "pt_peer_disconnect")

    # -- statistics / stage hooks --
    def stats(self):
        st = abi.Stats()
        self._ck(self.L.pt_get_stats(self.h, C.byref(st)), "pt_get_stats")
        return st

    def reset_stats(self):
        self._ck(self.L.pt_reset_stats(self.h), "pt_reset_stats")

    STAGES = ("extend", "shade", "connect", "guiding", "other", "tail")

    def set_profiling(self, on):
        self._ck(self.L.pt_set_profiling(self.h, int(on)), "pt_set_profiling")

    def profile(self):
        """{stage: (device ms summed over launches, launches)} since set_profiling(True); call after Finish()."""
        ms = (C.c_double * len(self.STAGES))()
        n = (C.c_uint64 * len(self.STAGES))()
        self._ck(self.L.pt_get_profile(self.h, ms, n), "pt_get_profile")
        return {k: (ms[i], int(n[i])) for i, k in enumerate(self.STAGES)}

    def update_instances(self, instances, tlas):
        inst = np.ascontiguousarray(instances)
        tl = np.ascontiguousarray(tlas)
        self._ck(self.L.pt_update_instances(self.h, abi.ptr(inst), len(inst), abi.ptr(tl), len(tl)), "pt_update_instances")

    def rand_state(self):
        a, b = C.c_uint32(), C.c_uint32()
        self.L.pt_get_rand_state(self.h, C.byref(a), C.byref(b))
        return a.value, b.value

    def stream(self):
        return self.L.pt_cuda_stream(self.h)

    def debug_primary_rays(self, camera, rand_idx):
        out = np.zeros(self.cfg.width * self.cfg.height, abi.RAY)
        cam = np.ascontiguousarray(camera)
        self._ck(self.L.pt_debug_primary_rays(self.h, abi.ptr(cam), rand_idx, abi.ptr(out)), "pt_debug_primary_rays")
        return out

    def debug_trace(self, rays, any_hit=False):
        rays = np.ascontiguousarray(rays)
        out = np.zeros(len(rays), abi.HIT)
        self._ck(self.L.pt_debug_trace(self.h, abi.ptr(rays), len(rays), int(any_hit), abi.ptr(out)), "pt_debug_trace")
        return out

    def trace_static(self, rays, any_hit=False):
        """pt_trace_static: pt_debug_trace without any device allocation (scratch = the library's static device array;
        at most abi.STATIC_TRACE_MAX_RAYS rays).  No counterpart in the reference."""
        rays = np.ascontiguousarray(rays)
        out = np.zeros(len(rays), abi.HIT)
        self._ck(self.L.pt_trace_static(self.h, abi.ptr(rays), len(rays), int(any_hit), abi.ptr(out)), "pt_trace_static")
        return out

    def debug_trace_timed(self, rays, any_hit=False, repeats=5):
        rays = np.ascontiguousarray(rays)
        out = np.zeros(len(rays), abi.HIT)
        ms = C.c_float()
        self._ck(self.L.pt_debug_trace_resident(self.h, abi.ptr(rays), len(rays), int(any_hit), repeats, C.byref(ms), abi.ptr(out)),
                 "pt_debug_trace_resident")
        return out, ms.value

    def close(self):
        if self.h:
            self.L.pt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def build_bvh(triangles, tri_offset=0, device=0):
    """createBVHBinned (reference src/bvhBuilder.h:35-274) on the GPU: `triangles` = abi.TRIANGLE_V records (or an (n, 9)
    float array).  Returns (nodes [abi.BVH_NODE], order [uint32: input index per output position], device_ms, levels).
    The caller permutes its TriangleV / TriangleD arrays with `order` (permuteTriangles, bvhBuilder.h:17-27)."""
    L = load_library()
    if not (isinstance(triangles, np.ndarray) and triangles.dtype == abi.TRIANGLE_V):
        v = np.ascontiguousarray(triangles, np.float32).reshape(-1, 9)
        t = np.zeros(len(v), abi.TRIANGLE_V)
        t["v0"], t["v1"], t["v2"] = v[:, 0:3], v[:, 3:6], v[:, 6:9]
        triangles = t
    triangles = np.ascontiguousarray(triangles)
    n = len(triangles)
    nodes = np.zeros(max(1, 2 * n - 1), abi.BVH_NODE)
    order = np.zeros(n, np.uint32)
    used, levels, ms = C.c_uint32(), C.c_uint32(), C.c_float()
    r = L.pt_build_bvh(int(device), abi.ptr(triangles), n, int(tri_offset), abi.ptr(nodes), C.byref(used), abi.ptr(order), C.byref(ms), C.byref(levels))
    if r != 0:
        raise PtError("pt_build_bvh failed (%d): %s" % (r, (L.pt_build_bvh_last_error() or b"").decode()))
    return nodes[:used.value].copy(), order, ms.value, levels.value


def bounds_checked_build():
    """True when the loaded library was built with -DPT_BOUNDS_CHECK (pt_version says so)."""
    return b"bounds-checked" in load_library().pt_version()


def check_report():
    """(failures, code, index, limit) of the bounds-checked build since the last call; None for other builds."""
    if not bounds_checked_build():
        return None
    out = (C.c_uint32 * 4)()
    r = load_library().pt_debug_check_report(None, out)
    if r != 0:
        raise PtError("pt_debug_check_report failed (%d)" % r)
    return tuple(int(x) for x in out)


def host_wide_bvh(bvh2, triangles, optimal=True):
    """Host-only (no GPU): (nodes [n, 20] uint32, triangle records [m, 12] uint32, sah_cost, max_depth) of the 8-wide BVH the host
    converter cuts from one model's BVH2 (abi.BVH_NODE array) over the scene's triangle array (abi.TRIANGLE_V)."""
    L = load_library()
    bvh2, triangles = np.ascontiguousarray(bvh2), np.ascontiguousarray(triangles)
    nn, nt, cost, depth = C.c_uint32(), C.c_uint32(), C.c_double(), C.c_uint32()
    r = L.pt_host_wide_bvh(abi.ptr(bvh2), len(bvh2), abi.ptr(triangles), int(bool(optimal)), None, C.byref(nn), None, C.byref(nt), None, None)
    if r != 0:
        raise PtError("pt_host_wide_bvh failed (%d)" % r)
    nodes, tris = np.zeros((nn.value, 20), np.uint32), np.zeros((nt.value, 12), np.uint32)
    r = L.pt_host_wide_bvh(abi.ptr(bvh2), len(bvh2), abi.ptr(triangles), int(bool(optimal)), abi.ptr(nodes), C.byref(nn), abi.ptr(tris), C.byref(nt), C.byref(cost), C.byref(depth))
    if r != 0:
        raise PtError("pt_host_wide_bvh failed (%d)" % r)
    return nodes, tris, cost.value, depth.value
