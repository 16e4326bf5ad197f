#!/usr/bin/env python
"""bench.py -- Mrays/s (and spp/s) of the wavefront path tracer on the sibenik 1080p configuration
(BASELINE.json configs[2]: NEE + path guiding, MAX_RAY_DEPTH 8), tile-sharded over N GPUs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one converge sample of the whole 1920x1080 image (Pathtracer::Render(shouldClear=false),
reference src/pathtracer.h:255-301): primary rays, <= 8 x (extend, shade, connect), radiance-cache update
(+ all-reduce over ranks) + propagate, accumulate.  A ray = one entry processed by extend or connect.
`value` is timed on the device (CUDA events on the context's stream) with the scene resident in HBM;
`e2e` goes through the public Pathtracer API with host buffers: per step the camera goes in and the whole
accumulated image comes back device->host (asynchronously, ring of pinned buffers); `e2e.blocking` is the same
with an instances/TLAS upload, Finish and a blocking image read every step.
--impl reference times the CPU implementation of the same path on the host cores (bounded sample).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, DEPTH = 1920, 1080, 8
METRIC = "Mrays/s on sibenik 1080p (NEE + path guiding, depth 8)"
WORKLOAD = "sibenik-proxy (cathedral+statue+emissive cube+2 spheres, 158k tris) 1920x1080, NEE on, guiding on, MAX_RAY_DEPTH 8"
# identical in both arms (ours and --impl reference): what is measured, not how
CONFIG = {"workload": WORKLOAD, "resolution": "1920x1080", "max_ray_depth": 8, "nee": True, "path_guiding": True, "guiding_frames": 100,
          "cameras": "16-camera sweep inside the nave, one camera per step", "global_pixels": W * H,
          "l2": "per-step working set (rays+hits+state+sample cache, ~500 MB) exceeds the 126 MB L2"}
CPU_W, CPU_H = 1920, 1080    # bounded CPU sample: the same workload, a few samples (about 10-20 s on 16 cores)
CPU_SAMPLES = 16


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows if len(r) > 2 + i)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def build_scene():
    from cuda_pathtracer_b200 import scenes
    return scenes.sibenik_scene()


def camera_for(step, w, h):
    from cuda_pathtracer_b200 import scenes
    cams = scenes.camera_sweep("sibenik", w, h, 16)
    return cams[step % len(cams)]


def cpu_leg(scene_arrays, threads, samples=1):
    """Runs the CPU oracle on the bounded sample; returns (Mrays/s, per-ray traversal counters, seconds, rays)."""
    from cuda_pathtracer_b200 import abi
    from oracle.oracle_py import OracleLib
    cfg = abi.default_config(CPU_W, CPU_H, max_ray_depth=DEPTH, interactive_depth=1)
    ora = OracleLib(cfg)
    desc, keep = abi.make_scene_desc(scene_arrays)
    ora.upload_scene(desc, keep)
    ora.set_threads(threads)
    cam = camera_for(0, CPU_W, CPU_H)
    ora.render(cam, True)                   # clearing frame: syncs the toggles (pathtracer.h:242-243)
    ora.reset_stats()
    t0 = time.time()
    for _ in range(samples):
        ora.render(cam, False)
    dt = time.time() - t0
    st, cnt = ora.stats()
    rays = st.extend_rays + st.shadow_rays
    per_ray = {"ext_inner": cnt[0] / max(1, st.extend_rays), "ext_tris": cnt[1] / max(1, st.extend_rays),
               "conn_inner": cnt[3] / max(1, st.shadow_rays), "conn_tris": cnt[4] / max(1, st.shadow_rays),
               "ext_per_pixel": st.extend_rays / (CPU_W * CPU_H * samples), "shadow_per_pixel": st.shadow_rays / (CPU_W * CPU_H * samples)}
    ora.close()
    return rays / dt / 1e6, per_ray, dt, rays



def reference_gpu_leg(rays_per_sample, frames=40):
    """The reference's OWN device path (src/pathtracer.h + kernels.h, unmodified, built for sm_100a with the reference's
    flags, oracle/_ref/refgpu_shipped_1920x1080x8) on this GPU: the same scene/camera/depth, NEE + guiding, 100 training
    frames then `frames` timed converge samples with one Finish per frame as src/main.cpp:308-319 does."""
    import tempfile
    from oracle import refgpu
    exe = refgpu.binary_path("shipped", W, H, DEPTH)
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/refgpu_shipped_%dx%dx%d not built (needs /root/reference at build time)" % (W, H, DEPTH)}
    try:
        with tempfile.TemporaryDirectory() as tmp:
            sp = refgpu.sibenik_spec(tmp)
            cam = camera_for(0, W, H)
            sp.camera = (tuple(float(x) for x in np.ravel(cam["eye"])[:3]), tuple(float(x) for x in np.ravel(cam["viewDir"])[:3]), 1.5, 5.0, 0.01)
            blue = (np.random.default_rng(3).random((64, 64)) > 0.5).astype(np.uint8) * 255
            _, _, ms, _ = refgpu.run_reference(exe, sp, tmp, W, H, 1, 100, True, True, blue, timed_frames=frames, timeout=900)
        return {"ms_per_sample": ms, "samples_per_s": 1e3 / ms, "mrays_s": rays_per_sample / ms / 1e3, "timed_samples": frames,
                "what": "reference src/pathtracer.h + src/kernels.h unmodified, nvcc --use_fast_math -O3 (CMakeLists.txt:21-25) for sm_100a, GL interop -> plain cudaArray; camera 0 of the sweep; Mrays/s uses this run's rays per sample"}
    except Exception as e:                       # a baseline leg must never take the bench line down
        return {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:300])}


class _QuietStdout:
    """The reference's host code prints from inside (bvhBuilder.h:266 ...): keep the bench's stdout one JSON line."""

    def __enter__(self):
        import ctypes
        sys.stdout.flush()
        self.libc = ctypes.CDLL(None)
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *exc):
        self.libc.fflush(None)
        os.dup2(self.saved, 1)
        os.close(self.saved); os.close(self.null)


def reference_cpu_traversal_leg(threads, repeats=3):
    """The reference's OWN CPU traversal (traverseTopLevel<false>, src/kernels.h:277-387, compiled where it lies into
    oracle/_ref/libptref_1920x1080.so) over the primary rays of camera 0 of the bench workload, parallelised as the reference's
    CPU renderer does (src/raytracer.h:66, `omp parallel for schedule(dynamic)`), on the scene built by the reference's own
    createBVHBinned / BuildTopLevelBVH.  kind "reference" for the traversal stage; the whole-path CPU number stays the port."""
    try:
        from cuda_pathtracer_b200 import scenes
        from oracle import oracle_py
        if not oracle_py.ref_available(W, H):
            return {"unavailable": "oracle/_ref/libptref_%dx%d.so not built (needs /root/reference at build time)" % (W, H)}
        with _QuietStdout():
            return _reference_cpu_traversal(threads, repeats)
    except Exception as e:                       # a baseline leg must never take the bench line down
        return {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:300])}


def _reference_cpu_traversal(threads, repeats):
    if True:
        from cuda_pathtracer_b200 import scenes
        from oracle import oracle_py
        ref = oracle_py.RefLib(W, H)
        rs = oracle_py.RefScene(ref)
        t0 = time.time()
        m = rs.add_material(diffuse=(0.2, 0.2, 0.2))
        rs.add_object(rs.add_model_soup(scenes.cathedral_soup(1), material=m), pos=(0, 12, 0))      # scenes.sibenik_scene(), same soups
        rs.add_object(rs.add_model_soup(scenes.statue_soup(4), material=m))
        rs.add_object(rs.add_model_soup(scenes.cube_soup(), material=m), material=m, pos=(0, 3, 0))
        rs.add_sphere((-2, -1, -3), 2, m); rs.add_sphere((-2, -1, 3), 2, m)
        rs.finalize()
        build_s = time.time() - t0
        cam = camera_for(0, W, H)
        f = lambda k: [float(x) for x in np.ravel(cam[k])]
        rays = ref.primary_rays(f("eye")[:3], f("viewDir")[:3], f("d")[0], f("focalLength")[0], f("aperture")[0], 0)
        best = None
        for _ in range(repeats):
            t0 = time.time()
            hits = rs.trace(rays, False, threads)
            dt = time.time() - t0
            best = dt if best is None else min(best, dt)
        return {"mrays_s": len(rays) / best / 1e6, "rays": int(len(rays)), "seconds": best, "threads": threads, "kind": "reference",
                "hit_fraction": float(np.mean(hits["primitive_id"] != 0xFFFFFFFF)), "scene_build_s": build_s,
                "what": "reference traverseTopLevel<false> (closest hit) on the %dx%d primary rays of camera 0, best of %d" % (W, H, repeats)}


def reference_arm(args):
    """--impl reference: the CPU implementation of the path (the oracle port of the reference's kernels,
    kind "port": the reference has no CPU build of kernel_shade/connect) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    arrays = build_scene().arrays()
    for _ in range(min(args.warmup, 1)):
        cpu_leg(arrays, threads, 1)
    t0 = time.time()
    total_rays, total_s = 0, 0.0
    for _ in range(args.steps):
        _, _, dt, rays = cpu_leg(arrays, threads, 1)
        total_rays += rays
        total_s += dt
    v = total_rays / total_s / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total_s / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": CONFIG,
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port",
                             "sample": "oracle/pt_oracle.cpp, 1 converge sample at %dx%d per step, OpenMP %d threads" % (CPU_W, CPU_H, threads)},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "reference_traversal": reference_cpu_traversal_leg(threads), "wall_s": time.time() - t0}
    print(json.dumps(line), flush=True)


def gpu_arm(args):
    import zlib
    import torch
    import torch.distributed as dist
    from cuda_pathtracer_b200 import abi, api
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    arrays = build_scene().arrays()
    flags = abi.FLAG_SORT_RAYS if args.sort else 0
    cfg = abi.default_config(W, H, max_ray_depth=DEPTH, interactive_depth=1, device=local, rank=rank, world_size=world, flags=flags)
    pt = api.Pathtracer(arrays, cfg).Init()
    stream = torch.cuda.ExternalStream(pt.stream(), device=torch.device("cuda", local))
    exchange, peers = "none", False
    if world > 1:
        # radiance-cache exchange of the training frames: the library's fused reduce + propagate + broadcast kernel over NVLink
        # peer memory; --exchange nccl (or a failed peer mapping) selects the all-reduce hook + local propagate instead
        from cuda_pathtracer_b200.distributed import make_allreduce_hook, connect_peers
        exchange = "nccl all-reduce + propagate"
        if args.exchange == "peer" and connect_peers(pt, torch.device("cuda", local)):
            exchange, peers = "fused reduce+propagate over NVLink peer memory", True
        else:
            pt.set_allreduce_hook(make_allreduce_hook(torch.device("cuda", local)))   # NCCL sum over NVLink on the stream the sample runs on
    cams = [camera_for(i, W, H) for i in range(16)]
    warm = max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step = 0

    def restart():
        """Every measured pass starts from the same state: clearing frame (sampleIdx 0), then `warm` converge samples, so each
        pass covers the same samples -- and the same radiance-cache training phase -- as the timed region."""
        nonlocal step
        pt.Render(cams[0], shouldClear=True)
        step = 0
        for _ in range(warm):
            pt.Render(cams[step % 16]); step += 1
        pt.Finish()

    def timed_pass(n):
        nonlocal step
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(n):
            pt.Render(cams[step % 16]); step += 1
        pt.join(); b.record(stream); pt.Finish(); barrier()       # renders run on the library's private streams: join orders them before the end event
        t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- timed region: device time, scene + queues resident in HBM ----
    restart()
    pt.reset_stats()
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed_pass(args.steps)
    sampler.stop_flag = True
    st = pt.stats()
    timed_launches = int(st.kernel_launches)
    rays = torch.tensor([st.extend_rays + st.shadow_rays, st.extend_rays, st.shadow_rays], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(rays)
    total_rays, ext_rays, sh_rays = (float(x) for x in rays.tolist())
    value = total_rays / (ms * 1e-3) / 1e6
    rays_per_step = total_rays / args.steps
    # the same window again (20 steps are tens of milliseconds: one scheduler hiccup is several percent)
    windows = [value]
    for _ in range(4):
        restart()
        windows.append(total_rays / (timed_pass(args.steps) * 1e-3) / 1e6)
    # N > 1: every rank must hold the same radiance cache after the same frames (the exchange replicates it)
    exchange_check = None
    if world > 1:
        crc = zlib.crc32(pt.read_radiance_cache()["radianceCache"].tobytes())
        crcs = [None] * world
        dist.all_gather_object(crcs, crc)
        exchange_check = {"radiance_cache_crc32_per_rank": crcs, "replicated": len(set(crcs)) == 1}
    # ---- the two phases of the reference's frame loop, separately (pathtracer.h:292: the cache trains while sampleIdx < 100) ----
    phase_steps = max(10, min(50, args.steps))
    pt.set_guiding_frames(1 << 30)
    restart()
    training_ms = timed_pass(phase_steps) / phase_steps          # every frame: update buckets, exchange over ranks, propagate
    pt.set_guiding_frames(0)
    timed_pass(8)
    steady_ms = timed_pass(phase_steps) / phase_steps            # cache frozen, guiding still steers the samples; 4-8 samples in flight
    pt.set_guiding_frames(100)                                   # back to the reference's window for everything below
    # ---- per-stage device time of the same steps (separate pass: event pairs around every launch, samples serialised) ----
    restart()
    pt.set_profiling(True)
    pt.reset_stats()
    for _ in range(args.steps):
        pt.Render(cams[step % 16]); step += 1
    pt.Finish()
    prof = pt.profile()
    pst = pt.stats()
    pt.set_profiling(False)
    # ---- what the wide layout actually requests: node steps / triangle tests counted by the traversal kernels themselves ----
    count_steps = min(args.steps, 8)
    restart()
    pt.set_counting(True)
    pt.reset_stats()
    for _ in range(count_steps):
        pt.Render(cams[step % 16]); step += 1
    pt.Finish()
    tally, cst = pt.tally(), pt.stats()
    pt.set_counting(False)
    # ---- end to end through the public API with host buffers ----
    # N == 1: the whole accumulated image comes back every step.  N > 1 with connected peers: ONE framebuffer -- every rank's
    # tiles go over NVLink into rank 0's image ring, rank 0 ships the assembled (H, W, 4) image to its host.  N > 1 without
    # peers (--exchange nccl): every rank ships its own tiles.
    gathered = world > 1 and peers
    if world == 1 or gathered:
        img = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True).numpy() if rank == 0 else None
    else:
        img = torch.empty((pt.num_slots(), 4), dtype=torch.float32, pin_memory=True).numpy()
    inst, tlas = arrays["instances"], arrays["topBvh"]
    restart()
    pt.reset_stats()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        pt.update_instances(inst, tlas)            # host -> device: instances + TLAS (camera rides in the launch parameters)
        pt.Render(cams[step % 16]); step += 1
        if gathered:
            pt.wait_read(pt.gather_image_async(img))
        else:
            pt.Finish()
            if world == 1:
                pt.read_image(0, out=img)           # device -> host: the accumulated float4 image
            else:
                pt.read_tiles(0, out=img)           # device -> host: this rank's tiles of it
    pt.Finish()
    barrier()
    blocking_s = time.perf_counter() - t0
    bst = pt.stats()
    # streaming variant of the same loop (the headline e2e): the per-step host input of a converge frame is the camera
    # (the reference uploads nothing else unless the frame clears, pathtracer.h:239-246); every step's image comes back
    # asynchronously into a ring of 4 pinned buffers while the next samples render, and the host takes delivery of image
    # k-3 (pt_wait_read) before its buffer is reused.
    ring = [torch.empty(img.shape, dtype=torch.float32, pin_memory=True).numpy() if img is not None else None for _ in range(4)]
    tickets = []
    restart()
    pt.reset_stats()
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        pt.Render(cams[step % 16]); step += 1
        if gathered:
            tickets.append(pt.gather_image_async(ring[k % 4]))
        else:
            tickets.append(pt.read_image_async(ring[k % 4], tiles=(world > 1)))
        if k >= 3:
            pt.wait_read(tickets[k - 3])
    pt.Finish()
    barrier()
    e2e_s = time.perf_counter() - t0
    est = pt.stats()
    image_ok = None
    if rank == 0 and (world == 1 or gathered):      # the last image delivered is a full framebuffer: every pixel carries all samples
        last = ring[(args.steps - 1) % 4]
        image_ok = bool(np.all(last[..., 3] == float(1 + warm + args.steps)) and np.isfinite(last).all())
    # N > 1: the same loop with the framebuffer in process-shared HOST memory (pt_host_ring_*): every rank stores its own tiles
    # through its own PCIe link, rank 0's process takes delivery of whole (H, W, 4) frames in place -- rank 0's link no longer
    # carries all 33 MB of every frame.
    ring_s, ring_rays, ring_ok, ring_note, hr = None, None, None, None, None
    if gathered:
        from cuda_pathtracer_b200.distributed import HostRing
        try:
            hr = HostRing(pt, slots=4)                 # raises on EVERY rank when /dev/shm cannot hold the ring (decided by rank 0)
        except RuntimeError as ex:
            hr, ring_note = None, str(ex)
    if gathered and hr is not None:
        epochs = []
        restart()
        pt.reset_stats()
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            pt.Render(cams[step % 16]); step += 1
            epochs.append(pt.deliver_image_async())
            if k >= 3 and rank == 0:
                pt.host_ring_wait(epochs[k - 3]); pt.host_ring_release(epochs[k - 3])
        if rank == 0:
            for e in epochs[-3:]:
                last_view = pt.host_ring_wait(e)
                if e != epochs[-1]:
                    pt.host_ring_release(e)
        pt.Finish()
        barrier()
        ring_s = time.perf_counter() - t0
        rst = pt.stats()
        ring_rays = rst.extend_rays + rst.shadow_rays
        if rank == 0:
            ring_ok = bool(np.all(last_view[..., 3] == float(1 + warm + args.steps)) and np.isfinite(last_view).all())
            pt.host_ring_release(epochs[-1])
            del last_view
        hr.close()
    e2 = torch.tensor([est.extend_rays + est.shadow_rays, e2e_s, bst.extend_rays + bst.shadow_rays, blocking_s], dtype=torch.float64, device="cuda")
    ring_value = None
    if gathered and hr is not None:
        rr = torch.tensor([ring_rays], dtype=torch.float64, device="cuda"); dist.all_reduce(rr)
        rt = torch.tensor([ring_s], dtype=torch.float64, device="cuda"); dist.all_reduce(rt, op=dist.ReduceOp.MAX)
        ring_value = float(rr.item()) / float(rt.item()) / 1e6
    use_ring = False
    if world > 1:
        r2 = e2[[0, 2]].clone(); dist.all_reduce(r2)
        t2 = e2[[1, 3]].clone(); dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e2 = torch.stack([r2[0], t2[0], r2[1], t2[1]])
    e2e_blocking = float(e2[2].item()) / float(e2[3].item()) / 1e6
    e2e_value = float(e2[0].item()) / float(e2[1].item()) / 1e6
    d2h = W * H * 16 if (world == 1 or gathered) else pt.num_slots() * 16 * world
    if rank == 0:
        use_ring = bool(ring_value and ring_ok and ring_value > e2e_value)
        l2_16, l2_32 = pt.measure_gather_bandwidth(32 << 20, 16), pt.measure_gather_bandwidth(32 << 20, 32)     # L2 gather ceiling
        ref_gpu = reference_gpu_leg(total_rays / args.steps)
        peak, peak_src = read_peaks()
        cpu_v, per_ray, cpu_s, cpu_rays = cpu_leg(arrays, os.cpu_count() or 1, CPU_SAMPLES)
        # ALGORITHMIC bytes of one extend launch: per ray 32 (RayPacked R) + 16 (HitInfoPacked W)
        # + 64 * N_inner + 48 * N_tri, N_* counted by the CPU oracle on the reference's BVH2 for the same scene/camera/depth
        bytes_per_ext_ray = 48.0 + 64.0 * per_ray["ext_inner"] + 48.0 * per_ray["ext_tris"]
        ext_ms, ext_n = prof["extend"]
        # only the rays the k_trace<false> launches traced (bounces below the tail bounce); k_tail's are not in ext_ms
        achieved = (pst.wave_extend_rays * bytes_per_ext_ray) / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0
        # the bytes THIS layout requests for the same rays, counted by the kernel: 80 B per wide-node step, 48 B per triangle
        # test, 32 B per top-level node, + the 48 B/ray stream; all extension rays of the counted frames (every bounce)
        en, et, etop = tally["extend"]
        n_cext = max(1, int(cst.extend_rays))
        wide_bytes_per_ray = 48.0 + (80.0 * en + 48.0 * et + 32.0 * etop) / n_cext
        wide_achieved = (pst.wave_extend_rays * wide_bytes_per_ray) / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0
        ncu = {}
        tp = os.path.join(ROOT, "profiles", "traffic.json")      # per launch, from the committed ncu --set full capture of this command
        if os.path.exists(tp):
            ncu = json.load(open(tp)).get("k_trace<false>", {})
        tot_prof_ms = sum(v[0] for v in prof.values())
        ws = sorted(windows)
        line = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": CONFIG,
                "run": {"tiles": "32x32 round-robin over ranks", "exchange": exchange, "samples_in_flight": 4 if pt.num_slots() > 1200000 else 8,
                        "ray_sorting": bool(args.sort), "phase_of_timed_steps": "samples %d..%d after a clearing frame: the radiance cache trains while sampleIdx < 100 (update + exchange + propagate every frame)" % (1 + warm, warm + args.steps)},
                "value_windows": {"n": len(ws), "min": ws[0], "median": ws[len(ws) // 2], "max": ws[-1], "what": "the timed region repeated from the same start state"},
                "spp_per_s": args.steps / (ms * 1e-3), "rays_per_step": rays_per_step,
                "phases": {"training_ms_per_step": training_ms, "training_mrays_s": rays_per_step / training_ms / 1e3,
                           "steady_ms_per_step": steady_ms, "steady_mrays_s": rays_per_step / steady_ms / 1e3,
                           "note": "training: guiding_frames unbounded, every frame updates/exchanges/propagates the cache and cannot overlap its predecessor; steady: cache frozen, samples overlap"},
                "clocks": sampler.summary(),
                "e2e": {"value": e2e_value if not use_ring else ring_value, "unit": "Mrays/s", "h2d_bytes_per_step": 76, "d2h_bytes_per_step": int(d2h),
                        "phase": "same samples as the timed region (restart + %d warm-up + %d steps)" % (warm, args.steps),
                        "image_complete": image_ok if not use_ring else ring_ok,
                        "delivery": None if not gathered else {
                            "through_rank0_link": {"value": e2e_value, "image_complete": image_ok, "how": "pt_gather_image_async: tiles over NVLink into rank 0's HBM image ring, one D2H copy per frame through rank 0's PCIe link"},
                            "through_all_links": {"value": ring_value, "image_complete": ring_ok, "skipped": ring_note, "how": "pt_host_ring_* (distributed.HostRing): the framebuffer ring lives in process-shared host memory, every rank stores its own tiles through its own PCIe link, rank 0's process takes delivery of frame k-3 each step in place"},
                            "value_is": "through_all_links" if use_ring else "through_rank0_link"},
                        "mode": ("streaming: per step Render(camera) + " + ("pt_deliver_image_async: every rank's tiles through its own PCIe link into ONE (H,W,4) float4 framebuffer in process-shared host memory, taken in place by rank 0's process" if use_ring else
                                 "pt_gather_image_async: every rank's tiles over NVLink into rank 0's image ring, ONE (H,W,4) float4 image to rank 0's pinned host memory" if gathered else
                                 "pt_read_image_async of the whole accumulated image" if world == 1 else "pt_read_tiles_async: every rank ships its own tiles (no peer mapping)") +
                                 ", ring of 4 pinned buffers, the host waits for image k-3 each step; all %d images land inside the timed region" % args.steps),
                        "blocking": {"value": e2e_blocking, "h2d_bytes_per_step": int(inst.nbytes + tlas.nbytes + 76), "d2h_bytes_per_step": int(d2h),
                                     "mode": "per step update_instances (instances + TLAS H2D), Render, image delivered to the host before the next step: nothing overlaps"}},
                "gpu_launches": timed_launches,
                "roofline": {"bound": "hbm", "kernel": "k_trace<false> (extend)", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": ncu.get("dram_bytes_per_launch"), "peak_source": peak_src,
                             "what": "reference-layout bytes (SURVEY 8(d)): per ray 48 B stream + 64 B x BVH2 inner visits + 48 B x triangle tests as the CPU oracle counts them on the reference's BVH2; NOT a bandwidth utilisation of this kernel (its 8-wide compressed nodes request ~5x fewer bytes, mostly from L1): the kernel is issue/divergence bound",
                             "algorithmic_bytes_per_launch": pst.wave_extend_rays * bytes_per_ext_ray / max(1, ext_n), "bytes_per_ray": bytes_per_ext_ray,
                             "avg_launch_ms": ext_ms / max(1, ext_n), "share_of_step": ext_ms / tot_prof_ms if tot_prof_ms else None,
                             "l2_gather_peak_gbs": {"16B": l2_16, "32B": l2_32, "how": "pt_measure_gather_bandwidth: random reads over a 32 MB (L2-resident) array, best of 5"},
                             "frac_of_l2_gather_peak": achieved / l2_32 if l2_32 else None,
                             "stream": {"bytes_per_ray": 48.0, "achieved_gbs": (pst.wave_extend_rays * 48.0) / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0,
                                        "frac_of_hbm_peak": (pst.wave_extend_rays * 48.0) / (ext_ms * 1e-3) / 1e9 / peak if ext_ms > 0 else 0.0},
                             "wide_layout": {"bytes_per_ray": wide_bytes_per_ray, "node_steps_per_ray": en / n_cext, "triangle_tests_per_ray": et / n_cext, "top_level_steps_per_ray": etop / n_cext,
                                             "achieved_gbs": wide_achieved, "frac_of_l2_gather_peak": wide_achieved / l2_32 if l2_32 else None, "frac_of_hbm_peak": wide_achieved / peak,
                                             "what": "bytes the 8-wide layout requests, counted by the traversal kernel itself (pt_set_counting: 80 B per node step, 48 B per triangle test, 32 B per top-level node, + 48 B/ray stream) over %d frames" % count_steps},
                             "ncu": {k: ncu.get(k) for k in ("lts_bytes_per_launch", "issue_active_pct", "lanes_per_inst", "l1_hit_pct", "dram_pct_of_peak", "source") if k in ncu},
                             "target_50pct_bandwidth": "not applicable: issue/divergence bound (see DESIGN.md 5)"},
                "stage_ms_per_step": {k: v[0] / args.steps for k, v in prof.items()},
                "exchange_check": exchange_check,
                "reference_on_this_gpu": ref_gpu,
                "reference_cpu_traversal": reference_cpu_traversal_leg(os.cpu_count() or 1),
                "cpu_baseline": {"value": cpu_v, "unit": "Mrays/s", "cores": os.cpu_count() or 1, "kind": "port",
                                 "sample": "oracle/pt_oracle.cpp: %d converge samples of the same scene/camera/depth at %dx%d (%.0f rays, %.1f s)" % (CPU_SAMPLES, CPU_W, CPU_H, cpu_rays, cpu_s)}}
        print(json.dumps(line), flush=True)
    if world > 1 and peers:
        from cuda_pathtracer_b200.distributed import disconnect_peers
        disconnect_peers(pt)
    pt.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sort", action="store_true", help="ray sorting between shade and trace (pt_set_ray_sorting); results identical")
    ap.add_argument("--exchange", default=os.environ.get("PTB200_EXCHANGE", "peer"), choices=["peer", "nccl"],
                    help="N>1: radiance-cache exchange of the training frames (peer = fused kernel over NVLink peer memory)")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
