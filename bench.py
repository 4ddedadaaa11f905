#!/usr/bin/env python
"""bench.py — Mrays/s of the volumetric ray-march hot path on B200 (one process per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c1|c3|c4] [--impl volrt|reference]

A "step" is one frame: one pass of the march over every pixel of the configured target.
Default workload = BASELINE.json configs[1]: 1920x1080 procedural fBm cloud, 128 primary steps per
ray + 6-step sun light-march, synthetic (no volume input at all).
  value      whole-job Mrays/s (W*H*frames / device time of the march kernels, CUDA events on the
             launching stream, L2 flushed between frames, max over ranks)
  e2e        the same through the public C ABI with HOST buffers: vr_set_params (UBO bytes, H2D) +
             vr_render into pinned host memory (frame D2H) per step, wall clock around the synchronous calls
  roofline   dominant kernel (march) against the FP32-FMA peak measured in the same process
  cpu_baseline  the CPU oracle (a port of frag.glsl; the reference has no CPU path) on a bounded,
             strided row sample of the same frame, all host threads
N > 1: the frame is cut into cost-balanced row bands (vr_row_bands), one band per GPU; each GPU's
march kernel stores its pixels straight into GPU 0's frame through an NVLink peer mapping
(CUDA IPC), so the gather is fused into the march ("scaling": "strong").
--impl reference: the oracle (the only CPU implementation of the path there is) timed on the host.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SUN = np.array([1.0, 1.0, 2.0], np.float32) / np.float32(np.sqrt(6.0))

WORKLOADS = {
    # name: (W, H, description, vr_params keyword overrides)
    "c1": (1280, 720, "default frag.glsl scene, 128^3 RGBA8 baked volume, 128 steps, texture-unit trilinear", dict()),
    "c2": (1920, 1080, "1920x1080 procedural fBm cloud (5-octave Perlin), 128 primary steps + 6-step sun light-march",
           dict(density_mode=2, max_steps=128, sigma_t=8.0, light_steps=6, light_step_len=2 * 4.0 / 128, ambient=0.15,
                sun_dir=SUN, sun_color=[1.0, 0.95, 0.85])),
    "c3": (3840, 2160, "3840x2160 procedural fBm cloud, 256 primary steps + 8-step sun light-march",
           dict(density_mode=2, max_steps=256, sigma_t=8.0, light_steps=8, light_step_len=2 * 4.0 / 256, ambient=0.15,
                sun_dir=SUN, sun_color=[1.0, 0.95, 0.85])),
    "c4": (1920, 1080, "512^3 fp16 smoke grid (device-baked fBm field), trilinear, 1 fetch/step, 512 steps",
           dict(density_mode=1, max_steps=512, sample_scale=4.0)),
    "c4s128": (1920, 1080, "512^3 fp16 smoke grid (device-baked fBm field), trilinear, 1 fetch/step, 128 steps",
               dict(density_mode=1, max_steps=128, sample_scale=16.0)),
    "c4s1024": (1920, 1080, "512^3 fp16 smoke grid (device-baked fBm field), trilinear, 1 fetch/step, 1024 steps",
                dict(density_mode=1, max_steps=1024, sample_scale=2.0)),
    "c5": (1920, 1080, "64 orbit views over a 1024^3 fp16 smoke grid (device-baked, thresholded), 512 steps, compositing with "
                       "empty-space skipping (8^3 bricks) and early termination (T < 1/512), views sharded over the GPUs",
           dict(density_mode=1, max_steps=512, sample_scale=1.0, sigma_t=400.0, early_term_eps=1.0 / 512, skip_empty=1,
                ambient=1.0, sun_color=[0.0, 0.0, 0.0])),
    "bake": (0, 0, "start-up bake of the 128^3 RGBA8 noise volume (TestMain.cpp:43-92: cellular x2, Perlin, simplex; min/max "
                   "normalisation, quirk Q1 included) — 4 fields x 128^3 noise points per step", dict()),
}
C5_VIEWS, C5_GRID, C5_THRESHOLD = 64, 1024, 0.07
BAKE_SIZE, BAKE_FIELDS = 128, 4
# algorithmic FP32 work of the bake per voxel: 2 cellular fields (27 cells x 14 flop) + Perlin (60) + simplex (4 corners x 20)
BAKE_FLOP_PER_VOXEL = 2 * 27 * 14.0 + 60.0 + 80.0
# algorithmic FP32 work: one Perlin octave = 60 flop, +20 per fBm sample for the
# remap/falloff, 26 flop per step for the integrator; FMA = 2 flop.
FLOP_PER_OCTAVE, FLOP_PER_SAMPLE, FLOP_PER_STEP = 60.0, 20.0, 26.0


def config_for(workload, world, layout, has_volume, single_process=False):
    """`config` of a bench line; the product arm and the reference arm print the very same dict for the same command"""
    W, H, desc, _ = WORKLOADS[workload]
    part = "single GPU"
    if world > 1:
        part = ("cost-balanced row bands, peer stores into GPU0's frame over NVLink; one process, one context over all devices"
                if single_process else "cost-balanced row bands, peer stores into GPU0's frame over NVLink (CUDA IPC)")
    return {"workload": "%s: %s" % (workload, desc), "width": W, "height": H,
            "l2": "flushed between frames (256 MiB written, then read back so that L2 is left clean); the procedural workloads read no input",
            "partition": part, "volume_layout": layout if has_volume else None}


def kernel_source_sha16():
    """identity of the kernels' source: roofline.traffic is only reported when the committed ncu capture was taken on this source"""
    import glob
    import hashlib
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "volumetricrenderer_b200", "csrc", "*.cu*")) +
                    glob.glob(os.path.join(ROOT, "volumetricrenderer_b200", "csrc", "*.h"))):
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:16]


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="volrt", choices=["volrt", "reference"])
    ap.add_argument("--layout", default="texture", choices=["texture", "linear", "brick"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the oracle sample")
    ap.add_argument("--no-extra", action="store_true", help="c2 only: skip the extra.c3 measurement")
    ap.add_argument("--single-process", action="store_true",
                    help="--gpus N without torchrun: ONE process, ONE context over N devices (vr_create_info.device_count)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU while the timed region runs (NVML)"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.nv = None

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.004)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except (ValueError, IndexError):
            return local
    return local


def oracle_params_for(workload):
    """vo_params of the workload (oracle side; used only by the cpu_baseline / reference legs)"""
    from oracle import oracle
    from volumetricrenderer_b200 import scene
    W, H, _, kw = WORKLOADS[workload]
    ubo0, ubo1 = scene.default_scene()
    scal = {k: v for k, v in kw.items() if not hasattr(v, "__len__")}
    vecs = {k: v for k, v in kw.items() if hasattr(v, "__len__")}
    p = oracle.params_default(**scal)
    oracle.set_params(p, **vecs)
    oracle.set_scene(p, ubo0, ubo1)
    return p


def oracle_volume_for(workload):
    from oracle import oracle
    if workload == "c1":
        vol = oracle.ref_bake_default(128)[0] if oracle.have_ref() else oracle.bake_default(128)
        return oracle.make_volume(vol, oracle.VOL_RGBA8)
    if workload.startswith("c4"):
        p = oracle.params_default()
        vol = oracle.bake_smoke_r16f(128, p, 0.0)          # CPU sample uses a 128^3 stand-in of the field
        return oracle.make_volume(vol, oracle.VOL_R16F)
    return None, None


def time_oracle(workload, target_seconds, threads=0):
    """Oracle on a bounded strided row sample -> dict(value Mrays/s, cores, sample, seconds)."""
    from oracle import oracle
    W, H, _, _ = WORKLOADS[workload]
    if workload == "c1" and oracle.have_ref_glsl():
        # config 1 has reference semantics, and the reference's own shaders compile here (oracle/glsl_shim.h): time THEM —
        # vert.glsl -> fixed-function rasteriser (oracle/ref_raster.cpp) -> frag.glsl on the reference's own FastNoise2 bake
        from volumetricrenderer_b200 import scene
        u0, u1 = scene.default_scene()
        vol = oracle.ref_bake_default(128)[0]
        cores = threads or oracle.lib().vo_hw_threads()
        oracle.ref_draw(u0, u1, vol, W, H, threads)          # warm the pages
        reps, t0 = 0, time.perf_counter()
        while True:
            _img, cov, _fp = oracle.ref_draw(u0, u1, vol, W, H, threads)
            reps += 1
            dt = time.perf_counter() - t0
            if dt >= min(target_seconds, 20.0) or reps >= 200:
                break
        return {"value": reps * W * H / dt / 1e6, "unit": "Mrays/s", "cores": int(cores), "kind": "reference",
                "sample": "%d whole %dx%d frames of the reference's vert.glsl + frag.glsl compiled as C++ (oracle/_ref), %.2f s, %d covered px per frame"
                          % (reps, W, H, dt, int(cov.sum())),
                "seconds": dt, "msamples_per_s": None}
    p = oracle_params_for(workload)
    vol, keep = oracle_volume_for(workload)
    cores = threads or oracle.lib().vo_hw_threads()
    probe = np.arange(8, H, 16, dtype=np.int32)              # calibration pass
    t0 = time.perf_counter()
    oracle.render_rows(vol, p, W, H, probe, threads)
    t_probe = time.perf_counter() - t0
    per_row = t_probe / len(probe)
    stride = int(max(1, min(64, np.ceil(H * per_row / max(target_seconds, 0.1)))))
    rows = np.arange(stride // 2, H, stride, dtype=np.int32)
    t0 = time.perf_counter()
    _, st = oracle.render_rows(vol, p, W, H, rows, threads)
    dt = time.perf_counter() - t0
    mrays = len(rows) * W / dt / 1e6
    return {"value": mrays, "unit": "Mrays/s", "cores": int(cores), "kind": "port",
            "sample": "%d of %d rows (every %d-th) of the %s frame, %.2f s, %d steps" % (len(rows), H, stride, workload, dt, st["steps"]),
            "seconds": dt, "msamples_per_s": st["steps"] / dt / 1e6}


# ---------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the CPU implementation of the path (the oracle port — the reference itself
    has only GLSL for it and no graphics stack exists here) on all host threads, rank 0 only."""
    if rank != 0:
        return
    W, H, desc, _ = WORKLOADS[args.workload]
    per_step = max(2.0, min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup)))
    vals, last = [], None
    for i in range(args.warmup + args.steps):
        last = time_oracle(args.workload, per_step)
        if i >= args.warmup:
            vals.append(last)
    secs = sum(v["seconds"] for v in vals)
    # rays of the sample / time of the sample, aggregated over the timed steps
    mrays = float(np.mean([v["value"] for v in vals]))
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": W * H / (mrays * 1e6) * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_for(args.workload, max(1, args.gpus), args.layout, args.workload == "c1" or args.workload.startswith("c4"),
                             args.single_process),
        "note": "ms_per_step = full-frame time extrapolated from the timed row sample; the CPU arm runs on rank 0's host cores whatever --gpus says",
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": last["cores"], "kind": last["kind"], "sample": last["sample"]},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "sample_seconds": secs,
    }
    print(json.dumps(line), flush=True)


def setup_volume(r, workload, layout):
    from volumetricrenderer_b200 import capi
    lay = {"texture": capi.LAYOUT_TEXTURE, "linear": capi.LAYOUT_LINEAR, "brick": capi.LAYOUT_BRICK}[layout]
    if layout == "brick" and not workload.startswith("c4"):
        lay = capi.LAYOUT_TEXTURE
    if workload == "c1":
        r.generate_volume(layout=lay)                                   # the reference's bake, on the device
        return 128 ** 3 * 4
    if workload.startswith("c4"):
        r.generate_volume(size=512, format=capi.VOL_R16F, layout=lay)   # fBm smoke field, fp16
        return 512 ** 3 * 2
    return 0


def run_volrt(args, rank, world, local):
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    line = measure_workload(args, args.workload, rank, world, local, dist, args.steps, full=True)
    if args.workload == "c2" and not args.no_extra:
        # BASELINE.json configs[2] (4K, 256 + 8) on the same record, at every N: the configuration north_star wants
        # "near-linear" on.  Fewer frames (a 4K frame is ~10x the work), same timing rules.
        extra = measure_workload(args, "c3", rank, world, local, dist, max(5, args.steps // 3), full=False)
        # BASELINE.json configs[4] (64 views over a 1024^3 fp16 grid, skipping + early termination, views sharded over the GPUs)
        extra5 = measure_views(args, rank, world, local, dist, max(3, args.steps // 6))
        # BASELINE.json configs[0]: the reference's own scene and shader semantics (1280x720, baked 128^3 RGBA8 volume); its CPU
        # baseline is the reference's shaders compiled as C++ (kind "reference")
        extra1 = measure_workload(args, "c1", rank, world, local, dist, args.steps, full=False)
        if rank == 0:
            if world == 1 and not args.no_cpu_baseline:
                try:
                    cb = time_oracle("c1", min(args.cpu_seconds, 5.0))
                    extra1["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
                except Exception as e:
                    extra1["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "reference", "sample": "failed: %r" % (e,)}
            line["extra"] = {"c1": extra1, "c3": extra, "c5": extra5}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def measure_workload(args, workload, rank, world, local, dist, K, full):
    """one workload on `world` ranks -> the bench line (rank 0; None elsewhere).  full: with roofline notes, cpu_baseline"""
    import volumetricrenderer_b200 as vr
    from volumetricrenderer_b200 import capi, scene
    W, H, desc, kw = WORKLOADS[workload]
    Wm = max(args.warmup, 3)
    ubo0, ubo1 = scene.default_scene()
    group = args.single_process and args.gpus > 1 and world == 1      # one process, one context over args.gpus devices
    n_gpus = args.gpus if group else world
    # N > 1: GPU 0 holds two frames, so the gather of frame k+1 (peer stores) overlaps the D2H copy of frame k
    r = vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, device=local, max_views=2 if world > 1 else 1,
                    devices=list(range(args.gpus)) if group else None)
    vol_bytes = setup_volume(r, workload, args.layout)
    r.set_params(ubo0, ubo1, **kw)

    # ---- partition: cost-balanced row bands, peer-mapped frame of rank 0
    from volumetricrenderer_b200 import multigpu
    bg = multigpu.BandGather(dist, rank, world, H, mode="peer")
    row_cost = r.row_cost() if world > 1 else None
    if world > 1:
        bg.plan(row_cost=row_cost)
    mapped = bg.connect_peer(r) if world > 1 else None

    def march(flags=0):
        rows = bg.rows
        if rows[1] <= rows[0]:
            return
        if mapped is not None:
            r.render_device(mapped, rows=rows, flags=capi.RENDER_DIRECT | flags)
        else:
            r.render(rows=rows, to_host=False, flags=flags)

    def barrier():
        r.wait()
        if dist is not None:
            import torch
            dist.barrier(device_ids=[local])
            torch.cuda.synchronize()

    # ---- one counted pass (device counters): work actually done per frame (a counted pass runs on one device)
    march(capi.RENDER_STATS)
    st = r.stats()
    counts = np.array([st["covered"], st["steps"], st["light_samples"], st["early_terminated"]], np.float64)
    if dist is not None:
        import torch
        t = torch.tensor(counts, device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        counts = t.cpu().numpy()
    covered, steps, light_samples = int(counts[0]), int(counts[1]), int(counts[2])

    # ---- device-timed region: K frames, L2 flushed between frames, CUDA events on the launching stream
    for _ in range(Wm + (6 if group else 0)):          # a group context re-cuts its bands from the measured times as it goes
        r.flush_l2(); march(); r.wait()
    if world > 1:
        # feedback re-cut of the bands from measured band times (a renderer does this every frame; here: during
        # warm-up only, the cut is frozen for the timed frames)
        # Kernel time is not linear in rows (waves, fixed costs), so the iteration hovers around the balanced cut instead of
        # landing on it: of the cuts tried, the one whose slowest band was fastest is kept (every rank sees the same times).
        best = (float("inf"), None)
        for _ in range(8):
            r.flush_l2(); march(); r.wait()
            bg.rebalance(row_cost, r.stats()["gpu_ms"] if bg.rows[1] > bg.rows[0] else 0.0)
            if max(bg.last_times) < best[0]:
                best = (max(bg.last_times), bg.last_bounds)
        bg.bounds = list(best[1])
    rows = bg.rows
    sampler = ClockSampler(physical_gpu_index(local))
    barrier()
    sampler.start()
    wall0 = time.perf_counter()
    ms, launches = [], 0
    for _ in range(K):
        r.flush_l2()
        march()
        st_k = r.stats() if rows[1] > rows[0] else {"gpu_ms": 0.0, "kernel_launches": 0}
        ms.append(st_k["gpu_ms"])
        launches += st_k["kernel_launches"]       # one march per device that had rows to render
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop()
    dev_ms = float(np.sum(ms))
    if dist is not None:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, dev_ms / K)
        if rank == 0:
            print("bands %s  ms per rank %s" % (bg.bounds, ["%.3f" % x for x in per_rank]), file=sys.stderr)
        import torch
        t = torch.tensor([dev_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
    ms_per_step = dev_ms / K
    value = W * H / (ms_per_step * 1e-3) / 1e6

    # ---- N > 1: is the gathered frame the right frame?  Rank 0 renders the whole frame alone and compares it with
    # the frame its peers have just stored their bands into (a band may take another lanes-per-ray / depth-segment mapping
    # than the full frame: re-association only, so the bar is 1/255)
    gather_check = None
    if group:
        gathered = r.render()[0].copy()
        with vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, device=0) as solo:
            setup_volume(solo, workload, args.layout)
            solo.set_params(ubo0, ubo1, **kw)
            alone = solo.render()[0].copy()
        d = np.abs(gathered.astype(np.int16) - alone.astype(np.int16))
        gather_check = {"max_abs8": int(d.max()), "pixels_differing": int((d.max(axis=-1) > 0).sum()),
                        "nonzero_pixels": int((alone[..., :3].max(axis=-1) > 0).sum()), "vs": "the same frame rendered by a single-GPU context"}
    if world > 1:
        march(); barrier()
        if rank == 0:
            gathered = r.render(flags=capi.RENDER_NO_MARCH)[0].copy()
            alone = r.render()[0].copy()
            d = np.abs(gathered.astype(np.int16) - alone.astype(np.int16))
            gather_check = {"max_abs8": int(d.max()), "pixels_differing": int((d.max(axis=-1) > 0).sum()),
                            "nonzero_pixels": int((alone[..., :3].max(axis=-1) > 0).sum()), "vs": "the same frame rendered by GPU 0 alone"}
        barrier()

    # ---- end to end through the public API with host buffers
    # every step: vr_set_params (the UBO bytes the reference host writes) + vr_render into page-locked
    # host memory (frame D2H).  N == 1: two frames in flight (VR_RENDER_ASYNC), as the reference's
    # frame loop has (MAX_FRAMES_IN_FLIGHT = 2, VulkanRenderer.cpp:13): frame k+1 marches while frame k is
    # copied out; all K frames are complete in host memory when the clock stops.  The synchronous
    # per-frame latency (set_params + render + copy, nothing overlapped) is reported beside it.
    hosts = [r.new_frame_buffer(pinned=True), r.new_frame_buffer(pinned=True)] if rank == 0 else [None, None]

    def e2e_step(i, flags=0):
        r.set_params(ubo0, ubo1, **kw)                     # UBO bytes -> device (H2D)
        if world == 1:
            r.render(out=hosts[i & 1], flags=flags)        # march + frame D2H into pinned memory
        else:
            march()
            barrier()                                      # every band has landed in GPU 0's frame
            if rank == 0:
                r.render(out=hosts[i & 1], flags=capi.RENDER_NO_MARCH)

    for i in range(3):
        e2e_step(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        e2e_step(i)
    barrier()
    sync_s = (time.perf_counter() - t0) / K
    e2e_s = sync_s
    e2e_check = None
    if world > 1:
        # two frames in flight across the GPUs, completed on the device (include/volrt.h: vr_gather_*): every rank marches
        # frame i into frame slot i & 1 of GPU 0 (peer stores; rank 0 into its own memory) and bumps the slot's arrival
        # counter behind its stores; GPU 0's copy stream waits for all the bands, copies the slot to the host and releases
        # it; a rank's render stream waits for that release before it marches into the slot again.  No host barrier per
        # frame: the hosts only enqueue, frame i+1 gathers while frame i leaves over PCIe.
        own_ptr, _pitch, vstride = r.frame_info()
        hosts2 = r.new_frame_buffer(views=2, pinned=True) if rank == 0 else None
        base = mapped if mapped is not None else own_ptr
        frame0 = mapped                                    # None on rank 0
        counter = [0]                                      # frames sent through the protocol so far (slot use = counter // 2)

        def e2e_step_mg(_i):
            i = counter[0]
            counter[0] += 1
            slot, releases, arrivals = multigpu.gather_ticket(i, world, 2)
            r.set_params(ubo0, ubo1, **kw)
            r.gather_acquire(frame0, 2, slot, releases)
            if rows[1] > rows[0]:
                r.render_device(base + slot * vstride, rows=rows, flags=capi.RENDER_DIRECT | capi.RENDER_ASYNC)
            r.gather_signal(frame0, 2, slot)
            if rank == 0:
                r.gather_collect(slot, arrivals, hosts2[slot])

        for i in range(4):
            e2e_step_mg(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(K):
            e2e_step_mg(i)
        barrier()                                          # every frame is in host memory
        e2e_s = (time.perf_counter() - t0) / K
        if rank == 0:                                      # both slots hold the frame the one-at-a-time path delivered
            e2e_check = bool(np.array_equal(hosts2[0], hosts[0][0]) and np.array_equal(hosts2[1], hosts[0][0]))
    if world == 1:
        for i in range(3):
            e2e_step(i, capi.RENDER_ASYNC)
        barrier()
        t0 = time.perf_counter()
        for i in range(K):
            e2e_step(i, capi.RENDER_ASYNC)
        barrier()                                          # vr_wait: every frame is in host memory
        e2e_s = (time.perf_counter() - t0) / K
        e2e_check = bool(np.array_equal(hosts[0], hosts[1]))
    e2e = {"value": W * H / e2e_s / 1e6, "unit": "Mrays/s", "ms_per_frame": e2e_s * 1e3,
           "sync_ms_per_frame": sync_s * 1e3, "frames_in_flight": 2,
           "h2d_bytes_per_step": 192 + 140, "d2h_bytes_per_step": W * H * 4, "frames_identical": e2e_check}

    line = None
    if rank == 0:
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": n_gpus, "steps": K, "warmup": Wm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_for(workload, n_gpus, args.layout, bool(vol_bytes), group),
            "e2e": e2e, "gpu_launches": launches,
            "clocks": clocks, "wall_ms_per_step_incl_flush": wall / K * 1e3,
            "covered_rays": covered, "steps_per_frame": steps, "light_samples_per_frame": light_samples,
            "msamples_per_s": (steps + light_samples) / (ms_per_step * 1e-3) / 1e6,
        }
        line["covered_mrays_per_s"] = covered / (ms_per_step * 1e-3) / 1e6      # rays that actually march (the rest are cleared)
        if gather_check is not None:
            line["gather_check"] = gather_check
        line["roofline"] = roofline(r, args, workload, kw, steps, light_samples, ms_per_step, W, H, vol_bytes, n_gpus)
        line["ray_split"] = os.environ.get("VOLRT_RAY_SPLIT_COMPOSITE", "auto")
        if full and not args.no_cpu_baseline and world == 1:
            try:
                cb = time_oracle(workload, args.cpu_seconds)
                line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
            except Exception as e:                          # the baseline is a report, not the product
                line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": 0, "kind": "port", "sample": "failed: %r" % e}
    bg.disconnect_peer(r)
    barrier()
    r.close()
    return line


def run_views(args, rank, world, local):
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    line = measure_views(args, rank, world, local, dist, args.steps)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def measure_views(args, rank, world, local, dist, K):
    """C5: a batch of independent views, sharded contiguously over the ranks — or, with --single-process, over the devices of
    one group context (no exchange: every GPU stores its own views).  value = W*H*views / max-over-GPUs device time."""
    import volumetricrenderer_b200 as vr
    from volumetricrenderer_b200 import capi, scene, multigpu
    group = args.single_process and args.gpus > 1 and world == 1
    n_gpus = args.gpus if group else world
    W, H, desc, kw = WORKLOADS["c5"]
    Wm = max(args.warmup, 3)
    v0, v1 = multigpu.shard_views(C5_VIEWS, world)[rank]
    nv = v1 - v0
    views = scene.orbit_views(C5_VIEWS)[v0:v1]
    r = vr.Renderer(W, H, capi.OUT_RGBA8_UNORM, device=local, max_views=max(nv, 1), devices=list(range(args.gpus)) if group else None)
    grid = int(os.environ.get("VOLRT_C5_GRID", C5_GRID))
    r.generate_volume(size=grid, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE, threshold=C5_THRESHOLD)
    vol_bytes = grid ** 3 * 2
    r.set_params(view_ubos=views, **kw)

    def barrier():
        r.wait()
        if dist is not None:
            import torch
            dist.barrier(device_ids=[local])
            torch.cuda.synchronize()

    r.render(to_host=False, flags=capi.RENDER_STATS)
    st = r.stats()
    counts = np.array([st["covered"], st["steps"], st["early_terminated"], st["skipped_steps"]], np.float64)
    if dist is not None:
        import torch
        t = torch.tensor(counts, device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        counts = t.cpu().numpy()
    for _ in range(Wm):
        r.flush_l2(); r.render(to_host=False); r.wait()
    sampler = ClockSampler(physical_gpu_index(local))
    barrier()
    sampler.start()
    ms = []
    for _ in range(K):
        r.flush_l2()
        r.render(to_host=False)
        ms.append(r.stats()["gpu_ms"])
    barrier()
    clocks = sampler.stop()
    dev_ms = float(np.sum(ms))
    if dist is not None:
        import torch
        t = torch.tensor([dev_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
    ms_per_step = dev_ms / K
    value = W * H * C5_VIEWS / (ms_per_step * 1e-3) / 1e6
    # end to end: the view block (H2D) + every view of the shard read back into page-locked memory, two batches in flight
    hosts = [r.new_frame_buffer(views=nv, pinned=True), r.new_frame_buffer(views=nv, pinned=True)]
    def e2e_step(i):
        r.set_params(view_ubos=views, **kw)
        r.render(out=hosts[i & 1], flags=capi.RENDER_ASYNC)
    for i in range(2):
        e2e_step(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        e2e_step(i)
    barrier()
    e2e_s = (time.perf_counter() - t0) / K
    line = None
    if rank == 0:
        covered, steps, early, skipped = (int(x) for x in counts)
        peaks = r.measure_peaks()
        t = ms_per_step * 1e-3
        fetches = steps - skipped
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": n_gpus, "steps": K, "warmup": Wm,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": "c5: %s" % desc, "width": W, "height": H, "views": C5_VIEWS, "grid": grid,
                           "l2": "flushed between batches; the %d MiB grid exceeds L2" % (vol_bytes >> 20),
                           "partition": ("contiguous view shards over the devices of one context, peer stores into GPU0's frames" if group
                                         else "contiguous view shards, no exchange")},
                "e2e": {"value": W * H * C5_VIEWS / e2e_s / 1e6, "unit": "Mrays/s", "ms_per_frame": e2e_s * 1e3, "frames_in_flight": 2,
                        "h2d_bytes_per_step": 336 * C5_VIEWS, "d2h_bytes_per_step": W * H * 4 * C5_VIEWS},
                "gpu_launches": K, "clocks": clocks, "covered_rays": covered, "steps_per_frame": steps,
                "early_terminated_frac": early / max(covered, 1), "skipped_step_frac": skipped / max(steps, 1),
                "roofline": {"bound": "tex", "achieved": fetches / t / 1e9, "peak": peaks["tex_gfetch"] * n_gpus, "unit": "Gfetch/s",
                             "frac": fetches / t / 1e9 / (peaks["tex_gfetch"] * n_gpus) if peaks["tex_gfetch"] else None,
                             "note": "fetches = steps not skipped by the brick grid, before early termination (upper bound of the executed ones)",
                             "traffic": None, "peaks_measured_here": peaks}}
    barrier()
    r.close()
    return line


def time_cpu_bake(repeats):
    """the reference's own bake (FastNoise2 compiled from its sources, oracle/_ref) when present, else the oracle port"""
    from oracle import oracle
    n, pts = BAKE_SIZE, BAKE_FIELDS * BAKE_SIZE ** 3
    use_ref = oracle.have_ref()
    ts = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        if use_ref:
            oracle.ref_bake_default(n)
        else:
            oracle.bake_default(n, threads=1)
        ts.append(time.perf_counter() - t0)
    t = float(np.median(ts))
    return {"value": pts / t / 1e6, "unit": "Mpoints/s", "cores": 1, "kind": "reference" if use_ref else "port",
            "sample": "%d full 128^3 bakes, median; %s, one thread like the reference's start-up code" %
                      (repeats, "FastNoise2 (reference sources, best SIMD level of this CPU)" if use_ref else "scalar oracle port"),
            "ms": t * 1e3}


def run_bake(args, rank, world, local):
    """The start-up bake as its own bench line.  N > 1: replicas (every rank needs the whole volume; 'weak')."""
    W, H, desc, _ = WORKLOADS["bake"]
    n, pts = BAKE_SIZE, BAKE_FIELDS * BAKE_SIZE ** 3
    K, Wm = args.steps, max(args.warmup, 3)
    if args.impl == "reference":
        if rank != 0:
            return
        cb = time_cpu_bake(max(3, min(K + Wm, 40)))
        print(json.dumps({
            "impl": "reference", "metric": "Mpoints/s", "value": cb["value"], "unit": "Mpoints/s", "n_gpus": args.gpus, "steps": K,
            "warmup": Wm, "ms_per_step": cb["ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": "bake: " + desc, "size": n},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "Mpoints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}), flush=True)
        return
    import volumetricrenderer_b200 as vr
    from volumetricrenderer_b200 import capi
    from volumetricrenderer_b200.renderer import PinnedArray
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    r = vr.Renderer(64, 64, capi.OUT_RGBA8_UNORM, device=local)
    host = PinnedArray((n, n, n, 4), np.uint8)

    def barrier():
        r.wait()
        if dist is not None:
            dist.barrier(device_ids=[local])

    for _ in range(Wm):
        r.generate_volume(size=n)
    launches = r.stats()["kernel_launches"]
    sampler = ClockSampler(physical_gpu_index(local))
    barrier()
    sampler.start()
    dev_ms, t0 = 0.0, time.perf_counter()
    for _ in range(K):
        r.flush_l2()
        r.generate_volume(size=n)
        dev_ms += r.stats()["gpu_ms"]
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    # e2e: the call a host makes (synchronous bake) + the texels read back to pinned host memory
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        r.generate_volume(size=n)
        r.download_volume(out=host.array)
    barrier()
    e2e_s = (time.perf_counter() - t0) / K
    ms = dev_ms / K
    if dist is not None:
        import torch
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = float(t[0]), float(t[1])
    if rank == 0:
        peaks = r.measure_peaks()
        ach = BAKE_FLOP_PER_VOXEL * n ** 3 / (ms * 1e-3) / 1e12
        line = {
            "metric": "Mpoints/s", "value": world * pts / (ms * 1e-3) / 1e6, "unit": "Mpoints/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "bake: " + desc, "size": n, "l2": "flushed between bakes (256 MiB written, then read back)",
                       "partition": "replicas (every GPU bakes its own copy of the volume)" if world > 1 else "single GPU"},
            "e2e": {"value": world * pts / e2e_s / 1e6, "unit": "Mpoints/s", "ms_per_bake": e2e_s * 1e3,
                    "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4 * n ** 3},
            "gpu_launches": K * launches, "clocks": clocks, "wall_ms_per_step_incl_flush": wall / K * 1e3,
            "roofline": {"bound": "fp32", "achieved": ach, "peak": peaks["fp32_tflops"], "unit": "TFLOP/s",
                         "frac": ach / peaks["fp32_tflops"] if peaks["fp32_tflops"] else None, "traffic": None,
                         "note": "whole bake (fill x3, min/max-only x1, pack, texture copy, occupancy) against the FP32 FMA peak measured in this "
                                 "process; algorithmic %.0f flop/voxel; 128^3 is %d CTAs of work per field, launch latency is a visible share" %
                                 (BAKE_FLOP_PER_VOXEL, n ** 3 // 256)},
        }
        if not args.no_cpu_baseline and world == 1:
            cb = time_cpu_bake(20)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    barrier()
    host.close()
    r.close()
    if dist is not None:
        dist.destroy_process_group()


def roofline(r, args, workload, kw, steps, light_samples, ms_per_step, W, H, vol_bytes, world):
    """Dominant kernel = the march (the only kernel in the timed region).  Neither HBM nor the tensor cores bound
    this path: the procedural workloads read no input and the grid workloads' volumes are L2- or
    TEX-bound, so `bound` is "fp32" (algorithmic flop of SURVEY.md §8(d) / the FP32-FMA peak measured in this
    process) or "tex" (trilinear fetches / the fetch peak measured in this process).  The compulsory-HBM
    figure against MEASURED_PEAKS.json is reported beside it."""
    peaks = r.measure_peaks()
    t = ms_per_step * 1e-3
    measured = {}
    try:
        measured = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = measured.get("hbm_gbs", 6650.0)
    compulsory = vol_bytes + W * H * 4
    # dram bytes per launch from the committed ncu capture — only if that capture was taken on the kernels' current source
    # (profiles/traffic.json records the source hash; tools/update_traffic.py writes it); otherwise null, never a stale number
    traffic, traffic_note = None, "no ncu capture of this workload on the current kernel source"
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tj.get("source_sha16") == kernel_source_sha16():
            traffic = tj.get(workload)
            traffic_note = "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu capture %s" % tj.get("captures", {}).get(workload, "?")
    except Exception:
        pass
    out = {"hbm": {"algorithmic_bytes": compulsory, "achieved": compulsory / t / 1e9, "peak": hbm_peak * world, "unit": "GB/s",
                   "frac": compulsory / t / 1e9 / (hbm_peak * world),
                   "peak_source": "of measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in measured else "of fallback (6650 GB/s)"},
           "traffic": traffic, "traffic_note": traffic_note, "peaks_measured_here": peaks}
    if kw.get("density_mode") == 2:
        octaves, L = 5, kw.get("light_steps", 0)
        per_sample = octaves * FLOP_PER_OCTAVE + FLOP_PER_SAMPLE
        flop_per_step = (1 + L) * per_sample + FLOP_PER_STEP            # 2 266 (C2), 2 906 (C3)
        flops = steps * flop_per_step
        executed = (steps + light_samples) * per_sample + steps * FLOP_PER_STEP
        peak = peaks["fp32_tflops"] * world
        out.update({"bound": "fp32", "achieved": flops / t / 1e12, "peak": peak, "unit": "TFLOP/s",
                    "peak_source": "FP32 FMA micro-benchmark in this process (dependent-free FMA chains; nominal 74.5)",
                    "flop_per_step": flop_per_step, "flop_per_frame": flops,
                    "achieved_executed": executed / t / 1e12, "frac_executed": executed / t / 1e12 / peak if peak else None,
                    "note": "achieved = primary steps x SURVEY 8(d) flop/step (every step owes 1+L noise samples of 5 octaves x 60 + 20 flop, "
                            "+26 for the integrator). The kernel evaluates fewer (achieved_executed): steps whose density is exactly 0 "
                            "owe no light-march and light samples outside the box are not taken. Table fetches, byte extraction and "
                            "conversions share the issue ports and are not counted as flop: ncu shows the kernel 68 % issue-active and the "
                            "shared-memory pipe 75 % busy on 1.10 G warp instructions per C2 frame; the lattice hashes come from per-cell "
                            "tables built at vr_set_params (profiles/r02_march_fbm_cells_lds128.txt)."})
    else:
        # frag.glsl fetches four channels per step; a channel of the resident volume that holds one value everywhere
        # (the reference's default bake: G = 255, quirk Q1) is not fetched by the hardware-filtered march
        const_ch = bin(int(r.stats().get("constant_channels", 0)) & 15).count("1") if args.layout == "texture" else 0
        fetches = steps * ((4 - const_ch) if kw.get("density_mode", 0) == 0 else 1)
        peak = peaks["tex_gfetch"] * world
        out.update({"bound": "tex", "achieved": fetches / t / 1e9, "peak": peak, "unit": "Gfetch/s",
                    "peak_source": "trilinear-fetch micro-benchmark in this process (L1/L2-resident RGBA8, warp-coherent; nominal 581)",
                    "fetches_per_frame": fetches, "fetches_per_step_algorithmic": 4 if kw.get("density_mode", 0) == 0 else 1,
                    "constant_channels_elided": const_ch if kw.get("density_mode", 0) == 0 else 0,
                    "note": "the volume is cache-resident or streamed once; compulsory HBM traffic is under 'hbm'; achieved counts "
                            "the fetches actually issued (a constant channel of the volume is not fetched)"})
    if workload.startswith("c4") and args.layout == "texture":
        # what the texture path can stream out of DRAM at all (every texel of the resident grid fetched exactly once): the
        # ceiling of the 128-step march, whose fetches are nearly all first touches
        try:
            g = r.measure_tex_stream()
            out["hbm"]["tex_stream_gbs"] = g
            out["hbm"]["frac_of_tex_stream"] = out["hbm"]["achieved"] / (g * world) if g else None
            out["hbm"]["tex_stream_note"] = ("vr_measure_tex_stream in this process: the grid read once through tex3D (trilinear fetches "
                                             "on a 2-texel lattice, L2 flushed)")
        except Exception as e:
            out["hbm"]["tex_stream_gbs"] = None
            out["hbm"]["tex_stream_note"] = "failed: %r" % (e,)
    out["frac"] = out["achieved"] / out["peak"] if out["peak"] else None
    return out


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "bake":
        run_bake(args, rank, world, local)
        return
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.workload == "c5":
        run_views(args, rank, world, local)
        return
    run_volrt(args, rank, world, local)


if __name__ == "__main__":
    main()
