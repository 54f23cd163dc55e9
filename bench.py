#!/usr/bin/env python
"""bench.py -- Mrays/s and frames/s of the raytracer hot path at 3840x2160 (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload render_map_4k|spheres1024_4k]

One "step" = one frame of the workload.  At N=1 the workload is BASELINE.json configs[1]:
scenes/render.map at 3840x2160, all materials/textures, MAX_DEPTH 15, 2 soft-shadow samples.
"ray" = one segment iteration of raytracing.cl:42 plus each shadow ray (raytracing.cl:108); the count
per frame comes from the counted variant of the kernel (checked against the oracle's counters in
tests/test_parity_gpu.py), never from the timed launch.

  value     device-timed whole-job Mrays/s, scene/textures resident in HBM, frame left in HBM
  e2e       the same metric through the reference-facing cl_wrap_* calls (the sequence raypng.c:50-89
            makes per frame: 6 camera scalars, run raygen, run raytracer + blocking read-back of the
            frame into a pinned HOST buffer); at N>1: set camera, render band, NCCL gather, D2H on rank 0
  roofline  FP32 pipe: algorithmic flop (21/sphere test, +5 if D>=0, 14/plane test) over the
            kernel's CUDA-event time, against SMs*128*2*sm_max_mhz and against a measured FMA loop
  cpu_baseline / --impl reference
            the reference's own .cl text compiled for the host (oracle/_ref, kind "reference"; the
            plain-C port if that library is absent) on all host threads, on a bounded sample of the
            same frame (same scene and camera at a reduced resolution; rays counted at that resolution)
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

SM_LANES = 128  # FP32 lanes per SM
_JSON_FD = 1


def kernel_source_hash():
    """sha256 over the kernel sources: a committed ncu capture only describes the kernel it was taken from."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "example-gui-opencl-raytracer_b200", "csrc")
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".h")) and name != "rt_engine.cu":   # the kernels and what they include
            h.update(name.encode())
            h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()[:16]


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one render_frame launch from the committed `ncu --set full` capture
    of this same command (profiles/traffic.json, written by tools/ncu_traffic.py on the GPU box together with the hash of
    the kernel sources it profiled).  A capture of OTHER sources is refused: (None, why)."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(workload)
    except Exception:
        return None, "no profiles/traffic.json"
    if not rec:
        return None, "no ncu capture of this workload in profiles/traffic.json"
    if rec.get("source_hash") != kernel_source_hash():
        return None, "profiles/traffic.json was captured from other kernel sources (%s, now %s): refused as stale" % (
            rec.get("source_hash"), kernel_source_hash())
    return rec["dram_bytes"], ("DRAM bytes per launch (ncu --set full, %s); algorithmic bytes: 4 B per pixel of frame + <= 4 B per "
                               "texel fetch; the kernel is FP32-issue bound, not HBM bound" % rec.get("capture", "profiles/"))


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


# --------------------------------------------------------------------------------------------------
# workloads
def make_workload(name, host):
    if name == "render_map_4k":
        s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
        W, H = 3840, 2160
        camp = dict(host.RAYPNG_CAMERA)
        desc = "scenes/render.map (4 spheres, 2 planes, 3 lights) at 3840x2160, raypng camera, 4 textures + skybox"
    elif name == "render_map_800":
        s, p, l = host.load_map(os.path.join(ROOT, "scenes", "render.map"))
        W, H = 800, 600
        camp = dict(host.RAYPNG_CAMERA)
        desc = "scenes/render.map at the reference's default 800x600"
    elif name == "spheres1024_4k":
        s, p, l, camp = host.sphere_grid_scene(32)
        W, H = 3840, 2160
        desc = "synthetic 1,024 mixed-material spheres over a textured plane at 3840x2160 (counts widened)"
    elif name in ("mesh50k_8k", "mesh50k_4k", "mesh5k_4k"):
        n = 50 if name == "mesh5k_4k" else 158
        s, p, l, tris, camp = host.heightfield_mesh_scene(n)
        W, H = (7680, 4320) if name.endswith("8k") else (3840, 2160)
        desc = ("EXTENSION X1 (no reference parity): %d-triangle heightfield + ground plane + 2 spheres at %dx%d, "
                "brute-force closest hit" % (len(tris) // 112, W, H))
    elif name == "spheres255_8k":
        s, p, l, camp = host.sphere_grid_scene(16)
        s = s[:255 * 96]
        W, H = 7680, 4320
        desc = "255 mixed-material spheres (the most the reference's uchar counts allow) over a textured plane at 7680x4320"
    else:
        raise SystemExit("unknown workload " + name)
    if not name.startswith("mesh"):
        tris = b""
    return dict(name=name, spheres=s, planes=p, lights=l, tris=tris, W=W, H=H, cam=host.camera(W=W, H=H, **camp), camp=camp,
                desc=desc, camera_fn=host.camera)


# --------------------------------------------------------------------------------------------------
# clocks (sampled DURING the timed region)
class ClockSampler:
    BAD = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")

    def __init__(self, index):
        self.index, self.samples, self.reasons = index, [], set()
        self.stop_flag = threading.Event()
        self.h = None
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.h = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "hw_power_brake": 0x80, "sync_boost": 0x10}
        while not self.stop_flag.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((mhz, util))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.004)

    def __enter__(self):
        if self.h is not None:
            self.t = threading.Thread(target=self._loop, daemon=True)
            self.t.start()
        return self

    def __exit__(self, *a):
        self.stop_flag.set()
        if self.h is not None:
            self.t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        mhz = sorted(m for m, _ in self.samples)
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(mhz)}


# --------------------------------------------------------------------------------------------------
# CPU reference arm / baseline
def cpu_sample_shape(W, H, budget_s, probe):
    """Pick the divisor s so that the (W/s)x(H/s) frame fits the time budget, from a small probe run."""
    probe_px, probe_s = probe
    per_px = probe_s / probe_px
    for s in (1, 2, 3, 4, 5, 6, 8, 10, 12, 15, 16, 20, 24, 30, 40, 48, 60, 80, 120):
        if W % s or H % s:
            continue
        if per_px * (W // s) * (H // s) <= budget_s:
            return s
    return 120


def cpu_setup(wl):
    ref = entry.load_oracle()
    tex, sky = ref.default_textures(ROOT)
    scene = ref.Scene(wl["spheres"], wl["planes"], wl["lights"], tex, sky)
    wide = max(scene.S, scene.P, scene.L) > 255 or len(wl["tris"]) > 0
    if len(wl["tris"]):   # extension X1: only the port knows triangles
        lib_, kind = ref.port(), "port"
        lib_.set_triangles(wl["tris"])
    elif ref.have_ref():
        lib_, kind = ref.ref(), "reference"
    else:
        lib_, kind = ref.port(), "port"
    cores = os.cpu_count() or 1
    return ref, scene, wide, lib_, kind, cores


def cpu_camera_for(ref, wl, W, H):
    """Same view at another resolution (rgen_perspective recomputed for W x H)."""
    return wl["camera_fn"](W=W, H=H, **wl["camp"])


def cpu_time_frames(wl, steps, warmup, budget_total_s):
    """Returns (mrays_per_s, ms_per_step, cores, kind, sample description)."""
    ref, scene, wide, lib_, kind, cores = cpu_setup(wl)
    W, H = wl["W"], wl["H"]
    pw, ph = (W // 16, H // 16) if not len(wl["tris"]) else (W // 96, H // 90)   # brute-force meshes: tiny probe
    t0 = time.perf_counter()
    lib_.render(scene, cpu_camera_for(ref, wl, pw, ph), pw, ph, wide=wide, threads=cores, want_hit=False)
    probe = (pw * ph, max(time.perf_counter() - t0, 1e-4))
    per_step = budget_total_s / max(1, steps + warmup)
    s = cpu_sample_shape(W, H, per_step, probe)
    sw, sh = W // s, H // s
    cam = cpu_camera_for(ref, wl, sw, sh)
    _, _, cnt = ref.port().render_counted(scene, cam, sw, sh, wide=wide, threads=cores)  # accounting only
    for _ in range(warmup):
        lib_.render(scene, cam, sw, sh, wide=wide, threads=cores, want_hit=False)
    t0 = time.perf_counter()
    for _ in range(steps):
        lib_.render(scene, cam, sw, sh, wide=wide, threads=cores, want_hit=False)
    dt = (time.perf_counter() - t0) / steps
    sample = "same scene and camera at %dx%d (1/%d of the %dx%d pixels), %d rays/frame, %d steps" % (
        sw, sh, s * s, W, H, cnt["rays"], steps)
    return cnt["rays"] / dt / 1e6, dt * 1e3, cores, kind, sample, (sw * sh) / dt


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    pkg = entry.load_package()
    from example_gui_opencl_raytracer_b200 import host
    wl = make_workload(args.workload, host)
    mrays, ms, cores, kind, sample, pps = cpu_time_frames(wl, args.steps, args.warmup, budget_total_s=150.0)
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "frames_per_s_equiv": pps / (wl["W"] * wl["H"]),
        "config": {"workload": wl["name"], "desc": wl["desc"], "width": wl["W"], "height": wl["H"]},
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------------------------------
# GPU arm
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N ...")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    pkg = entry.load_package()
    eng = pkg.Engine(local)
    try:
        gpu_body(args, pkg, eng, world, rank, local)
    finally:
        # every tensor that touched the engine's stream or NCCL's is gone (gpu_body's locals) before either is torn down
        import gc
        gc.collect()
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        eng.close()
    return 0


class Ctx:
    """What every measurement of one bench.py process shares: the engine, its stream as a torch stream, the L2 flush
    buffer, the decoded image arrays."""

    def __init__(self, args, pkg, eng, world, rank, local):
        import torch
        import torch.distributed as dist
        from example_gui_opencl_raytracer_b200 import host
        self.args, self.pkg, self.eng, self.world, self.rank, self.local = args, pkg, eng, world, rank, local
        self.torch, self.dist, self.host = torch, dist, host
        self.dev = torch.device("cuda", local)
        self.stream = torch.cuda.ExternalStream(eng.stream(), device=self.dev)
        with torch.cuda.stream(self.stream):
            self.flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        self.tex, self.sky = host.default_images(ROOT)
        eng.upload_images(0, self.tex)
        eng.upload_images(1, self.sky)
        pk = peaks()
        self.sm_max = float(pk.get("sm_max_mhz", 1965.0))
        self.sms = torch.cuda.get_device_properties(local).multi_processor_count
        self.peak_tf = self.sms * SM_LANES * 2 * self.sm_max * 1e6 / 1e12
        self.dbg = os.environ.get("BENCH_DEBUG")

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()


class _Raw:
    """A raw device pointer as a CUDA array (for torch.as_tensor)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 3}


FLAG_STRIDE_WORDS = 16      # one 64-byte line per rank's completion flag


def measure(ctx, wl, steps, warmup, exchange="p2p", want_e2e=True, verify="full", kernel_reps=10, opts=None, count=True):
    """One workload on this process group: accounting launch, bit-identity check against one GPU (N>1), device-timed
    steps, kernel-only time for the roofline and (want_e2e) the end-to-end leg.  Returns a dict (complete on rank 0)."""
    torch, dist, eng, pkg = ctx.torch, ctx.dist, ctx.eng, ctx.pkg
    world, rank, dev, stream, args = ctx.world, ctx.rank, ctx.dev, ctx.stream, ctx.args
    W, H = wl["W"], wl["H"]
    eng.upload_scene(wl["spheres"], wl["planes"], wl["lights"])
    eng.upload_triangles(wl["tris"])
    eng.set_camera(wl["cam"])
    for kv in filter(None, os.environ.get("BENCH_ENGINE_OPTS", "").split(",")):   # A/B hook: name=value,...
        k, v = kv.split("=")
        eng.set_option(k, int(v))
    for k, v in (opts or {}).items():
        eng.set_option(k, v)

    band = args.band_rows
    part = pkg.Partition(band, world, rank) if world > 1 else None
    max_rows = max(int(pkg.lib().rtb_partition_rows(H, pkg.Partition(band, world, r))) for r in range(world)) if world > 1 else H
    exchange = exchange if world > 1 else "none"

    with torch.cuda.stream(stream):
        band_buf = torch.zeros(max_rows * W, dtype=torch.int32, device=dev)
        # rank 0 with --exchange nccl: one contiguous [world][max_rows*W] landing area for the gather
        landing = torch.empty(world * max_rows * W, dtype=torch.int32, device=dev) if (exchange == "nccl" and rank == 0) else None
        gathered = list(landing.view(world, max_rows * W).unbind(0)) if landing is not None else None
        frame = torch.empty(H * W, dtype=torch.int32, device=dev) if (rank == 0 and exchange != "p2p") else None
    stream.synchronize()

    # ---- accounting launch (untimed): rays and algorithmic flop of this rank's rows ----
    if count:
        cnt = eng.render_counted(W, H, part=part, d_out=band_buf.data_ptr())
    else:   # lines that report a frame time only (the literal counted launch of a 50k-triangle frame takes many seconds)
        cnt = dict(rays=0, flop=0, segments=0, shadow_rays=0, sphere_tests=0, plane_tests=0)
    totals = torch.tensor([cnt["rays"], cnt["flop"], cnt["segments"], cnt["shadow_rays"], cnt["sphere_tests"],
                           cnt["plane_tests"]], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(totals)
    frame_rays, frame_flop = int(totals[0]), int(totals[1])
    my_flop = cnt["flop"]

    # ---- band exchange ----
    # "p2p":  every rank's render kernel also stores its bands straight into rank 0's frame over NVLink (CUDA IPC mapping
    #         of a cudaMalloc'ed frame, rtb_set_mirror_frame) and its LAST CTA release-stores the frame number into that
    #         rank's flag word behind the frame (rtb_set_completion); rank 0 enqueues a one-warp wait kernel on those
    #         words (rtb_wait_flags).  No collective per frame at all.
    # "nccl": compact band buffers gathered to rank 0 with one NCCL gather, then one assemble_bands kernel.
    raw_frame = peer_frame = None
    flag_off = (H * W * 4 + 255) // 256 * 256            # bytes; the flags live behind the frame, same allocation
    target = None
    if exchange == "p2p":
        okf = torch.ones(1, dtype=torch.int32, device=dev)
        handle = torch.zeros(64, dtype=torch.uint8, device=dev)
        if rank == 0:
            try:
                if os.environ.get("BENCH_FORCE_NO_IPC"):
                    raise RuntimeError("forced by BENCH_FORCE_NO_IPC")
                raw_frame = eng.dev_alloc(flag_off + 4096)
                handle = torch.tensor(list(eng.ipc_export(raw_frame)), dtype=torch.uint8, device=dev)
            except Exception as ex:
                print("CUDA IPC export failed (%s): falling back to --exchange nccl" % ex, file=sys.stderr)
                okf.zero_()
        dist.broadcast(handle, 0)
        dist.broadcast(okf, 0)
        if int(okf.item()) == 1 and rank != 0:
            try:
                peer_frame = eng.ipc_open(bytes(handle.cpu().tolist()))
                target = peer_frame.value
            except Exception as ex:
                print("CUDA IPC open failed on rank %d (%s): falling back to --exchange nccl" % (rank, ex), file=sys.stderr)
                okf.zero_()
        dist.all_reduce(okf, op=dist.ReduceOp.MIN)
        if int(okf.item()) == 1:
            if rank == 0:
                target = raw_frame.value
                frame = torch.as_tensor(_Raw(target, H * W), device=dev)
                with torch.cuda.stream(stream):
                    torch.as_tensor(_Raw(target + flag_off, 1024), device=dev).zero_()
                stream.synchronize()
            dist.barrier()
        else:
            if peer_frame is not None:
                eng.ipc_close(peer_frame)
                peer_frame = None
            dist.barrier()
            if raw_frame is not None:
                eng.dev_free(raw_frame)
                raw_frame = None
            exchange = "nccl"
            with torch.cuda.stream(stream):
                landing = torch.empty(world * max_rows * W, dtype=torch.int32, device=dev) if rank == 0 else None
                gathered = list(landing.view(world, max_rows * W).unbind(0)) if landing is not None else None
                frame = torch.empty(H * W, dtype=torch.int32, device=dev) if rank == 0 else None
            stream.synchronize()

    dbg = [] if ctx.dbg else None
    frame_no = [0]

    def step_device():
        """One frame, everything on the engine's stream; the frame stays in HBM (rank 0)."""
        if exchange == "p2p":
            frame_no[0] += 1
            if dbg is not None:
                e0 = torch.cuda.Event(enable_timing=True); e0.record(stream)
            eng.set_mirror_frame(target)
            eng.set_completion(target + flag_off + 4 * FLAG_STRIDE_WORDS * rank, frame_no[0])
            eng.render(W, H, part=part, d_out=band_buf.data_ptr())
            if dbg is not None:
                e1 = torch.cuda.Event(enable_timing=True); e1.record(stream)
            if rank == 0:                         # completion only: the pixels already travelled with the stores
                eng.wait_flags(target + flag_off, FLAG_STRIDE_WORDS, world, frame_no[0])
            if dbg is not None:
                e2 = torch.cuda.Event(enable_timing=True); e2.record(stream)
                dbg.append((e0, e1, e2, eng.last_kernel_ms() if len(dbg) % 8 == 7 else None))
            return
        eng.render(W, H, part=part, d_out=(band_buf.data_ptr() if world > 1 else frame.data_ptr()))
        if world > 1:
            dist.gather(band_buf, gathered, dst=0)   # the one collective: bands -> rank 0 over NCCL/NVLink
            if rank == 0:
                eng.assemble_bands(landing.data_ptr(), max_rows * W, frame.data_ptr(), W, H, band, world)

    verified = None
    if world > 1:   # untimed: the N-GPU frame must be bit-identical to what ONE GPU renders
        with torch.cuda.stream(stream):
            step_device()
            ctx.barrier()
            if rank == 0:
                if verify == "full":
                    full = torch.empty(H * W, dtype=torch.int32, device=dev)
                    eng.render(W, H, d_out=full.data_ptr())
                    eng.sync()
                    verified = bool(torch.equal(full, frame))
                    del full
                else:   # the full frame on one GPU would take tens of seconds: every 16th band (8 rows), pixel ids global
                    sub = pkg.Partition(band, 16, 0)
                    nrows = eng.partition_rows(H, sub)
                    part_buf = torch.empty(nrows * W, dtype=torch.int32, device=dev)
                    eng.render(W, H, part=sub, d_out=part_buf.data_ptr())
                    eng.sync()
                    nb = (H + band - 1) // band
                    rows = torch.cat([torch.arange(b * band, min(H, (b + 1) * band)) for b in range(0, nb, 16)]).to(dev)
                    verified = {"equal": bool(torch.equal(part_buf.view(nrows, W), frame.view(H, W).index_select(0, rows))),
                                "checked": "every 16th band of %d rows (%d of %d rows) re-rendered by rank 0 alone" % (band, nrows, H)}
                    del part_buf
            ctx.barrier()

    def timed(fn, nsteps, nwarm, do_flush=True):
        """Per-step CUDA events on the launching stream, L2 flushed between steps (outside the events).
        Returns this rank's total ms over `nsteps`."""
        with torch.cuda.stream(stream):
            for _ in range(nwarm):
                fn()
            ctx.barrier()
            evs = []
            for _ in range(nsteps):
                if do_flush:
                    ctx.flush.fill_(1)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                fn()
                b.record(stream)
                evs.append((a, b))
            ctx.barrier()
        per = [a.elapsed_time(b) for a, b in evs]
        return sum(per), per

    # ---- device-resident value + kernel-only time for the roofline ----
    kernel_ms = []
    with ClockSampler(ctx.local) as clk:
        t_start = time.perf_counter()
        total_ms, per = timed(step_device, steps, warmup)
        wall = time.perf_counter() - t_start
        # kernel-only: the engine's own event pair brackets exactly the render_frame launch
        with torch.cuda.stream(stream):
            for _ in range(min(steps, kernel_reps)):
                ctx.flush.fill_(1)
                eng.render(W, H, part=part, d_out=band_buf.data_ptr())
                kernel_ms.append(eng.last_kernel_ms())
    clocks = clk.summary()
    if dbg:
        torch.cuda.synchronize()
        rs = [a.elapsed_time(b) for a, b, _, _ in dbg[-12:]]
        ar = [b.elapsed_time(c) for _, b, c, _ in dbg[-12:]]
        kin = [k for _, _, _, k in dbg if k is not None]
        print("%s rank %d: render span (memsets+hint+order+kernel) median %.3f max %.3f; completion-wait span median %.3f max %.3f; "
              "render_frame alone inside the loop %s; kernel-only reps %s; loop wall per step %.3f ms" %
              (wl["name"], rank, float(np.median(rs)), max(rs), float(np.median(ar)), max(ar), ["%.3f" % k for k in kin],
               ["%.3f" % k for k in kernel_ms], wall * 1e3 / (steps + warmup)), file=sys.stderr)
    timeouts = eng.wait_timeouts()
    if timeouts:
        raise SystemExit("rtb_wait_flags timed out %d times: a rank's completion flag never arrived" % timeouts)

    # ---- end to end through the public API with host buffers ----
    e2e_ms = h2d = d2h = 0
    rb_stats = e2e_api = None
    if want_e2e and world == 1:
        host_frame = torch.empty(H * W, dtype=torch.int32, pin_memory=True)
        e2e_ms, h2d, d2h, rb_stats = e2e_cl_wrap(pkg, wl, args, host_frame, steps)
        e2e_api = "cl_wrap_load_single_data x6 + cl_wrap_output(raygen) + cl_wrap_output(raytracer, read-back)"
        del host_frame
    elif want_e2e:
        e2e_ms, h2d, d2h, rb_stats, e2e_api = e2e_multi(ctx, wl, part, band_buf, frame, step_device, timed, steps, warmup, exchange)

    t_ms = torch.tensor([total_ms, e2e_ms, float(np.mean(kernel_ms))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, kern_max_ms = [float(x) for x in t_ms]
    ms_per_step = total_ms / steps
    my_kernel_ms = float(np.mean(kernel_ms))
    achieved_tf = (my_flop if world > 1 else frame_flop) / (my_kernel_ms * 1e-3) / 1e12
    # per frame and rank: tile_cost_history + tile_order + render_frame (+ render_tree when the frame has heavy tiles for it);
    # on rank 0 + wait_flags (p2p) or assemble_bands (nccl)
    launches_per_frame = (3 + (1 if eng.last_launch().get("tree_grid") else 0)) * world + (1 if exchange in ("nccl", "p2p") else 0)
    res = {
        "workload": wl["name"], "desc": wl["desc"], "width": W, "height": H, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_per_step, "value": frame_rays / (ms_per_step * 1e-3) / 1e6, "unit": "Mrays/s",
        "frames_per_s": 1e3 / ms_per_step, "mpixels_per_s": W * H / (ms_per_step * 1e-3) / 1e6,
        "rays_per_frame": frame_rays, "algorithmic_flop_per_frame": frame_flop,
        "segments": int(totals[2]), "shadow_rays": int(totals[3]),
        "exchange": exchange, "frame_equals_1gpu": verified, "launch": eng.last_launch(), "clocks": clocks,
        "gpu_launches": steps * launches_per_frame,
        "kernel_ms": my_kernel_ms, "kernel_ms_slowest_rank": kern_max_ms,
        "roofline": {"bound": "fp32", "achieved": achieved_tf, "peak": ctx.peak_tf, "unit": "TFLOP/s",
                     "frac": achieved_tf / ctx.peak_tf, "kernel": "render_frame", "kernel_ms": my_kernel_ms},
    }
    if want_e2e:
        res["e2e"] = {"value": frame_rays / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms,
                      "frames_per_s": 1e3 / e2e_ms, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "readback": rb_stats, "api": e2e_api}

    if exchange == "p2p":   # unmap the peers' views before rank 0 frees the frame
        torch.cuda.synchronize()
        dist.barrier()
        if peer_frame is not None:
            eng.ipc_close(peer_frame)
        dist.barrier()
        frame = None
        if raw_frame is not None:
            eng.dev_free(raw_frame)
    for k in (opts or {}):
        eng.set_option(k, 0)
    return res


def e2e_multi(ctx, wl, part, band_buf, frame, step_device, timed, steps, warmup, exchange):
    """e2e at N > 1: the frame must end up in HOST memory on rank 0.  Every rank maps the same POSIX shared-memory frame
    (each process page-locks its own mapping) and its render kernel stores its bands straight into it over its own PCIe
    link (rtb_set_mirror_frame with the host view) and, last CTA, its frame number into a flag word behind the frame
    (rtb_set_completion): no gather, no D2H copy, no collective.  Rank 0's HOST polls the flag words -- the frame is
    complete in the caller's memory when they all show the frame number.  Timed with the host clock (the call a user
    makes is synchronous), barrier + synchronize on both sides, max over ranks."""
    torch, dist, eng = ctx.torch, ctx.dist, ctx.eng
    world, rank, dev = ctx.world, ctx.rank, ctx.dev
    W, H = wl["W"], wl["H"]
    flag_off = (H * W * 4 + 255) // 256 * 256
    shm_words = (flag_off + 4096) // 4
    shm_path = "/dev/shm/rtb200_frame_%s_%s" % (os.environ.get("MASTER_PORT", "0"), wl["name"])
    ok = torch.ones(1, dtype=torch.int32, device=dev)
    shm = None
    if rank == 0:
        try:
            if os.environ.get("BENCH_FORCE_NO_SHM"):
                raise RuntimeError("forced by BENCH_FORCE_NO_SHM")
            shm = np.memmap(shm_path, dtype=np.uint32, mode="w+", shape=(shm_words,))
            shm[:H * W] = 0xFFFFFFFF
            shm[H * W:] = 0
        except Exception as ex:   # /dev/shm missing or too small for this frame
            print("shared host frame unavailable (%s): e2e falls back to a D2H copy on rank 0" % ex, file=sys.stderr)
            ok.zero_()
    dist.broadcast(ok, 0)
    use_shm = int(ok.item()) == 1
    if use_shm and rank != 0:
        try:
            shm = np.memmap(shm_path, dtype=np.uint32, mode="r+", shape=(shm_words,))
        except Exception:
            ok.zero_()
    if use_shm:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        use_shm = int(ok.item()) == 1
    if use_shm:
        host_view, registered = eng.host_map(shm.ctypes.data, shm_words * 4)
        flags = shm[flag_off // 4:]
        n = [0]

        def step_e2e():
            n[0] += 1
            eng.set_camera(wl["cam"])
            eng.set_mirror_frame(host_view)
            eng.set_completion(host_view + flag_off + 4 * FLAG_STRIDE_WORDS * rank, n[0])
            eng.render(W, H, part=part, d_out=band_buf.data_ptr())
            if rank == 0:                                    # the HOST waits for every rank's flag
                deadline = time.perf_counter() + 20.0
                for r in range(world):
                    while int(flags[FLAG_STRIDE_WORDS * r]) - n[0] < 0:
                        if time.perf_counter() > deadline:
                            raise SystemExit("e2e: rank %d's completion flag never reached frame %d" % (r, n[0]))
            else:
                eng.sync()                                   # a synchronous API: one frame in flight per rank
        for _ in range(max(3, warmup)):
            step_e2e()
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            step_e2e()
        eng.sync()
        t = (time.perf_counter() - t0) * 1e3
        ctx.barrier()
        e2e_ok = None
        if rank == 0:
            e2e_ok = bool(np.array_equal(shm[:H * W], frame.cpu().numpy().view(np.uint32)))
        dist.barrier()
        if registered:
            eng.host_unregister(shm.ctypes.data)
        rb_stats = {"shared_host_frame_equals_device_frame": e2e_ok}
        e2e_api = ("rtb_set_camera + rtb_render(band) on every rank with the frame mirrored into a shared HOST frame (POSIX shm, "
                   "page-locked by each rank) + per-rank completion words in that same host memory, polled by rank 0's host")
    else:
        host_frame = torch.empty(H * W, dtype=torch.int32, pin_memory=True) if rank == 0 else None

        def step_e2e():
            eng.set_camera(wl["cam"])
            step_device()
            if rank == 0:
                host_frame.copy_(frame, non_blocking=True)
        t, _ = timed(step_e2e, steps, warmup, do_flush=False)
        rb_stats = None
        e2e_api = "rtb_set_camera + rtb_render(band, %s exchange) + D2H of the frame on rank 0" % exchange
    flags = None
    shm = None
    dist.barrier()
    if rank == 0 and os.path.exists(shm_path):
        os.unlink(shm_path)
    return t / steps, 14 * 4, H * W * 4, rb_stats, e2e_api


def orbit_config(ctx):
    """BASELINE.json configs[4]: 240-frame camera sweep at 2560x1440 through the rayinteractive call sequence
    (drivers/orbit_bench.c replays rayinteractive.c:86-103,186-189 against cl_wrap_* with the reference's own
    cpu_ray.c / cpu_obj.c linked unchanged); every 24th frame is also counted for the roofline."""
    import subprocess
    exe = os.path.join(ROOT, "drivers", "_bin", "orbit_bench")
    if not os.path.exists(exe):
        return {"unavailable": "drivers/_bin/orbit_bench not built (needs the reference's host sources at build time)"}
    try:
        r = subprocess.run([exe, "2560", "1440", "240", "-", "24"], cwd=ROOT, capture_output=True, text=True, timeout=300,
                           env=dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", str(ctx.local))))
        line = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as ex:
        return {"unavailable": "orbit_bench failed: %s" % ex}
    out = {"workload": "orbit_1440p", "desc": "240-frame look-around + strafe sweep of scenes/render.map at 2560x1440 through "
           "cl_wrap_* (rayinteractive.c call sequence), 14.7 MB read-back per frame",
           "frames": line["frames"], "frame_ms_p50": line["frame_ms_p50"], "frame_ms_p99": line["frame_ms_p99"],
           "frame_ms_max": line["frame_ms_max"], "frames_per_s": line["frames_per_s"], "kernel_ms_mean": line["kernel_ms_mean"],
           "d2h_bytes_per_frame": line["d2h_bytes_per_frame"], "gpu_launches": 3 * line["frames"]}
    if line.get("sampled_frames"):
        k = line["sampled_frames"]
        tf = line["sampled_flop"] / (line["sampled_kernel_ms"] * 1e-3) / 1e12
        out["value"] = line["sampled_rays"] / k / (line["frame_ms_p50"] * 1e-3) / 1e6
        out["unit"] = "Mrays/s (mean rays of %d sampled frames / p50 frame time)" % k
        out["roofline"] = {"bound": "fp32", "achieved": tf, "peak": ctx.peak_tf, "unit": "TFLOP/s", "frac": tf / ctx.peak_tf,
                           "kernel": "render_frame", "kernel_ms": line["sampled_kernel_ms"] / k,
                           "sample": "every 24th frame counted by the literal launch (untimed)"}
    return out


def mesh_accel_config(ctx, host):
    """BASELINE.json configs[3]'s mesh at 3840x2160 on one GPU through the N4 hierarchy over the triangles
    (csrc/rt_trace_accel.cuh, option "accel"), next to ONE brute-force frame of the same engine, which is also the judge:
    every colour word and every hit id of the two frames must be equal."""
    eng = ctx.eng
    wl = make_workload("mesh50k_4k", host)
    W, H = wl["W"], wl["H"]
    eng.upload_scene(wl["spheres"], wl["planes"], wl["lights"])
    eng.upload_triangles(wl["tris"])
    eng.set_camera(wl["cam"])
    eng.set_option("accel", 0)
    b_rgb, b_hit = eng.render_to_host(W, H, want_hit=True)
    brute_ms = eng.last_render_ms()
    eng.set_option("accel", 1)
    try:
        t0 = time.perf_counter()
        a_rgb, a_hit = eng.render_to_host(W, H, want_hit=True)       # includes the one-off host build of the hierarchy
        first_ms = (time.perf_counter() - t0) * 1e3
        same = bool(np.array_equal(a_rgb, b_rgb) and np.array_equal(a_hit, b_hit))
        ms = []
        for i in range(3 + 10):
            with ctx.torch.cuda.stream(ctx.stream):
                ctx.flush.fill_(1)
            eng.render(W, H)
            eng.sync()
            if i >= 3:
                ms.append(eng.last_render_ms())
        a_rgb, a_hit = eng.render_to_host(W, H, want_hit=True)       # a history-scheduled frame
        same = same and bool(np.array_equal(a_rgb, b_rgb) and np.array_equal(a_hit, b_hit))
        launch = eng.last_launch()
    finally:
        eng.set_option("accel", 0)
        eng.upload_triangles(b"")
    m = float(np.mean(ms))
    return {"workload": "mesh50k_4k_accel", "desc": wl["desc"].replace("brute-force closest hit", "box hierarchy over the triangles (N4)"),
            "width": W, "height": H, "steps": len(ms), "warmup": 3, "ms_per_step": m, "ms_per_step_max": float(np.max(ms)),
            "frames_per_s": 1e3 / m, "brute_force_ms_per_step": brute_ms, "speedup_vs_brute_force": brute_ms / m,
            "first_frame_ms_incl_host_build": first_ms, "frame_equals_brute_force": same, "launch": launch,
            "gpu_launches": 3 * len(ms),
            "note": "EXTENSION N4 over EXTENSION X1 (neither is in the reference; parity unpinned): one brute-force frame (every tile in the "
                    "tree kernel, TMA-staged chunks) timed once and compared word for word, colour and hit ids, with the hierarchy's "
                    "first and last frame; times are the engine's CUDA events around a whole frame, L2 flushed between frames"}


def slim(res):
    """An other_configs entry: the measurement without the bulky keys."""
    keep = ("workload", "desc", "width", "height", "steps", "warmup", "ms_per_step", "value", "unit", "frames_per_s",
            "rays_per_frame", "algorithmic_flop_per_frame", "exchange", "frame_equals_1gpu", "launch", "gpu_launches",
            "kernel_ms", "kernel_ms_slowest_rank", "roofline")
    return {k: res[k] for k in keep if k in res}


def gpu_body(args, pkg, eng, world, rank, local):
    import torch
    ctx = Ctx(args, pkg, eng, world, rank, local)
    host = ctx.host
    wl = make_workload(args.workload, host)
    main = measure(ctx, wl, args.steps, args.warmup, exchange=args.exchange, want_e2e=True)

    others = {}
    want_others = args.other_configs == "on" or (args.other_configs == "auto" and args.workload == "render_map_4k")
    if want_others:
        # BASELINE.json configs[2] at every N; configs 4' / 5 on one GPU; configs[3] (extension X1) only where a frame
        # takes seconds instead of tens of seconds; the NCCL-gather exchange beside the peer-store default
        others["spheres1024_4k"] = slim(measure(ctx, make_workload("spheres1024_4k", host), 3, 3, exchange=args.exchange,
                                                want_e2e=False, kernel_reps=3))
        # SURVEY 8(f) N4 (extension, not the reference's algorithm): the same frame through the cluster walk; counted rays
        # and algorithmic flop are the brute-force ones, so "roofline" here is an equivalent rate, not executed work
        others["spheres1024_4k_accel"] = slim(measure(ctx, make_workload("spheres1024_4k", host), 5, 3, exchange=args.exchange,
                                                      want_e2e=False, kernel_reps=3, opts={"accel": 1}))
        rf = others["spheres1024_4k_accel"].pop("roofline", None)
        if rf:   # not a roofline claim: the numerator is the brute-force test count, which this walk does not execute
            others["spheres1024_4k_accel"]["brute_force_equivalent_tflops"] = rf["achieved"]
        others["spheres1024_4k_accel"]["note"] = ("EXTENSION N4: cluster acceleration structure (csrc/rt_trace_accel.cuh), frame word-identical "
                                                  "to the brute-force one (tests/test_accel_gpu.py); Mrays/s and the roofline figure count the "
                                                  "reference's brute-force rays and tests, which this walk mostly skips")
        if world == 1:
            others["spheres255_8k"] = slim(measure(ctx, make_workload("spheres255_8k", host), 3, 3, want_e2e=False, kernel_reps=3))
            others["orbit_1440p"] = orbit_config(ctx) if rank == 0 else None
            others["mesh50k_4k_accel"] = mesh_accel_config(ctx, host)
        if world > 1:
            alt = "nccl" if args.exchange == "p2p" else "p2p"
            others["render_map_4k_exchange_" + alt] = slim(measure(ctx, wl, min(args.steps, 10), 3, exchange=alt, want_e2e=False,
                                                                   kernel_reps=3))
        if world == 8:
            others["mesh50k_8k"] = slim(measure(ctx, make_workload("mesh50k_8k", host), 1, 1, exchange=args.exchange,
                                                want_e2e=False, verify="sampled", kernel_reps=1))
            others["mesh50k_8k"]["note"] = ("EXTENSION X1, parity unpinned (the reference has no triangle type); one warm-up and one "
                                            "timed step: a frame takes seconds")
            acc = slim(measure(ctx, make_workload("mesh50k_8k", host), 5, 3, exchange=args.exchange, want_e2e=False,
                               verify="sampled", kernel_reps=3, opts={"accel": 1}, count=False))
            for k in ("value", "unit", "rays_per_frame", "algorithmic_flop_per_frame", "roofline"):
                acc.pop(k, None)
            acc["workload"] = "mesh50k_8k_accel"
            acc["note"] = ("EXTENSION N4 over EXTENSION X1: the same 8K frame through the box hierarchy over the triangles; frame time only "
                           "(no counted launch); equality with brute force: tests/test_accel_gpu.py and the N=1 line mesh50k_4k_accel")
            others["mesh50k_8k_accel"] = acc

    if rank == 0:
        W, H = wl["W"], wl["H"]
        exchange = main["exchange"]
        fma_tf = eng.fp32_peak_tflops(4096)
        traffic, traffic_note = ncu_traffic(wl["name"])
        line = {
            "metric": "Mrays/s", "value": main["value"], "unit": "Mrays/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": main["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "frames_per_s": main["frames_per_s"], "mpixels_per_s": main["mpixels_per_s"],
            "config": {"workload": wl["name"], "desc": wl["desc"], "width": W, "height": H,
                       "rays_per_frame": main["rays_per_frame"], "algorithmic_flop_per_frame": main["algorithmic_flop_per_frame"],
                       "segments": main["segments"], "shadow_rays": main["shadow_rays"],
                       "parallelism": ("row bands of %d rows dealt cyclically over %d GPUs, " % (args.band_rows, world)) +
                                      ("bands stored into rank 0's frame over NVLink by the render kernel (CUDA IPC); completion = one "
                                       "release-stored word per rank behind the frame, awaited by a one-warp kernel on rank 0; no "
                                       "collective per frame" if exchange == "p2p" else "NCCL gather to rank 0 + assemble kernel")
                       if world > 1 else "1 GPU",
                       "frame_equals_1gpu": main["frame_equals_1gpu"],
                       "l2": "flushed between timed steps (256 MiB fill outside the event pair)",
                       "launch": main["launch"]},
            "clocks": main["clocks"],
            "e2e": main["e2e"],
            "gpu_launches": main["gpu_launches"],
            "roofline": dict(main["roofline"], traffic=traffic, traffic_note=traffic_note,
                             peak_source="%d SMs x 128 lanes x 2 flop x %.0f MHz (MEASURED_PEAKS.json sm_max_mhz); no tensor/HBM bound applies"
                                         % (ctx.sms, ctx.sm_max),
                             measured_fma_tflops=fma_tf, frac_of_measured_fma=main["roofline"]["achieved"] / fma_tf,
                             kernel_ms_slowest_rank=main["kernel_ms_slowest_rank"],
                             note="kernel is compiled without FMA contraction (parity contract), so mul+add pairs cap it at 0.5 of an FMA peak"),
        }
        if others:
            line["other_configs"] = others
        if world == 1 and not args.no_cpu:
            mrays, ms, cores, kind, sample, _ = cpu_time_frames(wl, 1, 0, budget_total_s=20.0)
            line["cpu_baseline"] = {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample}
        sys.stdout.flush()
        os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


def e2e_cl_wrap(pkg, wl, args, host_frame, steps):
    """The per-frame call sequence of raypng.c:50-57,86-89 / rayinteractive.c:98-103,186-189 with HOST
    buffers; set-up (scene, images, output slot) happens once, as in the reference drivers."""
    from example_gui_opencl_raytracer_b200.wrap_session import WrapSession
    W, H = wl["W"], wl["H"]
    ses = WrapSession(W, H, wl["spheres"], wl["planes"], wl["lights"], wl["cam"], root=ROOT, triangles=wl["tris"])
    hp = host_frame.data_ptr()

    def frame():
        ses.set_camera(wl["cam"])
        ses.frame(hp)
    for _ in range(max(3, args.warmup)):
        frame()
    t0 = time.perf_counter()  # the API is synchronous (clFinish + blocking read): host clock == device done
    for _ in range(steps):
        frame()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    stats = ses.readback_stats()
    ses.release()
    return ms, 4 * 16 + 2 * 4, 4 * W * H, stats


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="render_map_4k")
    ap.add_argument("--band-rows", type=int, default=8)
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="N>1: how bands reach rank 0")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--other-configs", default="auto", choices=["auto", "on", "off"],
                    help="also measure BASELINE.json's other configs in this run (auto: with the default workload)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference_arm(args)
    # libraries (NCCL's version banner) write to fd 1: keep stdout for the ONE JSON line
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
