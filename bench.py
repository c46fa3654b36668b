#!/usr/bin/env python
"""bench.py — the hot-path benchmark (driver contract).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

A step = one frame of BASELINE.json configs[1] ("glass/mirror scene with Schlick Fresnel, patterns
and shadows at 1920x1080, depth 5" = scene-parser/examples/reflect-refract.yml at 1080p) rendered
by `Camera::render` semantics. metric = Mrays/s (rays = primary + shadow + reflection + refraction,
counted on the device), ms_per_step = frame ms.

  value   scene resident in HBM, K frames timed with CUDA events on the launching stream (max over
          ranks), L2 flushed between frames.
  e2e     the reference-facing call with HOST buffers, the same at every N: flatten(World) → H2D scene upload →
          render of the rank's tiles → canvas in pinned host memory, every step on every rank (public API:
          DeviceScene / rtc_render; at N = 1 the kernel stores its pixels straight into the pinned canvas over PCIe;
          at N > 1 the canvas is one shared pinned segment and every rank's tiles cross its own PCIe link).
  N > 1   one process per GPU (torchrun); the frame's row tiles are sharded cyclically. `value`: each rank's
          tiles land in rank 0's DEVICE canvas over NVLink (CUDA IPC peer mapping: kernel stores, or one strided
          DMA per rank); falls back to a grouped NCCL send/recv of the tiles if peer mapping is unavailable.
          scaling = strong. The e2e host canvas of rank 0 is compared with the oracle on a bounded block of rows.
  --impl reference   the reference's CPU algorithm (oracle/ C++ restatement; rustc is not in the
          image, so the Rust binary itself cannot run) on all host threads, same frame.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "raytracer-challenge-rs_b200"

WORKLOADS = {
    # name: (scene function, kwargs, description)
    "reflect_refract_1080p": ("reflect_refract", dict(width=1920, height=1080), "configs[1]: reflect-refract.yml @1920x1080 depth 5 aa1"),
    "chapter7_800x400": ("chapter7", dict(width=800, height=400), "configs[0]: chapter-7 world @800x400"),
    "cover_1080p": ("cover", dict(width=1920, height=1080), "configs[2a]: cover.yml @1920x1080"),
    "primitives_csg_1080p": ("primitives_csg", dict(width=1920, height=1080), "configs[2b]: cubes/cylinders/cones/CSG @1920x1080"),
    "teapot_standin_1080p": ("teapot_standin", dict(width=1920, height=1080), "configs[3]: OBJ mesh via Group::divide @1920x1080"),
    "mesh_and_spheres_4k": ("mesh_and_spheres", dict(width=3840, height=2160), "configs[4]: 1M triangles + 200 spheres @3840x2160"),
    "smoke_96x54": ("reflect_refract", dict(width=96, height=54), "tiny reflect-refract frame"),
}


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of one GPU, sampled in-process through NVML every ~2 ms for as long as the GPU
    section runs (a 10-40 ms timed region gives a 100 ms nvidia-smi poll nothing to see); `mark()` brackets the
    timed region so the line can say how many samples fell inside it. nvidia-smi is the fallback."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.stop_flag, self.marks = gpu_index, [], False, []
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def mark(self):
        self.marks.append(time.perf_counter())

    def _nvml_sample(self):
        n = self.nvml
        sm = n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(self.h, n.NVML_CLOCK_SM)
        try:
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        reasons = []
        for name, bit in (("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40), ("sw_power_cap", 0x4)):
            if r & bit:
                reasons.append(name)
        return (time.perf_counter(), float(sm), float(mx), reasons)

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            if self.nvml is not None:
                try:
                    self.samples.append(self._nvml_sample())
                except Exception:
                    pass
                time.sleep(0.002)
                continue
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in out.split(",")]
                self.samples.append((time.perf_counter(), float(f[0]), float(f[1]),
                                     [n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")]))
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        lo, hi = (self.marks[0], self.marks[-1]) if len(self.marks) >= 2 else (0.0, float("inf"))
        timed = [x for x in self.samples if lo <= x[0] <= hi]
        use = timed if len(timed) >= 5 else self.samples   # a region shorter than 5 polls: the whole GPU section
        sm = sorted(x[1] for x in use)
        reasons = sorted({r for x in use for r in x[3]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max((x[2] for x in use), default=None),
                "reasons": reasons, "samples": len(use), "samples_in_timed_region": len(timed),
                "source": "nvml, in-process, 2 ms poll" if self.nvml is not None else "nvidia-smi, 50 ms poll"}


def load_oracle(pkg):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import conftest
    return conftest.Oracle(pkg)


L2_NOTE = "GPU arm: L2 flushed between timed frames (256 MB fill); CPU arm: not applicable"


def config_of(args, desc, cam, aa):
    """The workload, in the same keys and values for both arms (the driver compares the two dicts)."""
    return {"workload": args.workload, "desc": desc, "width": cam.hsize, "height": cam.vsize, "aa": aa, "max_depth": 5,
            "l2": L2_NOTE if not getattr(args, "no_l2_flush", False) else "NOT flushed (--no-l2-flush: a profiling run, not a bench value)",
            "tile_rows": args.tile_rows}


def oracle_sample(orc, cam, world, aa, threads, budget_ms, H):
    """A bounded, centred block of rows of the frame on the oracle: probe a few rows, then as many as fit the budget
    (at least 16 — through the middle of the frame, where the glass / the mesh silhouette is). Returns (canvas, info, r0, rows)."""
    nprobe = min(H, max(4, min(2 * max(1, threads), 32)))
    p0 = (H - nprobe) // 2
    _, probe = orc.render(cam, world, aa=aa, threads=threads, rows=(p0, p0 + nprobe))
    per_row = max(probe["ms"] / nprobe, 1e-6)
    rows = H if per_row * H <= budget_ms else max(min(16, H), min(H, int(budget_ms / per_row)))
    r0 = (H - rows) // 2
    canvas, info = orc.render(cam, world, aa=aa, threads=threads, rows=(r0, r0 + rows))
    return canvas, info, r0, rows


def run_reference(args, pkg, scenes):
    """--impl reference: the reference's CPU render (C++ restatement: rustc is not in the image) of the same frame on
    all host threads, render_multithreaded semantics; each step a bounded sample of rows when the frame is too long."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    orc = load_oracle(pkg)
    fn, kw, desc = WORKLOADS[args.workload]
    world, cam = getattr(scenes, fn)(orc.api, **kw)
    cores = os.cpu_count() or 1
    aa = getattr(cam, "aa_samples", 1)
    h = cam.vsize
    if args.ref_rows:
        rows = min(h, args.ref_rows)
    else:
        # bound a step to about 2 s: probe 2 x cores rows, extrapolate
        nprobe = min(h, 2 * cores)
        p0 = (h - nprobe) // 2
        _, probe = orc.render(cam, world, aa=aa, threads=cores, rows=(p0, p0 + nprobe))
        per_row = max(probe["ms"] / nprobe, 1e-6)
        rows = h if per_row * h <= 2000.0 else max(min(16, h), int(2000.0 / per_row))
    r0 = 0 if rows >= h else (h - rows) // 2
    times, rays = [], 0
    for it in range(args.warmup + args.steps):
        _, info = orc.render(cam, world, aa=aa, threads=cores, rows=(r0, r0 + rows))
        if it >= args.warmup:
            times.append(info["ms"])
            rays = info["rays"]
    ms = sum(times) / len(times)
    value = rays / (ms * 1e-3) / 1e6
    sample = "rows [%d,%d) of %d per step, render_multithreaded semantics (camera.rs:150), %d threads; oracle/ C++ restatement (rustc absent)" % (r0, r0 + rows, h, cores)
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms * (h / rows), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(args, desc, cam, aa),
            "sample": {"rows_rendered": rows, "rows_total": h, "ms_per_sample": ms},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="reflect_refract_1080p")
    ap.add_argument("--gather", default="ipc", choices=["ipc", "nccl"])
    ap.add_argument("--tile-rows", type=int, default=8)
    ap.add_argument("--ref-rows", type=int, default=0, help="reference arm: render only this many rows (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=0.0, help="seconds of all-cores oracle work for the cpu_baseline / parity leg (default 15 at N=1, 6 at N>1)")
    ap.add_argument("--no-l2-flush", action="store_true", help="profiling only (ncu traffic capture): no 256 MB fill ahead of the frames; such a run's numbers are not bench values")
    ap.add_argument("--no-push-tiles", action="store_true", help="N >= 3: let every rank's kernels store into rank 0's canvas themselves")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    pkg = importlib.import_module(PKG)
    scenes = importlib.import_module(PKG + ".scenes")
    if args.impl == "reference":
        return run_reference(args, pkg, scenes)

    import numpy as np
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or pkg.cuda_lib().rtc_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the render path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world_size > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"    # the version banner goes to stdout, ahead of the one JSON line
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
    mg = importlib.import_module(PKG + ".multigpu")
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()

    fn, kw, desc = WORKLOADS[args.workload]
    api = pkg.scene_api()
    t0 = time.perf_counter()
    world, cam = getattr(scenes, fn)(api, **kw)
    build_ms = (time.perf_counter() - t0) * 1e3
    aa = getattr(cam, "aa_samples", 1)
    t0 = time.perf_counter()
    flat = pkg.FlatScene(world)
    flatten_ms = (time.perf_counter() - t0) * 1e3
    H, W = cam.vsize, cam.hsize

    gather = args.gather
    try:
        sr = mg.ShardedRenderer(pkg, flat, cam, rank, world_size, local_rank, tile_rows=args.tile_rows, gather=gather, dist=dist, push_tiles=not args.no_push_tiles)
    except Exception as e:  # peer mapping unavailable → grouped NCCL send/recv of the tiles
        if world_size > 1 and gather == "ipc":
            gather = "nccl"
            sr = mg.ShardedRenderer(pkg, flat, cam, rank, world_size, local_rank, tile_rows=args.tile_rows, gather=gather, dist=dist, push_tiles=not args.no_push_tiles)
        else:
            raise
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")  # 256 MB > 126 MB L2

    def step():
        sr.render_async(sptr, aa=aa)
        sr.gather_nccl()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    st = sr.finish(sptr)
    barrier()
    rays_rank = st.rays

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    if sampler:
        sampler.mark()
    wall0 = time.perf_counter()
    for i in range(args.steps):
        if not args.no_l2_flush:
            flush.fill_(float(i))        # L2 flush, outside the per-step events
        if dist is not None:
            dist.barrier()               # all ranks start the frame together
        ev[i][0].record(stream)
        step()
        ev[i][1].record(stream)
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3
    if sampler:
        sampler.mark()
    st = sr.finish(sptr)
    # a wavefront frame whose queues overflowed would have been timed incomplete (only the last one is re-rendered)
    overflowed = torch.tensor([1.0 if st.wave_overflowed else 0.0], dtype=torch.float64, device="cuda")
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([dev_ms, float(rays_rank)], dtype=torch.float64, device="cuda")
    if dist is not None:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dist.all_reduce(overflowed, op=dist.ReduceOp.MAX)
        dev_ms, rays_frame = float(tmax[0]), int(tsum[1])
    else:
        rays_frame = int(rays_rank)
    if float(overflowed[0]) != 0.0:
        raise SystemExit("bench.py: a timed wavefront frame overflowed its queues (RTC_STAT_WAVE_OVERFLOWED): the timing is not of whole frames")
    ms_per_step = dev_ms / args.steps
    value = rays_frame / (ms_per_step * 1e-3) / 1e6

    # ---- e2e, ONE definition at every N: the reference-facing call with HOST buffers, every step, on every rank —
    # flatten(World) → rtc_scene_create (H2D of the tables) → rtc_render of the rank's tiles into the HOST canvas →
    # rtc_scene_destroy. N = 1: a pinned host canvas (the kernel stores straight into it). N > 1: one shared pinned
    # canvas mapped by every rank (rtc_host_canvas); every rank's tiles cross its own PCIe link.
    e2e_steps = max(3, min(args.steps, 10))
    cam_rtc = pkg.camera_to_rtc(cam)
    opts = pkg.make_opts(aa=aa, tile_rows=args.tile_rows, rank=rank, nranks=world_size)
    stats = pkg.RtcStats()
    lib = pkg.cuda_lib()
    shared = None
    if world_size == 1:
        host_t = torch.zeros(H * W * 3, dtype=torch.float64).pin_memory()
        host_ptr, host_np = host_t.data_ptr(), host_t.numpy().reshape(H, W, 3)
    else:
        name = [None]
        if rank == 0:
            name[0] = "/rtc_canvas_%d_%d" % (os.getpid(), int(time.time()) & 0xffff)
            shared = mg.SharedHostCanvas(pkg, name[0], H * W * 3, local_rank, create=True)
        dist.broadcast_object_list(name, src=0)
        if rank != 0:
            shared = mg.SharedHostCanvas(pkg, name[0], H * W * 3, local_rank, create=False)
        shared.place(W, H, args.tile_rows, rank, world_size)   # each rank's tiles on its GPU's NUMA node, then pin
        dist.barrier()
        host_ptr, host_np = shared.ptr, shared.numpy().reshape(H, W, 3)
        dist.barrier()
    e2e_fallbacks = [0]
    phase = {"flatten": 0.0, "scene_create": 0.0, "render_call": 0.0, "render_kernels": 0.0, "scene_destroy": 0.0}

    def e2e_step():
        t_a = time.perf_counter()
        f = pkg.FlatScene(world)
        t_b = time.perf_counter()
        ds = pkg.DeviceScene(f, device=local_rank)            # H2D of the flattened scene
        t_c = time.perf_counter()
        pkg._check(lib.rtc_render(ds.h, C.byref(cam_rtc), C.byref(opts), host_ptr, C.byref(stats)))  # render + canvas to the host
        t_d = time.perf_counter()
        e2e_fallbacks[0] += 1 if stats.flags & pkg.RTC_STAT_WAVE_FALLBACK else 0
        ds.close()
        t_e = time.perf_counter()
        for k, v in (("flatten", t_b - t_a), ("scene_create", t_c - t_b), ("render_call", t_d - t_c), ("scene_destroy", t_e - t_d)):
            phase[k] += v * 1e3
        phase["render_kernels"] += stats.kernel_ms
        return f.info["bytes"]
    for _ in range(2):
        h2d = e2e_step()
    barrier()
    e2e_fallbacks[0] = 0
    for k in phase:
        phase[k] = 0.0
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        h2d = e2e_step()
        if shared is not None:                                # the frame is complete when every rank's tiles are in:
            shared.arrive(rank, k + 1)                        # hand-over through the shared segment (microseconds,
            shared.wait(world_size, k + 1)                    # no collective); every rank waits for the whole frame
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    if dist is not None:
        tt = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt[0])
    owned = int(np.count_nonzero(mg.owned_row_mask(H, args.tile_rows, rank, world_size)))
    e2e = {"value": rays_frame / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": (int(h2d) + 19 * 8) * world_size, "d2h_bytes_per_step": H * W * 24,
           "d2h_bytes_per_step_rank0": owned * W * 24, "wavefront_fallbacks": e2e_fallbacks[0],
           "rank0_ms_per_step": {k: v / e2e_steps for k, v in phase.items()},
           "definition": "every step, every rank: flatten(World) + rtc_scene_create (H2D) + rtc_render of the rank's tiles into the "
                         "host canvas (N=1: pinned; N>1: one shared pinned canvas, each rank over its own PCIe link) + rtc_scene_destroy; "
                         "frame hand-over per step through the segment (rtc_host_canvas_arrive / _wait on every rank); max over ranks"}

    # ---- parity of the e2e frame on rank 0 (any N), roofline, cpu baseline ------------------------------------------------
    roofline, cpu_baseline = None, None
    if rank == 0:
        peak_nofma = pkg.measure_fp64(local_rank, fused=False)
        peak_fma = pkg.measure_fp64(local_rank, fused=True)
        flops_frame = None
        if not args.no_cpu_baseline:
            orc = load_oracle(pkg)
            wo, co = getattr(scenes, fn)(orc.api, **kw)
            cores = os.cpu_count() or 1
            budget_ms = 1e3 * (args.cpu_budget_s if args.cpu_budget_s > 0 else (15.0 if world_size == 1 else 6.0))
            ref_canvas, info, r0, rows = oracle_sample(orc, co, wo, aa, cores, budget_ms, H)
            cpu_mrays = info["rays"] / (info["ms"] * 1e-3) / 1e6
            flops_per_ray = info["flops"] / max(1, info["rays"])
            flops_frame = flops_per_ray * rays_frame
            cpu_baseline = {"value": cpu_mrays, "unit": "Mrays/s", "cores": cores, "kind": "port",
                            "frame_ms_extrapolated": info["ms"] * H / rows,
                            "sample": "rows [%d,%d) of %d, render_multithreaded semantics (contiguous row blocks), %d threads; oracle/ C++ restatement of the reference (rustc absent)" % (r0, r0 + rows, H, cores)}
            # parity of the e2e host canvas (all ranks' tiles) against the oracle rows (not timed)
            got = host_np[r0:r0 + rows]
            want = ref_canvas[r0:r0 + rows]
            q = lambda a: np.clip(np.round(a * 255.0), 0, 255).astype(np.uint8)
            cpu_baseline["parity_rows"] = [r0, r0 + rows]
            cpu_baseline["parity_identical_8bit"] = float(np.mean(np.all(q(got) == q(want), axis=-1)))
            cpu_baseline["parity_max_8bit_diff"] = int(np.max(np.abs(q(got).astype(np.int32) - q(want).astype(np.int32))))
            cpu_baseline["parity_max_linear_diff"] = float(np.max(np.abs(got - want)))
            cpu_baseline["rays_match"] = bool(rows < H or info["rays"] == rays_frame)
            if world_size == 1:
                # Camera::render (camera.rs:133), one thread: a few rows spread over the same block (about 4 s of work)
                srows = min(rows, max(4, int(rows * 4000.0 / max(info["ms"] * cores, 1.0))))
                step_r = max(1, rows // srows)
                picks = list(range(r0 + step_r // 2, r0 + rows, step_r))[:srows]
                s_ms, s_rays = 0.0, 0
                for y in picks:
                    _, sinfo = orc.render(co, wo, aa=aa, threads=0 if aa == 1 else 1, rows=(y, y + 1))
                    s_ms += sinfo["ms"]
                    s_rays += sinfo["rays"]
                cpu_baseline["serial"] = {"value": s_rays / (s_ms * 1e-3) / 1e6, "unit": "Mrays/s", "cores": 1,
                                          "frame_ms_extrapolated": s_ms * H / len(picks),
                                          "sample": "%d rows spread over [%d,%d) of %d, Camera::render, 1 thread" % (len(picks), r0, r0 + rows, H)}
        # dram bytes per launch of the render kernel from the committed ncu --set full capture, if any
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
                traffic = json.load(fh).get(args.workload, {}).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
        flops_source = "counted by the oracle in this run"
        if flops_frame is None:  # cpu_baseline leg skipped: the oracle's committed per-ray figure for this workload
            try:
                with open(os.path.join(ROOT, "profiles", "algorithmic_flops.json")) as fh:
                    fpr = json.load(fh).get(args.workload)
                if fpr:
                    flops_frame = float(fpr) * rays_frame
                    flops_source = "profiles/algorithmic_flops.json (oracle count of an earlier N=1 run) x rays of this run"
            except Exception:
                pass
        # BVH workloads: the traversal's bytes against the measured HBM rate (the reference's flop count says nothing there:
        # it tests every triangle for every ray)
        abytes = None
        try:
            with open(os.path.join(ROOT, "profiles", "algorithmic_bytes.json")) as fh:
                abytes = json.load(fh).get(args.workload)
        except Exception:
            abytes = None
        if abytes:
            hbm = None
            try:
                with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
                    hbm = float(json.load(fh)["hbm_gbs"])
                peak_src = "MEASURED_PEAKS.json hbm_gbs (driver-measured copy rate) x n_gpus"
            except Exception:
                hbm, peak_src = 6550.0, "fallback 6550 GB/s (B200_PROFILING.md): MEASURED_PEAKS.json absent"
            bytes_frame = abytes["bytes_per_ray"] * rays_frame + 24.0 * W * H
            achieved = bytes_frame / (ms_per_step * 1e-3) / 1e9
            roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm * world_size, "unit": "GB/s", "frac": achieved / (hbm * world_size),
                        "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_frame": bytes_frame,
                        "bytes_definition": "BVH nodes visited x 64 B + leaf prims tested x 96 B (device-counted per ray, profiles/algorithmic_bytes.json) x rays of this run + 24 B per pixel",
                        "fp64_peak_tflops_nofma": peak_nofma / 1e3 * world_size}
        elif flops_frame is not None:
            achieved = flops_frame / (ms_per_step * 1e-3) / 1e12
            peak_nofma *= world_size   # whole-job achieved against the N-GPU roof (rank 0's measured peak x N)
            peak_fma *= world_size
            roofline = {"bound": "fp64", "achieved": achieved, "peak": peak_nofma / 1e3, "unit": "TFLOP/s",
                        "frac": achieved / (peak_nofma / 1e3), "traffic": traffic, "flops_source": flops_source,
                        "peak_source": "measured in this run: rtc_measure_fp64 (DMUL+DADD stream, non-fused — the library is built --fmad=false); MEASURED_PEAKS.json has no FP64 figure",
                        "peak_dfma_tflops": peak_fma / 1e3, "frac_of_dfma_peak": achieved / (peak_fma / 1e3),
                        "algorithmic_flops_per_frame": flops_frame,
                        "flop_definition": "one per + - * / sqrt as written in the reference (SURVEY 8-a constants), counted by the oracle"}
        else:
            roofline = {"bound": "fp64", "achieved": None, "peak": peak_nofma / 1e3, "unit": "TFLOP/s", "frac": None, "traffic": traffic,
                        "peak_dfma_tflops": peak_fma / 1e3}

    if sampler:
        sampler.stop_flag = True
    if rank == 0:
        if st.flags & pkg.RTC_STAT_WAVEFRONT:
            renderer = ("wavefront, one barrier-free persistent launch per frame (rt_stream.cuh)" if st.launches == 1
                        else "wavefront, %d level-synchronous launches per frame (rt_wave.cuh)" % st.launches)
        else:
            renderer = "recursive kernel, 1 launch/frame"
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": config_of(args, desc, cam, aa),
                "run": {"rays_per_frame": rays_frame, "sharding": "cyclic %d-row tiles, gather=%s" % (args.tile_rows, sr.gather),
                        "renderer": renderer, "timed_frames_overflowed": False,
                        "scene_bytes": flat.info["bytes"], "scene_build_ms": build_ms, "flatten_ms": flatten_ms,
                        "wall_ms_timed_region": wall_ms},
                "e2e": e2e, "gpu_launches": args.steps * int(st.launches), "clocks": sampler.summary() if sampler else None,
                "roofline": roofline, "cpu_baseline": cpu_baseline}
        print(json.dumps(line), flush=True)
    if shared is not None:
        dist.barrier()
        shared.close()
    sr.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
