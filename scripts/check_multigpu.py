#!/usr/bin/env python
"""check_multigpu.py — run under torchrun on N GPUs. The sharded frame must be bit-identical to the single-GPU frame and
the per-rank ray counts must add up to the frame's, for every way the tiles come together:
  ipc    every rank's tiles into rank 0's DEVICE canvas over NVLink (kernel stores, or one strided DMA per rank)
  nccl   tiles packed, one group of ncclSend / ncclRecv, unpacked on rank 0
  host   every rank's rtc_render writes its tiles into ONE shared pinned HOST canvas (rtc_host_canvas) over its own PCIe link
and for a scene whose wavefront queues OVERFLOW on every rank (nested glass): the frames that were re-rendered
recursively inside finish() must reach rank 0 all the same (the tiles are pushed / sent again).
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/check_multigpu.py [workload]"""
import ctypes as C
import importlib
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("raytracer-challenge-rs_b200")
scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
mg = importlib.import_module("raytracer-challenge-rs_b200.multigpu")

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
name = sys.argv[1] if len(sys.argv) > 1 else "reflect_refract"
api = pkg.scene_api()


def nested_glass():
    """every hit spawns a reflection AND a refraction ray: a level needs more than the 2 queue slots per pixel"""
    w = api.world()
    w.add_light((-4, 6, -8), (1, 1, 1))
    for k in range(4):
        r = 3.0 - 0.5 * k
        w.add_object(api.glass_sphere().set_transform(pkg.transform.scaling(r, r, r)).set_material(
            pkg.Material(color=(0.1, 0.1, 0.1), reflective=0.9, transparency=0.9, refractive_index=1.1 + 0.1 * k)))
    c = api.camera(640, 360, 0.9)
    c.set_transform(pkg.transform.view_transform((0, 0, -2.9), (0, 0, 0), (0, 1, 0)))
    return w, c


ok = True
cases = [(name,) + scenes.CONFIGS[name](api, width=640, height=360), ("nested_glass (queues overflow)",) + nested_glass()]
for label, world_, cam in cases:
    flat = pkg.FlatScene(world_)
    want = st1 = None
    if rank == 0:
        ds = pkg.DeviceScene(flat, device=local)
        want, st1 = ds.render(cam, wavefront=False)
        ds.close()
    n = cam.hsize * cam.vsize * 3
    for gather in ("ipc", "nccl"):
        for force_wave in ((None, True) if world == 2 else (None,)):   # 2 ranks default to the recursive kernel
            sr = mg.ShardedRenderer(pkg, flat, cam, rank, world, local, tile_rows=8, gather=gather, dist=dist)
            s = torch.cuda.current_stream().cuda_stream
            sr.render_async(s, **(dict(wavefront=True) if force_wave else {}))
            sr.gather_nccl()
            st = sr.finish(s)
            dist.barrier()
            torch.cuda.synchronize()
            info = torch.tensor([st.rays, 1 if st.wave_fallback else 0], dtype=torch.int64, device="cuda")
            dist.all_reduce(info)
            if rank == 0:
                host = torch.empty(n, dtype=torch.float64).pin_memory()
                sr.download(host)
                got = host.numpy().reshape(cam.vsize, cam.hsize, 3)
                same = np.array_equal(got, want)
                print("%-32s gather=%-4s n=%d wave=%s identical=%s rays %d vs %d, ranks that fell back: %d" %
                      (label, gather, world, bool(force_wave) or world >= 3, same, int(info[0]), st1.rays, int(info[1])), flush=True)
                ok = ok and same and int(info[0]) == st1.rays
            dist.barrier()
            sr.close()
    # the shared HOST canvas: every rank renders its tiles through rtc_render
    shm = [None]
    if rank == 0:
        shm[0] = "/rtc_check_%d_%d" % (os.getpid(), int(time.time()) & 0xffff)
    dist.broadcast_object_list(shm, src=0)
    if rank == 0:
        canvas = mg.SharedHostCanvas(pkg, shm[0], n, local, create=True)
    dist.barrier()
    if rank != 0:
        canvas = mg.SharedHostCanvas(pkg, shm[0], n, local, create=False)
    canvas.place(cam.hsize, cam.vsize, 8, rank, world)
    dist.barrier()
    ds = pkg.DeviceScene(flat, device=local)
    cam_rtc = pkg.camera_to_rtc(cam)
    opts = pkg.make_opts(tile_rows=8, rank=rank, nranks=world)
    stats = pkg.RtcStats()
    pkg._check(pkg.cuda_lib().rtc_render(ds.h, C.byref(cam_rtc), C.byref(opts), canvas.ptr, C.byref(stats)))
    ds.close()
    canvas.arrive(rank, 1)
    canvas.wait(world, 1)
    if rank == 0:
        got = canvas.numpy().reshape(cam.vsize, cam.hsize, 3)
        same = np.array_equal(got, want)
        print("%-32s shared host canvas n=%d identical=%s" % (label, world, same), flush=True)
        ok = ok and same
    dist.barrier()
    canvas.close()
if rank == 0:
    print("MULTIGPU_CHECK", "OK" if ok else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
