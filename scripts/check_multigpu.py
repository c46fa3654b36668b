#!/usr/bin/env python
"""check_multigpu.py — run under torchrun on N GPUs: the sharded frame (cyclic row tiles, every rank's
kernel storing straight into rank 0's canvas over NVLink, or the NCCL reduce variant) must be
bit-identical to the single-GPU frame, and the per-rank ray counts must add up to the frame's.
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/check_multigpu.py [workload]"""
import importlib
import os
import sys

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
kw = dict(width=640, height=360)
world_, cam = scenes.CONFIGS[name](pkg.scene_api(), **kw)
flat = pkg.FlatScene(world_)
ok = True
for gather in ("ipc", "nccl"):
    sr = mg.ShardedRenderer(pkg, flat, cam, rank, world, local, tile_rows=8, gather=gather, dist=dist)
    s = torch.cuda.current_stream().cuda_stream
    sr.render_async(s)
    sr.gather_nccl()
    st = sr.finish(s)
    dist.barrier()
    torch.cuda.synchronize()
    rays = torch.tensor([st.rays], dtype=torch.int64, device="cuda")
    dist.all_reduce(rays)
    if rank == 0:
        host = torch.empty(cam.hsize * cam.vsize * 3, dtype=torch.float64).pin_memory()
        sr.download(host)
        got = host.numpy().reshape(cam.vsize, cam.hsize, 3)
        ds = pkg.DeviceScene(flat, device=local)
        want, st1 = ds.render(cam)
        ds.close()
        same = np.array_equal(got, want)
        print("gather=%s n=%d identical=%s rays %d vs %d" % (gather, world, same, int(rays[0]), st1.rays), flush=True)
        ok = ok and same and int(rays[0]) == st1.rays
    dist.barrier()
    sr.close()
if rank == 0:
    print("MULTIGPU_CHECK", "OK" if ok else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
