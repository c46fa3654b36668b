"""N>1 path on CPU: world_size-2 `gloo` processes shard one frame exactly as the multi-GPU run does
(cyclic row tiles, rtc_opts.rank/nranks/tile_rows), each renders ITS tiles with the host build of the
kernel source, and the canvases are combined with the same disjoint-support reduce the NCCL gather
variant uses. The result must be bit-identical to the unsharded frame and to the oracle."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world_size, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import conftest
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    pkg = importlib.import_module("raytracer-challenge-rs_b200")
    scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
    mg = importlib.import_module("raytracer-challenge-rs_b200.multigpu")
    emul = conftest.Emul(pkg)
    w, c = scenes.reflect_refract(pkg.scene_api(), 64, 36)
    flat = pkg.FlatScene(w)
    H, W, tile_rows = 36, 64, 8
    # this rank's tiles only; untouched pixels stay zero
    part, info = emul.render(c, w, flat=flat, threads=2, tile_rows=tile_rows, rank=rank, nranks=world_size)
    mask = mg.owned_row_mask(H, tile_rows, rank, world_size)
    assert np.all(part[~mask] == 0.0), "a rank wrote rows it does not own"
    t = torch.from_numpy(part.copy())
    rays = torch.tensor([info["rays"]], dtype=torch.int64)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)        # disjoint supports → the sum is the gather
    dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    if rank == 0:
        q.put((t.numpy(), int(rays[0])))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_frame_equals_full_frame(pkg, prod, orc, emul):
    import torch.multiprocessing as mp
    scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 200)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered, rays = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    w, c = scenes.reflect_refract(prod, 64, 36)
    full, info = emul.render(c, w)
    assert np.array_equal(gathered, full)
    assert rays == info["rays"]
    wo, co = scenes.reflect_refract(orc.api, 64, 36)
    want, iw = orc.render(co, wo, threads=0)
    assert np.array_equal(gathered, want) and rays == iw["rays"]


def test_tile_ownership_partitions_the_frame(pkg):
    mg = importlib.import_module("raytracer-challenge-rs_b200.multigpu")
    for H, tr, n in [(1080, 8, 8), (400, 8, 3), (37, 4, 2), (5, 8, 4)]:
        seen = np.zeros(H, dtype=np.int32)
        for r in range(n):
            rows = mg.owned_rows(H, tr, r, n)
            assert rows == list(np.nonzero(mg.owned_row_mask(H, tr, r, n))[0])
            seen[rows] += 1
        assert np.all(seen == 1)
