"""N>1 path on CPU: world_size-2 `gloo` processes shard one frame exactly as the multi-GPU run does
(cyclic row tiles, rtc_opts.rank/nranks/tile_rows), each renders ITS tiles with the host build of the
kernel source, and the tiles are gathered by the very function the NCCL gather variant uses (pack → one group
of send / recv → unpack, multigpu.gather_tiles). The result must be bit-identical to the unsharded frame and to the
oracle; a ragged frame height (36 rows, 8-row tiles, 2 ranks) is part of the case."""
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
    rows = mg.padded_rows(H, tile_rows, world_size)
    t = torch.zeros(rows, W * 3, dtype=torch.float64)
    t[:H] = torch.from_numpy(part.reshape(H, W * 3))
    rays = torch.tensor([info["rays"]], dtype=torch.int64)
    mg.gather_tiles(dist, t, rank, world_size, tile_rows)
    dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    if rank == 0:
        q.put((t[:H].numpy().reshape(H, W, 3).copy(), int(rays[0])))
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


def test_pack_unpack_round_trip(pkg):
    import torch
    mg = importlib.import_module("raytracer-challenge-rs_b200.multigpu")
    for H, tr, n in [(1080, 8, 8), (36, 8, 2), (37, 4, 3)]:
        rows = mg.padded_rows(H, tr, n)
        assert rows >= H and rows % (tr * n) == 0 and rows - H < tr * n
        src = torch.arange(rows * 6, dtype=torch.float64).view(rows, 6)
        dst = torch.zeros_like(src)
        for r in range(n):
            p = mg.pack_tiles(src, r, tr, n)
            own = [y for y in range(rows) if (y // tr) % n == r]
            assert torch.equal(p.reshape(-1, 6), src[own])
            mg.unpack_tiles(dst, p, r, tr, n)
        assert torch.equal(dst, src)


def test_tile_ownership_partitions_the_frame(pkg):
    mg = importlib.import_module("raytracer-challenge-rs_b200.multigpu")
    for H, tr, n in [(1080, 8, 8), (400, 8, 3), (37, 4, 2), (5, 8, 4)]:
        seen = np.zeros(H, dtype=np.int32)
        for r in range(n):
            rows = mg.owned_rows(H, tr, r, n)
            assert rows == list(np.nonzero(mg.owned_row_mask(H, tr, r, n))[0])
            seen[rows] += 1
        assert np.all(seen == 1)
