"""Row-tile sharding of one frame over the GPUs of a box (one process per GPU, torch.distributed).

The reference shards contiguous row blocks over OS threads (camera.rs:155-172) and sends each
block back over an mpsc channel into the host Canvas (:195-207). Here rows are grouped in tiles of
`tile_rows` rows and tile t belongs to rank t % nranks (cyclic, so expensive image regions spread over
all GPUs); the scene is replicated. No data-path collective:

  device gather (`ShardedRenderer`, gather="ipc")   rank 0 owns a device canvas, exports it with CUDA IPC, and every
      rank's kernel stores its finished pixels straight into it over NVLink (2 ranks), or renders into a local canvas
      and pushes its tiles with one strided DMA (`rtc_canvas_push_tiles`, >= 3 ranks, where the wavefront runs);
  host canvas (`SharedHostCanvas`)   what the caller of render_multithreaded gets: ONE pinned shared-memory canvas
      mapped by every rank; each rank's rtc_render writes ITS tiles over ITS OWN PCIe link;
  gather="nccl"   the plain-NCCL line item and the fallback when peer mapping is unavailable: every rank packs its
      tiles (one strided copy) and rank 0 receives them with one group of ncclRecv (ranks: ncclSend) and unpacks.
"""
import ctypes as C
import os

import numpy as np


def tile_owner(row, tile_rows, nranks):
    return (row // tile_rows) % nranks


def owned_rows(vsize, tile_rows, rank, nranks):
    """Frame rows rendered by `rank` (sorted)."""
    return [y for y in range(vsize) if tile_owner(y, tile_rows, nranks) == rank]


def owned_row_mask(vsize, tile_rows, rank, nranks):
    y = np.arange(vsize)
    return ((y // tile_rows) % nranks) == rank


def padded_rows(vsize, tile_rows, nranks):
    """vsize rounded up to whole rounds of `nranks` tiles: a canvas of that many rows views as
    [rounds, nranks, tile_rows * row_elems], and [:, r] is rank r's tiles — the packed form the NCCL gather sends."""
    step = tile_rows * nranks
    return (vsize + step - 1) // step * step


def pack_tiles(canvas_rows, rank, tile_rows, nranks):
    """canvas_rows: torch tensor [padded_rows, row_elems] → contiguous [rounds, tile_rows * row_elems] of rank's tiles."""
    rows, row_elems = canvas_rows.shape
    return canvas_rows.view(rows // (tile_rows * nranks), nranks, tile_rows * row_elems)[:, rank].contiguous()


def unpack_tiles(canvas_rows, packed, rank, tile_rows, nranks):
    rows, row_elems = canvas_rows.shape
    canvas_rows.view(rows // (tile_rows * nranks), nranks, tile_rows * row_elems)[:, rank].copy_(packed)


def gather_tiles(dist, canvas_rows, rank, nranks, tile_rows, recv=None):
    """Every rank's tiles into rank 0's canvas with one group of send / recv (NCCL: ncclSend / ncclRecv in one
    ncclGroup; gloo on CPU tensors in the tests). canvas_rows: [padded_rows, row_elems]; recv: rank 0's per-rank
    packed buffers (allocated here when None)."""
    if rank == 0:
        if recv is None:
            rounds = canvas_rows.shape[0] // (tile_rows * nranks)
            recv = [canvas_rows.new_empty(rounds, tile_rows * canvas_rows.shape[1]) for _ in range(nranks)]
        ops = [dist.P2POp(dist.irecv, recv[r], r) for r in range(1, nranks)]
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        for r in range(1, nranks):
            unpack_tiles(canvas_rows, recv[r], r, tile_rows, nranks)
    else:
        packed = pack_tiles(canvas_rows, rank, tile_rows, nranks)
        for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, packed, 0)]):
            w.wait()


class SharedHostCanvas:
    """rtc_host_canvas: one POSIX shm segment mapped by every rank. create/open → place() on every rank → the caller's
    barrier → .ptr (page-locks and maps the segment for this process's GPU)."""

    def __init__(self, pkg, name, n_doubles, device, create):
        self.pkg, self.n = pkg, n_doubles
        self.h = C.c_void_p()
        self._ptr = None
        pkg._check(pkg.cuda_lib().rtc_host_canvas_create(name.encode(), n_doubles, device, 1 if create else 0, C.byref(self.h)))

    def place(self, hsize, vsize, tile_rows, rank, nranks):
        """This rank's tiles onto the NUMA node of its GPU (best effort) and zero-filled; before the first .ptr."""
        self.pkg._check(self.pkg.cuda_lib().rtc_host_canvas_place(self.h, hsize, vsize, tile_rows, rank, nranks))

    @property
    def ptr(self):
        if self._ptr is None:
            self._ptr = self.pkg.cuda_lib().rtc_host_canvas_ptr(self.h)
            if not self._ptr:
                raise self.pkg.RtcError(-2, (self.pkg.cuda_lib().rtc_last_error() or b"").decode())
        return self._ptr

    def numpy(self):
        return np.ctypeslib.as_array((C.c_double * self.n).from_address(self.ptr))

    def arrive(self, rank, frame):
        """This rank's tiles of frame `frame` (1, 2, …) are in the canvas."""
        self.pkg._check(self.pkg.cuda_lib().rtc_host_canvas_arrive(self.h, rank, frame))

    def wait(self, nranks, frame, timeout_s=60.0):
        """Returns once every rank has arrived at `frame`: the hand-over of the frame, through the segment itself."""
        self.pkg._check(self.pkg.cuda_lib().rtc_host_canvas_wait(self.h, nranks, frame, timeout_s))

    def close(self):
        if self.h:
            self.pkg.cuda_lib().rtc_host_canvas_destroy(self.h)
            self.h = None


class ShardedRenderer:
    """One rank's side of the sharded render into a DEVICE canvas on rank 0. Needs torch with CUDA and an initialised
    process group."""

    def __init__(self, pkg, flat, camera, rank, nranks, device, tile_rows=8, gather="ipc", dist=None, push_tiles=True):
        import torch
        self.torch, self.pkg, self.dist = torch, pkg, dist
        self.rank, self.nranks, self.device, self.tile_rows = rank, nranks, device, tile_rows
        self.camera = pkg.camera_to_rtc(camera)
        self.scene = pkg.DeviceScene(flat, device=device)
        self.n = self.camera.hsize * self.camera.vsize * 3
        self.push_tiles = push_tiles
        self.gather = gather if nranks > 1 else "local"
        self.canvas = None       # rtc_canvas handle (ipc) or torch tensor
        self.canvas_ptr = None
        self.local = None
        self._setup_canvas()

    def _setup_canvas(self):
        torch, pkg, lib = self.torch, self.pkg, self.pkg.cuda_lib()
        if self.gather == "ipc":
            handle = torch.zeros(64, dtype=torch.uint8)
            h = C.c_void_p()
            if self.rank == 0:
                pkg._check(lib.rtc_canvas_create(self.device, self.n, C.byref(h)))
                buf = C.create_string_buffer(64)
                pkg._check(lib.rtc_canvas_export(h, buf))
                handle = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
            hd = handle.to("cuda:%d" % self.device) if self.dist.get_backend() == "nccl" else handle
            self.dist.broadcast(hd, src=0)
            if self.rank != 0:
                raw = bytes(hd.cpu().numpy().tobytes())
                pkg._check(lib.rtc_canvas_open(self.device, raw, self.n, C.byref(h)))
            self.canvas = h
            self.canvas_ptr = lib.rtc_canvas_ptr(h)
            # >= 3 ranks: scenes that recurse are rendered as a wavefront, whose pixels complete in 24-byte pieces —
            # ranks other than 0 then render into a local canvas and push their tiles to rank 0 with one strided DMA
            # (rtc_canvas_push_tiles) instead of storing across NVLink piecewise
            if self.rank != 0 and self.nranks >= 3 and self.push_tiles:
                self.local = torch.zeros(self.n, dtype=torch.float64, device="cuda:%d" % self.device)
        else:
            rows = padded_rows(self.camera.vsize, self.tile_rows, max(1, self.nranks))
            self.row_elems = self.camera.hsize * 3
            self.canvas = torch.zeros(rows * self.row_elems, dtype=torch.float64, device="cuda:%d" % self.device)
            self.canvas_ptr = self.canvas.data_ptr()
            if self.gather == "nccl" and self.rank == 0:
                rounds = rows // (self.tile_rows * self.nranks)
                self.recv = [torch.empty(rounds, self.tile_rows * self.row_elems, dtype=torch.float64, device="cuda:%d" % self.device)
                             for _ in range(self.nranks)]

    def render_async(self, stream_ptr, aa=1, max_depth=5, **renderer):
        """Queue this rank's tiles on `stream_ptr` (a cudaStream_t as int; torch's current stream). `renderer`:
        wavefront= / wave_stream= of make_opts (default: the library's choice)."""
        local = self.local
        self.scene.render_device(self.camera, local.data_ptr() if local is not None else self.canvas_ptr, stream=stream_ptr,
                                 stats=False, aa=aa, max_depth=max_depth, tile_rows=self.tile_rows, rank=self.rank,
                                 nranks=self.nranks, **renderer)
        if local is not None:
            self._push(stream_ptr)

    def _push(self, stream_ptr):
        self.pkg._check(self.pkg.cuda_lib().rtc_canvas_push_tiles(self.canvas, self.local.data_ptr(), self.camera.hsize,
                                                                  self.camera.vsize, self.tile_rows, self.rank,
                                                                  self.nranks, stream_ptr))

    def gather_nccl(self):
        """gather="nccl": owned tiles packed by one strided copy, one group of ncclSend / ncclRecv, unpack on rank 0."""
        if self.gather != "nccl":
            return
        gather_tiles(self.dist, self.canvas.view(-1, self.row_elems), self.rank, self.nranks, self.tile_rows,
                     self.recv if self.rank == 0 else None)

    def finish(self, stream_ptr, regather=True):
        """Wait for this rank's frame; if its wavefront queues overflowed, finish() has rendered it again with the
        recursive kernel AFTER the tiles left: send them again (every rank must call with the same `regather`)."""
        st = self.scene.finish(stream_ptr)
        if self.nranks > 1 and regather:
            flag = self.torch.tensor([1 if st.wave_fallback else 0], dtype=self.torch.int32,
                                     device="cuda:%d" % self.device if self.dist.get_backend() == "nccl" else "cpu")
            self.dist.all_reduce(flag, op=self.dist.ReduceOp.MAX)
            if int(flag[0]):
                if self.local is not None and st.wave_fallback:
                    self._push(stream_ptr)
                self.gather_nccl()
                self.torch.cuda.synchronize()
        return st

    def download(self, host_tensor):
        """rank 0: device canvas → pinned host tensor (float64[n])."""
        if self.gather == "ipc":
            self.pkg._check(self.pkg.cuda_lib().rtc_canvas_download(self.canvas, host_tensor.data_ptr(), 0, self.n))
        else:
            host_tensor.copy_(self.canvas[:self.n])
            self.torch.cuda.synchronize()

    def close(self):
        if self.gather == "ipc" and self.canvas is not None:
            self.pkg.cuda_lib().rtc_canvas_destroy(self.canvas)
            self.canvas = None
        self.scene.close()
