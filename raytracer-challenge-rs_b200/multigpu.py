"""Row-tile sharding of one frame over the GPUs of a box (one process per GPU, torch.distributed).

The reference shards contiguous row blocks over OS threads (camera.rs:155-172) and sends each
block back over an mpsc channel (:195-207). Here rows are grouped in tiles of `tile_rows` rows and
tile t belongs to rank t % nranks (cyclic, so expensive image regions spread over all GPUs); the
scene is replicated. The gather of the canvas to rank 0 is fused into the render kernel: rank 0
owns the device canvas, exports it with CUDA IPC, and every rank's kernel stores its finished
pixels straight into it over NVLink (peer-mapped pointer) — no collective launch on the data path.
`gather="nccl"` is the plain variant: each rank renders into a zeroed local canvas and the canvases
are summed onto rank 0 with one NCCL reduce (disjoint supports → the sum is the gather).
"""
import numpy as np


def tile_owner(row, tile_rows, nranks):
    return (row // tile_rows) % nranks


def owned_rows(vsize, tile_rows, rank, nranks):
    """Frame rows rendered by `rank` (sorted)."""
    return [y for y in range(vsize) if tile_owner(y, tile_rows, nranks) == rank]


def owned_row_mask(vsize, tile_rows, rank, nranks):
    y = np.arange(vsize)
    return ((y // tile_rows) % nranks) == rank


class ShardedRenderer:
    """One rank's side of the sharded render. Needs torch with CUDA and an initialised process group."""

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
        self._setup_canvas()

    def _setup_canvas(self):
        import ctypes as C
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
            # >= 3 ranks: scenes that recurse are rendered as a wavefront, whose pixels complete in 24-byte pieces
            # from many passes — ranks other than 0 then render into a local canvas and push their tiles to
            # rank 0 with one strided DMA (rtc_canvas_push_tiles) instead of storing across NVLink piecewise
            self.local = None
            if self.rank != 0 and self.nranks >= 3 and self.push_tiles:
                self.local = torch.zeros(self.n, dtype=torch.float64, device="cuda:%d" % self.device)
        else:
            self.canvas = torch.zeros(self.n, dtype=torch.float64, device="cuda:%d" % self.device)
            self.canvas_ptr = self.canvas.data_ptr()

    def render_async(self, stream_ptr, aa=1, max_depth=5):
        """Queue this rank's tiles on `stream_ptr` (a cudaStream_t as int; torch's current stream)."""
        if self.gather == "nccl":
            self.canvas.zero_()  # rank 0 holds last step's gathered frame; the reduce needs disjoint supports
        local = getattr(self, "local", None)
        self.scene.render_device(self.camera, local.data_ptr() if local is not None else self.canvas_ptr, stream=stream_ptr,
                                 stats=False, aa=aa, max_depth=max_depth, tile_rows=self.tile_rows, rank=self.rank,
                                 nranks=self.nranks)
        if local is not None:
            self.pkg._check(self.pkg.cuda_lib().rtc_canvas_push_tiles(self.canvas, local.data_ptr(), self.camera.hsize,
                                                                      self.camera.vsize, self.tile_rows, self.rank,
                                                                      self.nranks, stream_ptr))

    def gather_nccl(self):
        if self.gather == "nccl":
            self.dist.reduce(self.canvas, dst=0, op=self.dist.ReduceOp.SUM)

    def finish(self, stream_ptr):
        st = self.scene.finish(stream_ptr)
        local = getattr(self, "local", None)
        if local is not None and st.launches == 1:
            # one launch: the recursive kernel ran — possibly as the overflow fallback inside finish(), after
            # the tiles were pushed; push what is in the local canvas now
            lib = self.pkg.cuda_lib()
            self.pkg._check(lib.rtc_canvas_push_tiles(self.canvas, local.data_ptr(), self.camera.hsize, self.camera.vsize,
                                                      self.tile_rows, self.rank, self.nranks, stream_ptr))
            self.torch.cuda.synchronize()
        return st

    def download(self, host_tensor):
        """rank 0: device canvas → pinned host tensor (float64[n])."""
        import ctypes as C
        if self.gather == "ipc":
            self.pkg._check(self.pkg.cuda_lib().rtc_canvas_download(self.canvas, host_tensor.data_ptr(), 0, self.n))
        else:
            host_tensor.copy_(self.canvas)
            self.torch.cuda.synchronize()

    def close(self):
        if self.gather == "ipc" and self.canvas is not None:
            self.pkg.cuda_lib().rtc_canvas_destroy(self.canvas)
            self.canvas = None
        self.scene.close()
