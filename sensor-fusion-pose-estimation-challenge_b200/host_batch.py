"""Host-buffer batch entry: pinned host sweeps in, pinned host images out.

This is the "reference-facing" way to use the library for many sweeps (what
extract_rosbag.py:162-187 does one message at a time on the CPU): the batch is cut into
chunks that flow through three CUDA streams -- H2D copy, kernels, D2H copy -- with two
device slots, so both PCIe directions and the SMs work concurrently.  torch is used for
pinned/device memory, streams and events only.
"""
from concurrent.futures import ThreadPoolExecutor

from . import _lib
from .projector import LidarProjector


def host_threads_for_this_rank():
    """Host threads one rank may use: the cores this process may run on (its affinity mask), divided
    among the ranks that share the machine when the launcher did not already partition the cores
    (LOCAL_WORLD_SIZE, set by torchrun)."""
    import os
    try:
        cores = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        cores = os.cpu_count() or 1
    local_world = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    if cores == (os.cpu_count() or cores):      # not partitioned by affinity: share by rank count
        cores = max(1, cores // local_world)
    return cores


class HostBatchPipeline:
    def __init__(self, projector: LidarProjector, n_frames, points_per_frame, chunk=32, bev_transport="dense",
                 unpack_threads=0, unpack_workers=2, cmap="gray"):
        """bev_transport "dense": the float BEV crosses PCIe as it is (4 bytes per cell).
        "u8": one byte per cell (sfpe_bev_pack_u8 on the device after the projection), expanded into
        the same pinned float image by host threads while later chunks are in flight
        (sfpe_bev_unpack_host) -- the host image is bit-identical, the D2H volume 3x smaller.
        "rgb": the result type of the reference's own BEV entry (birds_eye_view_generator.py:48-54):
        the same one byte per cell over PCIe, expanded on the host into the uint8 (nbx, nby, 3)
        picture h_bev_rgb through the colours of the 65 density values (colormap `cmap`, parity
        unpinned like every RGB output); no float image is written on the host at all."""
        import torch
        if bev_transport not in ("dense", "u8", "rgb"):
            raise ValueError("bev_transport must be 'dense', 'u8' or 'rgb'")
        self.torch = torch
        self.pr = projector
        self.n_frames, self.n = int(n_frames), int(points_per_frame)
        self.chunk = max(1, min(int(chunk), self.n_frames))
        dev = torch.device("cuda", projector.device)
        self.dev = dev
        self.s_in, self.s_k, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
        self.d_in = [torch.empty((self.chunk * self.n, 4), dtype=torch.float32, device=dev) for _ in range(2)]
        self.d_out = [projector.alloc_outputs(self.chunk) for _ in range(2)]
        p = projector
        self.h_range = torch.empty((self.n_frames, p.H, p.W, 3), dtype=torch.float32).pin_memory() \
            if p.params.enable_range else None
        self.rgb = bev_transport == "rgb" and bool(p.params.enable_bev)
        self.h_bev = torch.empty((self.n_frames, p.nbx, p.nby), dtype=torch.float32).pin_memory() \
            if p.params.enable_bev and not self.rgb else None
        self.h_bev_rgb = torch.empty((self.n_frames, p.nbx, p.nby, 3), dtype=torch.uint8).pin_memory() if self.rgb else None
        self.h_status = torch.empty((self.n_frames,), dtype=torch.int32).pin_memory()
        self.u8 = (bev_transport == "u8" and self.h_bev is not None) or self.rgb
        self.cells = p.nbx * p.nby
        self._fut = []
        # events of the two device slots persist across run() calls: the first chunks of a run wait
        # for the last kernels / copies of the previous run on their slot
        self._ev_k = [None, None]
        self._ev_out = [None, None]
        if self.u8:
            self.d_codes = [torch.empty((self.chunk, p.nbx, p.nby), dtype=torch.uint8, device=dev) for _ in range(2)]
            self.d_flag = torch.zeros(1, dtype=torch.int32, device=dev)
            self.h_codes = torch.empty((self.n_frames, p.nbx, p.nby), dtype=torch.uint8).pin_memory()
            self.lut = projector.density_lut()
            # two chunks expand side by side (one worker waits for its copy while the other writes); the
            # cores are split between them.  Measured on the B200 box (16 cores): host DRAM is the limit.
            workers = max(1, int(unpack_workers))
            self.unpack_threads = int(unpack_threads) if unpack_threads > 0 else max(1, host_threads_for_this_rank() // workers)
            self._pool = ThreadPoolExecutor(max_workers=max(1, int(unpack_workers)))
            self._unpack = projector.lib.sfpe_bev_unpack_host
            self._unpack_rgb = projector.lib.sfpe_bev_unpack_rgb_host
            if self.rgb:
                from . import colormap
                import numpy as np
                lut_dev = torch.from_numpy(self.lut).to(dev)
                self.rgb_table = np.ascontiguousarray(
                    colormap.device_rgb(projector, lut_dev, cmap, 0, 1500 * projector.params.bev_res_y).cpu().numpy())
        self.h2d_bytes = self.n_frames * self.n * 16
        bev_bytes = 0 if not p.params.enable_bev else self.n_frames * self.cells * (1 if self.u8 else 4)
        self.d2h_bytes = (self.h_range.numel() * 4 if self.h_range is not None else 0) + bev_bytes + self.n_frames * 4

    def host_outputs(self):
        """The pipeline's own pinned host outputs as the dict run(out=...) takes."""
        return {"range": self.h_range, "bev": self.h_bev, "bev_rgb": self.h_bev_rgb, "status": self.h_status,
                "codes": getattr(self, "h_codes", None)}

    def alloc_host_outputs(self):
        """A second set of pinned host outputs, so that successive runs need not overwrite each other."""
        pin = lambda t: None if t is None else self.torch.empty_like(t).pin_memory()
        return {k: pin(v) for k, v in self.host_outputs().items()}

    def _expand(self, ev, f0, nf, out):
        """Worker thread: wait for the chunk's D2H, then rebuild its images (float or RGB) on the host."""
        ev.synchronize()
        n = nf * self.cells
        if self.rgb:
            _lib.check(self._unpack_rgb(out["codes"].data_ptr() + f0 * self.cells, n, self.rgb_table.ctypes.data,
                                        out["bev_rgb"].data_ptr() + f0 * self.cells * 3, self.unpack_threads))
        else:
            _lib.check(self._unpack(out["codes"].data_ptr() + f0 * self.cells, n, self.lut.ctypes.data,
                                    out["bev"].data_ptr() + f0 * self.cells * 4, _lib.SFPE_F32, self.unpack_threads))

    def close(self):
        """Stop the expanding workers (after the last synchronize())."""
        if self.u8 and self._pool is not None:
            self._pool.shutdown(wait=True)
            self._pool = None

    def synchronize(self):
        """Everything of the runs issued so far is in the pinned host outputs."""
        for f in self._fut:
            if f is not None:
                f.result()
        self.torch.cuda.synchronize(self.dev)
        if self.u8 and int(self.d_flag.item()) != 0:
            raise RuntimeError("sfpe_bev_pack_u8 met a value that is not a density value")

    def run(self, pinned_points, out=None):
        """pinned_points: pinned CPU float32 tensor (n_frames*n, 4).  Asynchronous; results are in
        self.h_range / self.h_bev (or h_bev_rgb) / self.h_status -- or in the tensors of `out`
        (alloc_host_outputs()) -- after self.synchronize().  Runs may be issued back to back: the
        device slots are ordered across runs by their events."""
        torch = self.torch
        if out is None:
            out = self.host_outputs()
        ev_in = [None, None]; ev_k = self._ev_k; ev_out = self._ev_out
        n_chunks = (self.n_frames + self.chunk - 1) // self.chunk
        if len(self._fut) != n_chunks:
            self._fut = [None] * n_chunks
        for c in range(n_chunks):
            s = c & 1
            f0 = c * self.chunk
            nf = min(self.chunk, self.n_frames - f0)
            with torch.cuda.stream(self.s_in):
                if ev_k[s] is not None:
                    self.s_in.wait_event(ev_k[s])            # kernels of chunk c-2 are done with d_in[s]
                self.d_in[s][: nf * self.n].copy_(pinned_points[f0 * self.n:(f0 + nf) * self.n], non_blocking=True)
                ev_in[s] = torch.cuda.Event(); ev_in[s].record(self.s_in)
            with torch.cuda.stream(self.s_k):
                self.s_k.wait_event(ev_in[s])
                if ev_out[s] is not None:
                    self.s_k.wait_event(ev_out[s])           # D2H of chunk c-2 is done with d_out[s]
                self.pr.project(self.d_in[s], n_frames=nf, uniform_n=self.n, out=self.d_out[s],
                                stream=self.s_k.cuda_stream)
                if self.u8:
                    self.pr.pack_bev_u8(self.d_out[s]["bev"][:nf], codes=self.d_codes[s][:nf], flag=self.d_flag,
                                        stream=self.s_k.cuda_stream)
                ev_k[s] = torch.cuda.Event(); ev_k[s].record(self.s_k)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(ev_k[s])
                o = self.d_out[s]
                if out["range"] is not None:
                    out["range"][f0:f0 + nf].copy_(o["range"][:nf], non_blocking=True)
                if self.u8:
                    if self._fut[c] is not None:
                        self._fut[c].result()                # an earlier run's expansion has read codes[f0:]
                    out["codes"][f0:f0 + nf].copy_(self.d_codes[s][:nf], non_blocking=True)
                elif out["bev"] is not None:
                    out["bev"][f0:f0 + nf].copy_(o["bev"][:nf], non_blocking=True)
                out["status"][f0:f0 + nf].copy_(o["status"][:nf], non_blocking=True)
                ev_out[s] = torch.cuda.Event(); ev_out[s].record(self.s_out)
                if self.u8:
                    self._fut[c] = self._pool.submit(self._expand, ev_out[s], f0, nf, out)
        return self
