"""LidarProjector: Python handle on one sfpe_ctx (one device).

torch tensors are only the buffer carrier (device memory + streams); all work is done
by libsfpe.so through the C ABI (include/sfpe.h).  The batched entry replaces, for many
sweeps at once, what the reference does per sweep in
modules/lidar/process/extract_rosbag_lidar.py:80-86 and
modules/lidar/common/birds_eye_view_generator.py:41-46.
"""
import ctypes as C

import numpy as np

from . import _lib

_NP_OF = {_lib.SFPE_F32: np.float32, _lib.SFPE_F64: np.float64}


def _torch():
    import torch
    return torch


class LidarProjector:
    def __init__(self, device=0, bev_res=(0.2, 0.2), max_range=120, enable_range=True, enable_bev=True,
                 extrinsic=None, slice_edges=None, key_prefix_bits=0, path=0, debug_flags=0, firing_stride=0, **grid):
        """bev_res = (res_y, res_x) like the reference's `res` tuple
        (birds_eye_view_generator.py:43 uses res[1] for x and res[0] for y).
        firing_stride: beams per firing (32 / 64 / 128) when sweeps arrive firing by firing -- a
        performance hint for the fused kernel, never changes a result."""
        self.lib = _lib.load()
        p = _lib.default_params()
        p.bev_res_y, p.bev_res_x = float(bev_res[0]), float(bev_res[1])
        p.bev_range = float(max_range)
        p.enable_range, p.enable_bev = int(enable_range), int(enable_bev)
        for k, v in grid.items():
            if not hasattr(p, k):
                raise TypeError("unknown grid parameter %r" % k)
            setattr(p, k, v)
        if extrinsic is not None:
            m = np.asarray(extrinsic, dtype=np.float64)
            if m.shape == (4, 4):
                m = m[:3]
            if m.shape != (3, 4):
                raise ValueError("extrinsic must be 3x4 or 4x4")
            p.use_extrinsic = 1
            for i, v in enumerate(m.ravel()):
                p.extrinsic[i] = float(v)
        if slice_edges is not None:
            se = [float(v) for v in slice_edges]
            if not 2 <= len(se) <= _lib.MAX_SLICES + 1:
                raise ValueError("need 2..%d slice edges" % (_lib.MAX_SLICES + 1))
            p.n_slices = len(se) - 1
            for i, v in enumerate(se):
                p.slice_edges[i] = v
        p.ablation = int(debug_flags)
        p.firing_stride = int(firing_stride)   # beams per firing of a firing-ordered sweep (performance hint only)
        p.key_prefix_bits = int(key_prefix_bits)
        p.path = int(path)
        self.params = p
        self.H, self.W, self.nbx, self.nby = _lib.grid_dims(p)
        self.n_slices = p.n_slices
        self.device = int(device)
        ctx = C.c_void_p()
        _lib.check(self.lib.sfpe_create(self.device, C.byref(p), C.byref(ctx)))
        self._ctx = ctx
        # the *_host entries of one ctx share its stream and staging buffers: one caller at a time
        # (the drop-in free functions share one projector per parameter set, see _session.py)
        import threading
        self._host_lock = threading.Lock()

    # ------------------------------------------------------------------ lifecycle
    def close(self):
        if getattr(self, "_ctx", None):
            self.lib.sfpe_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self):
        return int(self.lib.sfpe_launch_count(self._ctx))

    @property
    def last_path(self):
        """Code path of the last project(): "fused", "fused+ext" (fused kernel + the height-slice / intensity
        kernels of the extension) or "generic"."""
        return self.lib.sfpe_last_path(self._ctx).decode()

    PHASES = ("key_init", "point_pass", "deferred_fp64", "range_resolve", "bev_compose", "bev_redo", "bev_line_write", "settle",
              "wait_store_warp", "bev_line_list", "tail_write", "unused")

    def phase_profile(self, enable=True):
        """Switch the fused kernel's per-phase SM-cycle counters on (and reset them) or off."""
        _lib.check(self.lib.sfpe_phase_profile(self._ctx, int(bool(enable))), self._ctx)

    def phase_cycles(self):
        """dict phase -> SM cycles summed over all CTAs since phase_profile(True)."""
        buf = (C.c_int64 * len(self.PHASES))()
        _lib.check(self.lib.sfpe_phase_cycles(self._ctx, buf), self._ctx)
        return dict(zip(self.PHASES, [int(v) for v in buf]))

    def selftest_fastpath(self, points):
        """points: CUDA float32 (N,4) contiguous.  Returns the four guard-band utilisation
        fractions (column, row, range bucket, BEV bin) of sfpe_selftest_fastpath."""
        buf = (C.c_double * 4)()
        _lib.check(self.lib.sfpe_selftest_fastpath(self._ctx, points.data_ptr(), int(points.shape[0]), buf), self._ctx)
        return [float(v) for v in buf]

    # ------------------------------------------------------------------ batched, device resident
    def alloc_outputs(self, n_frames, range_dtype="f32", range_layout="hwc", bev_dtype="f32",
                      winner=False, counts=False):
        """Allocate the output tensors of project() on this projector's device."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        tdt = {"f32": torch.float32, "f64": torch.float64}
        out = {}
        p = self.params
        if p.enable_range:
            shape = (n_frames, self.H, self.W, 3) if range_layout == "hwc" else (n_frames, 3, self.H, self.W)
            out["range"] = torch.empty(shape, dtype=tdt[range_dtype], device=dev)
            if winner:
                out["winner"] = torch.empty((n_frames, self.H, self.W), dtype=torch.int32, device=dev)
        if p.enable_bev:
            out["bev"] = torch.empty((n_frames, self.nbx, self.nby), dtype=tdt[bev_dtype], device=dev)
            if counts:
                out["counts"] = torch.empty((n_frames, self.nbx, self.nby), dtype=torch.int32, device=dev)
            if self.n_slices:
                out["bev_height"] = torch.empty((n_frames, self.n_slices, self.nbx, self.nby),
                                                dtype=torch.float32, device=dev)
                out["bev_intensity"] = torch.empty((n_frames, self.nbx, self.nby), dtype=torch.float32, device=dev)
        out["status"] = torch.empty((n_frames,), dtype=torch.int32, device=dev)
        out["dropped"] = torch.empty((n_frames,), dtype=torch.int32, device=dev)
        out["_layout"] = range_layout
        return out

    @staticmethod
    def _points_desc(points, columns=(0, 1, 2, 3)):
        """sfpe_points for a 2-D CUDA tensor whose rows are points."""
        torch = _torch()
        if points.dim() != 2 or points.shape[1] <= max(columns):
            raise ValueError("points must be (P, >=4)")
        if points.dtype == torch.float32:
            dt, es = _lib.SFPE_F32, 4
        elif points.dtype == torch.float64:
            dt, es = _lib.SFPE_F64, 8
        else:
            raise TypeError("points must be float32 or float64")
        if points.stride(1) != 1:
            raise ValueError("point fields must be contiguous within a row")
        d = _lib.Points()
        d.data = points.data_ptr()
        d.stride_bytes = points.stride(0) * es if points.shape[0] > 1 else points.shape[1] * es
        d.dtype = dt
        d.off_x, d.off_y, d.off_z, d.off_i = [c * es for c in columns]
        return d

    def project(self, points, n_frames, frame_offsets=None, uniform_n=None, max_n=None, out=None,
                stream=None, **alloc_kw):
        """Rasterise `n_frames` sweeps stored back to back in `points` (CUDA tensor (P,4)).

        frame_offsets: CUDA int64 (n_frames+1) for ragged batches, else every frame has
        `uniform_n` points.  Asynchronous on `stream` (default: torch's current stream).
        Returns the dict of output tensors (see alloc_outputs)."""
        torch = _torch()
        if not points.is_cuda or points.device.index != self.device:
            raise ValueError("points must live on cuda:%d" % self.device)
        if out is None:
            out = self.alloc_outputs(n_frames, **alloc_kw)
        desc = self._points_desc(points)
        if frame_offsets is not None:
            if frame_offsets.dtype != torch.int64 or not frame_offsets.is_cuda or frame_offsets.numel() != n_frames + 1:
                raise ValueError("frame_offsets must be a CUDA int64 tensor of n_frames+1 entries")
            if max_n is None:
                raise ValueError("max_n (upper bound of points per frame) is required with frame_offsets")
            offs_ptr, un = frame_offsets.data_ptr(), 0
        else:
            if uniform_n is None:
                if n_frames == 0 or points.shape[0] % max(n_frames, 1):
                    raise ValueError("uniform_n required")
                uniform_n = points.shape[0] // n_frames
            if uniform_n * n_frames > points.shape[0]:
                raise ValueError("points tensor too small for n_frames x uniform_n")
            offs_ptr, un, max_n = None, int(uniform_n), int(uniform_n)
        o = _lib.Outputs()
        ptr = lambda k: out[k].data_ptr() if k in out and out[k] is not None else None
        o.range = ptr("range")
        if "range" in out:
            o.range_dtype = _lib.SFPE_F32 if out["range"].dtype == torch.float32 else _lib.SFPE_F64
        o.range_layout = _lib.RANGE_HWC if out.get("_layout", "hwc") == "hwc" else _lib.RANGE_CHW
        o.range_winner = ptr("winner")
        o.bev = ptr("bev")
        if "bev" in out:
            o.bev_dtype = _lib.SFPE_F32 if out["bev"].dtype == torch.float32 else _lib.SFPE_F64
        o.bev_counts = ptr("counts")
        o.bev_height = ptr("bev_height")
        o.bev_intensity = ptr("bev_intensity")
        o.status = ptr("status")
        o.dropped = ptr("dropped")
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        rc = self.lib.sfpe_project_batch(self._ctx, C.byref(desc), offs_ptr, un, int(max_n), int(n_frames),
                                         C.byref(o), C.c_void_p(int(stream)))
        _lib.check(rc, self._ctx)
        return out

    # ------------------------------------------------------------------ PointCloud2 blob -> packed points
    def decode_pointcloud2(self, blob, n, point_step, offsets=(0, 4, 8, 16), dtype="f32", out=None, stream=None):
        """blob: CUDA uint8 tensor holding a sensor_msgs/PointCloud2 `data` buffer (n points of
        point_step bytes; x,y,z,intensity at `offsets`, FLOAT32 or FLOAT64).  Returns a CUDA
        float32 (n,4) tensor -- what np.array(list(read_points(msg))) yields on the host in
        team_sf_rosnode/scripts/lidar_predict.py:64-65, without the Python per-point loop."""
        torch = _torch()
        if not blob.is_cuda or blob.device.index != self.device or blob.numel() < n * point_step:
            raise ValueError("blob must be a CUDA byte tensor of at least n*point_step bytes on cuda:%d" % self.device)
        if out is None:
            out = torch.empty((n, 4), dtype=torch.float32, device=blob.device)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        dt = _lib.SFPE_F32 if dtype == "f32" else _lib.SFPE_F64
        rc = self.lib.sfpe_decode_pointcloud2(self._ctx, blob.data_ptr(), int(n), int(point_step), *[int(o) for o in offsets],
                                              dt, out.data_ptr(), C.c_void_p(int(stream)))
        _lib.check(rc, self._ctx)
        return out

    # ------------------------------------------------------------------ one-byte BEV transport (host consumers)
    def density_lut(self):
        """The 65 density values the kernels write (birds_eye_view_generator.py:31-37), as float64."""
        lut = np.empty(65, dtype=np.float64)
        _lib.check(self.lib.sfpe_density_lut(self._ctx, lut.ctypes.data), self._ctx)
        return lut

    def pack_bev_u8(self, bev, codes=None, flag=None, stream=None):
        """bev: contiguous CUDA float32/float64 density image(s) written by project().  Returns
        (codes, flag): uint8 class per cell and a device int32 that becomes 1 if a value is not a
        density value.  unpack_bev_host(codes.cpu(), ...) gives the image back bit for bit."""
        torch = _torch()
        if not bev.is_cuda or bev.device.index != self.device or not bev.is_contiguous():
            raise ValueError("bev must be a contiguous CUDA tensor on cuda:%d" % self.device)
        if bev.dtype not in (torch.float32, torch.float64):
            raise ValueError("bev must be float32 or float64")
        if codes is None:
            codes = torch.empty(bev.shape, dtype=torch.uint8, device=bev.device)
        if flag is None:
            flag = torch.zeros(1, dtype=torch.int32, device=bev.device)
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        dt = _lib.SFPE_F32 if bev.dtype == torch.float32 else _lib.SFPE_F64
        rc = self.lib.sfpe_bev_pack_u8(self._ctx, bev.data_ptr(), dt, int(bev.numel()), codes.data_ptr(), flag.data_ptr(),
                                       C.c_void_p(int(stream)))
        _lib.check(rc, self._ctx)
        return codes, flag

    # ------------------------------------------------------------------ one sweep, host in / host out
    def project_sweep_host(self, points, columns=(0, 1, 2, 3), range_dtype=np.float64, range_layout="chw",
                           bev_dtype=np.float64, want_range=True, want_bev=True, want_winner=False):
        """points: C-contiguous numpy (N, >=4) float32/float64 on the host.  Returns
        (range, winner, bev, status, dropped) as numpy arrays (None where not wanted).
        dtype contract: every decision is evaluated in float64 on the VALUES given, as the reference
        does on the float64 array its callers build (np.array(list(read_points(msg))),
        lidar_predict.py:64-65); a float32 array is widened first and gives the same images as the
        float64 array of the same values (numpy itself would evaluate a float32 array in float32)."""
        a = np.asarray(points)
        if a.ndim != 2 or a.shape[1] <= max(columns):
            raise ValueError("points must be (N, >=4)")
        if a.dtype not in (np.float32, np.float64):
            a = a.astype(np.float64)
        a = np.ascontiguousarray(a)
        es = a.dtype.itemsize
        d = _lib.Points()
        d.stride_bytes = a.shape[1] * es
        d.dtype = _lib.SFPE_F32 if a.dtype == np.float32 else _lib.SFPE_F64
        d.off_x, d.off_y, d.off_z, d.off_i = [c * es for c in columns]
        p = self.params
        rdt = _lib.SFPE_F32 if np.dtype(range_dtype) == np.float32 else _lib.SFPE_F64
        bdt = _lib.SFPE_F32 if np.dtype(bev_dtype) == np.float32 else _lib.SFPE_F64
        rng = win = bev = None
        if want_range and p.enable_range:
            shape = (3, self.H, self.W) if range_layout == "chw" else (self.H, self.W, 3)
            rng = np.empty(shape, dtype=_NP_OF[rdt])
            if want_winner:
                win = np.empty((self.H, self.W), dtype=np.int32)
        if want_bev and p.enable_bev:
            bev = np.empty((self.nbx, self.nby), dtype=_NP_OF[bdt])
        st, dr = C.c_int32(0), C.c_int32(0)
        vp = lambda x: x.ctypes.data_as(C.c_void_p) if x is not None else None
        with self._host_lock:
            rc = self.lib.sfpe_project_sweep_host(
                self._ctx, vp(a) if len(a) else None, C.byref(d), len(a), vp(rng), rdt,
                _lib.RANGE_CHW if range_layout == "chw" else _lib.RANGE_HWC, vp(win), vp(bev), bdt,
                C.byref(st), C.byref(dr))
        _lib.check(rc, self._ctx)
        return rng, win, bev, st.value, dr.value



def unpack_bev_host(codes, lut, out=None, dtype=np.float32, threads=0):
    """Host side of the one-byte BEV transport: out[i] = dtype(lut[codes[i]]) on `threads` host
    threads (0: all cores).  codes: C-contiguous uint8 array; lut: LidarProjector.density_lut()."""
    lib = _lib.load()
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    lut = np.ascontiguousarray(lut, dtype=np.float64)
    if lut.size != 65:
        raise ValueError("lut must hold 65 values")
    if out is None:
        out = np.empty(codes.shape, dtype=dtype)
    if out.size != codes.size or not out.flags.c_contiguous or out.dtype not in (np.float32, np.float64):
        raise ValueError("out must be a C-contiguous float32/float64 array of codes.size elements")
    dt = _lib.SFPE_F32 if out.dtype == np.float32 else _lib.SFPE_F64
    _lib.check(lib.sfpe_bev_unpack_host(codes.ctypes.data, int(codes.size), lut.ctypes.data, out.ctypes.data, dt, int(threads)))
    return out
