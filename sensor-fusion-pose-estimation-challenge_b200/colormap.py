"""Colormap stage on the device (sfpe_colormap_host).  PARITY UNPINNED -- see include/sfpe.h:
the reference delegates this to matplotlib (extract_rosbag_lidar.py:41,49-53,75;
birds_eye_view_generator.py:48-53), which it neither vendors nor pins and which is not
installed here."""
import ctypes as C

import numpy as np

from . import _lib, _session

CMAPS = {"gray": 0, "jet": 1}


def to_rgb_bytes(img, cmap, vmin=None, vmax=None):
    """float image (any shape) -> uint8 image of shape img.shape + (3,)."""
    if cmap not in CMAPS:
        raise ValueError("colormap %r is not available (have: %s)" % (cmap, ", ".join(sorted(CMAPS))))
    a = np.ascontiguousarray(img)
    if a.dtype not in (np.float32, np.float64):
        a = a.astype(np.float64)
    pr = _session.projector(enable_bev=False)
    out = np.empty(a.shape + (3,), dtype=np.uint8)
    auto = vmin is None or vmax is None
    with pr._host_lock:
        rc = pr.lib.sfpe_colormap_host(
            pr._ctx, a.ctypes.data_as(C.c_void_p), _lib.SFPE_F32 if a.dtype == np.float32 else _lib.SFPE_F64,
            a.size, CMAPS[cmap], int(auto), 0.0 if auto else float(vmin), 0.0 if auto else float(vmax),
            out.ctypes.data_as(C.c_void_p))
    _lib.check(rc, pr._ctx)
    return out


def device_rgb(pr, img, cmap, vmin=None, vmax=None, stream=None):
    """CUDA float32/float64 image (any shape, contiguous) -> CUDA uint8 image img.shape + (3,) through
    sfpe_colormap on `pr`'s context: the float image never leaves the device."""
    import torch
    if cmap not in CMAPS:
        raise ValueError("colormap %r is not available (have: %s)" % (cmap, ", ".join(sorted(CMAPS))))
    if not img.is_cuda or not img.is_contiguous() or img.dtype not in (torch.float32, torch.float64):
        raise ValueError("img must be a contiguous CUDA float32/float64 tensor")
    out = torch.empty(tuple(img.shape) + (3,), dtype=torch.uint8, device=img.device)
    auto = vmin is None or vmax is None
    if stream is None:
        stream = torch.cuda.current_stream(pr.device).cuda_stream
    rc = pr.lib.sfpe_colormap(
        pr._ctx, C.c_void_p(img.data_ptr()), _lib.SFPE_F32 if img.dtype == torch.float32 else _lib.SFPE_F64,
        img.numel(), CMAPS[cmap], int(auto), 0.0 if auto else float(vmin), 0.0 if auto else float(vmax),
        C.c_void_p(out.data_ptr()), C.c_void_p(int(stream)))
    _lib.check(rc, pr._ctx)
    return out
