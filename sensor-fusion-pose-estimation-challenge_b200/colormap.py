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
    rc = pr.lib.sfpe_colormap_host(
        pr._ctx, a.ctypes.data_as(C.c_void_p), _lib.SFPE_F32 if a.dtype == np.float32 else _lib.SFPE_F64,
        a.size, CMAPS[cmap], int(auto), 0.0 if auto else float(vmin), 0.0 if auto else float(vmax),
        out.ctypes.data_as(C.c_void_p))
    _lib.check(rc, pr._ctx)
    return out
