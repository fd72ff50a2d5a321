"""Drop-in for modules/lidar/process/extract_rosbag_lidar.py:18-86 (projection functions only).

Same names, arguments, return layout and error behaviour as the reference; the work is
done on the GPU by libsfpe.so (one H2D copy of the sweep, kernels, one D2H copy of the
three float64 (32,1801) images).  The bag-walking CLI of the reference module is out of
scope.

dtype contract: the reference's callers hand over float64 arrays (np.array(list(read_points(msg))));
every cell decision here is made in float64 on the values given.  A float32 array is widened first,
i.e. it gives the images of the float64 array of the same values (numpy would evaluate a float32
array in float32 arithmetic, which no caller of the reference does).  The functions may be called
from several threads (one projector per parameter set, calls serialised per projector).
"""
import numpy as np

from .. import _lib, _session, colormap
from . import globals  # noqa: A004  (the reference's module name)

_CH = {"distance": 0, "height": 1, "intensity": 2}


def _range_images(points):
    """float64 (3,32,1801) [distance, height, intensity] of one sweep; raises IndexError
    where the reference's fancy assignment would (extract_rosbag_lidar.py:71)."""
    pts = np.asarray(points)
    if pts.ndim != 2 or pts.shape[1] < 4:
        raise IndexError("points must be (N, >=4): too many indices for array")
    pr = _session.projector(enable_bev=False)
    rng, _, _, status, _ = pr.project_sweep_host(pts, range_dtype=np.float64, range_layout="chw", want_bev=False)
    if status:
        what = []
        if status & _lib.FRAME_ROW_OOB:
            what.append("elevation cell out of bounds for axis 0 with size %d" % (globals.Y_MAX + 1))
        if status & _lib.FRAME_COL_OOB:
            what.append("azimuth cell out of bounds for axis 1 with size %d" % (globals.X_MAX + 1))
        if status & _lib.FRAME_NAN:
            what.append("NaN cell coordinate (NaN input or a point at the origin)")
        raise IndexError("index out of bounds: " + "; ".join(what))
    return rng


def _finish(img, kind, cmap):
    if cmap is None:
        return img
    if kind == "height":
        return colormap.to_rgb_bytes(img, cmap, vmin=globals.LIDAR_MIN_HEIGHT, vmax=globals.LIDAR_MAX_HEIGHT)
    return colormap.to_rgb_bytes(img, cmap)  # autoscale to the image's min/max


def lidar_2d_front_view(points, res, fov, type, cmap=None):  # noqa: A002
    assert len(res) == 2, "res must be list/tuple of length 2"
    assert len(fov) == 2, "fov must be list/tuple of length 2"
    assert fov[0] <= 0, "first element in fov must be 0 or negative"
    assert type in ['intensity', 'height', 'distance'], "type must be 'intensity', 'height', or 'distance'"
    # like the reference, the grid comes from `globals`; `res` / `fov` are only validated
    img = np.ascontiguousarray(_range_images(points)[_CH[type]])
    return _finish(img, type, cmap), img


def generate_lidar_2d_front_view(points, cmap=None):
    rng = _range_images(points)  # one device pass yields all three channels
    out = {}
    for kind in ("intensity", "distance", "height"):
        img = np.ascontiguousarray(rng[_CH[kind]])
        out[kind] = _finish(img, kind, cmap)
        out[kind + "_float"] = img
    return out
