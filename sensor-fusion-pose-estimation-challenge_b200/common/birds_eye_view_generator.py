"""Drop-in for modules/lidar/common/birds_eye_view_generator.py:31-54.

`generate_birds_eye_view` keeps the reference signature and returns the uint8 (nb,nb,3)
image; `birds_eye_view_density` exposes the pre-colormap float64 image, which is the part
of the result the reference pins (the colormap bytes come from matplotlib, see colormap.py).
"""
import numpy as np

from .. import _session, colormap

LIDAR_MAX_DENSITY_SQ_METER = 1500


def birds_eye_view_density(arrPoints, res, max_range=120):
    """float64 (nbx, nby): 255*min(ln(N+1)/ln 64, 1) per cell, both axes reversed (:43-46)."""
    points = np.array(arrPoints)
    if points.ndim != 2 or points.shape[1] < 2:
        raise IndexError("too many indices for array")
    if points.shape[1] < 4:  # the reference only reads columns 0 and 1
        pad = np.zeros((points.shape[0], 4), dtype=np.float64)
        pad[:, :points.shape[1]] = points
        points = pad
    pr = _session.projector(enable_range=False, bev_res=(float(res[0]), float(res[1])), max_range=float(max_range))
    _, _, bev, _, _ = pr.project_sweep_host(points, bev_dtype=np.float64, want_range=False)
    return bev


def generate_birds_eye_view(arrPoints, t, res, max_range=120, cmap='gray'):
    density = birds_eye_view_density(arrPoints, res, max_range)
    return colormap.to_rgb_bytes(density, cmap, vmin=0, vmax=LIDAR_MAX_DENSITY_SQ_METER * res[0])
