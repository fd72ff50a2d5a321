"""Drop-in for modules/lidar/common/birds_eye_view_generator.py:31-54.

`generate_birds_eye_view` keeps the reference signature and returns the uint8 (nb,nb,3)
image; `birds_eye_view_density` exposes the pre-colormap float64 image, which is the part
of the result the reference pins (the colormap bytes come from matplotlib, see colormap.py).
"""
import numpy as np

from .. import _session, colormap

LIDAR_MAX_DENSITY_SQ_METER = 1500


def _points_and_projector(arrPoints, res, max_range):
    points = np.array(arrPoints)
    if points.ndim != 2 or points.shape[1] < 2:
        raise IndexError("too many indices for array")
    if points.shape[1] < 4:  # the reference only reads columns 0 and 1
        pad = np.zeros((points.shape[0], 4), dtype=np.float64)
        pad[:, :points.shape[1]] = points
        points = pad
    pr = _session.projector(enable_range=False, bev_res=(float(res[0]), float(res[1])), max_range=float(max_range))
    return points, pr


def birds_eye_view_density(arrPoints, res, max_range=120):
    """float64 (nbx, nby): 255*min(ln(N+1)/ln 64, 1) per cell, both axes reversed (:43-46)."""
    points, pr = _points_and_projector(arrPoints, res, max_range)
    _, _, bev, _, _ = pr.project_sweep_host(points, bev_dtype=np.float64, want_range=False)
    return bev


def generate_birds_eye_view(arrPoints, t, res, max_range=120, cmap='gray'):
    """uint8 (nbx, nby, 3) as :41-54.  One device pass: points up, histogram + normalize + colormap on
    the device, only the RGB bytes come back (the float image of :45 never crosses PCIe)."""
    if cmap not in colormap.CMAPS:
        raise ValueError("colormap %r is not available (have: %s)" % (cmap, ", ".join(sorted(colormap.CMAPS))))
    points, pr = _points_and_projector(arrPoints, res, max_range)
    vmax = LIDAR_MAX_DENSITY_SQ_METER * res[0]
    if len(points) == 0:
        return colormap.to_rgb_bytes(np.zeros((pr.nbx, pr.nby)), cmap, vmin=0, vmax=vmax)
    import torch
    if points.dtype not in (np.float32, np.float64):
        points = points.astype(np.float64)
    with pr._host_lock:       # the shared projector's scratch: one caller at a time
        d_pts = torch.from_numpy(np.ascontiguousarray(points)).to(torch.device("cuda", pr.device))
        out = pr.project(d_pts, n_frames=1, uniform_n=len(points), bev_dtype="f64")
        rgb = colormap.device_rgb(pr, out["bev"][0], cmap, 0, vmax)
        return rgb.cpu().numpy()
