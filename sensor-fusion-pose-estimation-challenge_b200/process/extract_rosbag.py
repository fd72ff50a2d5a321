"""The PointCloud2 branch of ROSBagExtractor.handle_msg (modules/lidar/process/extract_rosbag.py:162-187),
without the bag walking, file writing and pyglet display around it.

Per `/velodyne_points` message the reference decodes the message point by point in Python (:165-166), renders
the top-down density image (:169-170) and the three 360-degree images (:173), each through its own numpy pass
and matplotlib colormap.  Here the message's `data` bytes go to the device once; decode, both rasterisations
and the four colormap passes are kernels of libsfpe.so on one stream; only the uint8 images (and, with
`pickle=True`, the three float64 images) come back.  Colormap bytes: parity unpinned (include/sfpe.h).
"""
import ctypes as C

import numpy as np

from .. import _lib, _session, colormap, synth
from ..common.birds_eye_view_generator import LIDAR_MAX_DENSITY_SQ_METER
from . import globals  # noqa: A004


class VelodyneMessageRenderer:
    """Constructor arguments as ROSBagExtractor's (extract_rosbag.py:27-45) where they concern the lidar branch."""

    def __init__(self, topdown_res=.2, topdown_max_range=120, cmap=None, pickle=False,
                 point_step=synth.PC2_POINT_STEP, offsets=synth.PC2_OFFSETS, dtype="f32"):
        if cmap is not None and cmap not in colormap.CMAPS:
            raise ValueError("colormap %r is not available (have: %s)" % (cmap, ", ".join(sorted(colormap.CMAPS))))
        self.cmap = cmap
        self.topdown_res = (topdown_res, topdown_res)          # :42
        self.topdown_max_range = topdown_max_range
        self.pickle = pickle
        self.point_step, self.offsets, self.dtype = int(point_step), tuple(offsets), dtype
        self.pr = _session.projector(bev_res=(float(topdown_res), float(topdown_res)), max_range=float(topdown_max_range))

    def _rgb(self, torch, img, cmap, vmin=None, vmax=None):
        """device float64 image -> device uint8 image (..., 3) through sfpe_colormap."""
        out = torch.empty(tuple(img.shape) + (3,), dtype=torch.uint8, device=img.device)
        auto = vmin is None
        rc = self.pr.lib.sfpe_colormap(
            self.pr._ctx, C.c_void_p(img.data_ptr()), _lib.SFPE_F64, img.numel(), colormap.CMAPS[cmap], int(auto),
            0.0 if auto else float(vmin), 0.0 if auto else float(vmax), C.c_void_p(out.data_ptr()),
            C.c_void_p(torch.cuda.current_stream(self.pr.device).cuda_stream))
        _lib.check(rc, self.pr._ctx)
        return out

    def __call__(self, data, n_points=None):
        """data: the message's `data` (bytes or uint8 array).  Returns (density_top_down, lidar_images) as the
        branch computes them (:169-178): density_top_down uint8 (nb, nb, 3); lidar_images['intensity' | 'distance' |
        'height'] uint8 (32,1801,3) -- float64 (32,1801) when cmap is None, as in the reference -- plus the three
        '*_float' float64 images when `pickle`.  Raises IndexError where the reference's projection would."""
        import torch
        pr = self.pr
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1)
        n = len(buf) // self.point_step if n_points is None else int(n_points)
        dev = torch.device("cuda", pr.device)
        d_blob = torch.from_numpy(np.ascontiguousarray(buf[:n * self.point_step])).to(dev)
        pts = pr.decode_pointcloud2(d_blob, n, self.point_step, self.offsets, self.dtype)
        out = pr.project(pts, n_frames=1, uniform_n=n, range_dtype="f64", range_layout="chw", bev_dtype="f64")
        rng, bev = out["range"][0], out["bev"][0]              # (3,32,1801) distance/height/intensity, (nb,nb)
        # generate_birds_eye_view always colormaps (birds_eye_view_generator.py:48-53); with cmap=None matplotlib
        # would fall back to its rcParams default, which this library does not carry: gray is used instead
        density = self._rgb(torch, bev, self.cmap or "gray", 0, LIDAR_MAX_DENSITY_SQ_METER * self.topdown_res[0])
        images = {}
        for kind, ch in (("intensity", 2), ("distance", 0), ("height", 1)):
            if self.cmap is None:
                images[kind] = rng[ch]
            elif kind == "height":
                images[kind] = self._rgb(torch, rng[ch], self.cmap, globals.LIDAR_MIN_HEIGHT, globals.LIDAR_MAX_HEIGHT)
            else:
                images[kind] = self._rgb(torch, rng[ch], self.cmap)
        status = int(out["status"][0])                          # synchronises
        if status:
            raise IndexError("index out of bounds: sweep has points outside the range image (status %d)" % status)
        lidar_images = {k: v.cpu().numpy() for k, v in images.items()}
        if self.pickle:                                         # :175-178
            for kind, ch in (("intensity", 2), ("distance", 0), ("height", 1)):
                lidar_images[kind + "_float"] = rng[ch].cpu().numpy()
        return density.cpu().numpy().astype(np.uint8), lidar_images
