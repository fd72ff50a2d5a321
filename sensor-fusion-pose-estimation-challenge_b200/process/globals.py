"""Grid / FOV constants under the reference's names (modules/lidar/process/globals.py:3-13).

The values are read back from the library's default parameter block and its derived
grid (sfpe_default_params / sfpe_grid_dims), so Python callers and the CUDA kernels can
never disagree about the grid.
"""
import numpy as np

from .. import _lib

_p = _lib.default_params()
_H, _W, _, _ = _lib.grid_dims(_p)

LIDAR_MIN_HEIGHT = int(_p.height_bg)
LIDAR_MAX_HEIGHT = -LIDAR_MIN_HEIGHT

RES = (_p.res_v_deg, _p.res_h_deg)            # (vertical, horizontal) degrees per pixel
VFOV = (_p.vfov_lo_deg, _p.vfov_hi_deg)
RES_RAD = np.array(RES) * (np.pi / 180)
X_MIN = -360.0 / RES[1] / 2
Y_MIN = VFOV[0] / RES[0]
X_MAX = _W - 1
Y_MAX = _H - 1
