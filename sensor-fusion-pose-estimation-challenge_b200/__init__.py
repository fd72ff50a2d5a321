"""B200-native lidar rasterisation: BEV density + 360-degree range scan.

Drop-in for the projection hot path of J-Rojas/Sensor-Fusion-Pose-Estimation-Challenge
(modules/lidar/process/extract_rosbag_lidar.py:18-86,
modules/lidar/common/birds_eye_view_generator.py:31-54).  Everything numeric runs in
hand-written sm_100a CUDA kernels inside libsfpe.so (C ABI: include/sfpe.h); this
package is the Python host side that mirrors the reference's entry points.

Importing the package loads libsfpe.so and fails if it is missing -- there is no CPU
path.  (Creating a projector additionally needs a CUDA device.)
"""
from . import _lib

_lib.load()  # fail loudly if the CUDA library is not built

from .projector import LidarProjector  # noqa: E402
from .host_batch import HostBatchPipeline  # noqa: E402
from .streaming import StreamingFrontView  # noqa: E402
from . import synth  # noqa: E402,F401
from .sharding import shard_frames, shard_points  # noqa: E402,F401
from .process import globals as grid_globals  # noqa: E402,F401
from .process.extract_rosbag_lidar import generate_lidar_2d_front_view, lidar_2d_front_view  # noqa: E402,F401
from .common.birds_eye_view_generator import birds_eye_view_density, generate_birds_eye_view  # noqa: E402,F401

__all__ = ["LidarProjector", "HostBatchPipeline", "StreamingFrontView", "generate_lidar_2d_front_view", "lidar_2d_front_view",
           "generate_birds_eye_view", "birds_eye_view_density", "synth", "grid_globals", "shard_frames", "shard_points"]
