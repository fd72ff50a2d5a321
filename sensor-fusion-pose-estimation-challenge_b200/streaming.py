"""Streaming entry: one PointCloud2 message in, the FCN input tensor out.

Mirrors what the ROS callback of the reference does per message
(team_sf_rosnode/scripts/lidar_predict.py:60-70 -> lidar/pipeline.py:10-11 ->
train/predict.py:335-343): decode the message, build the three range images and stack them as
(32,1801,3) = (distance, height, intensity).  Here the `data` blob goes to the device as it is
(pinned staging buffer, one H2D copy), is unpacked by sfpe_decode_pointcloud2, rasterised by the
range-scan kernels and the (32,1801,3) float32 tensor comes back with one D2H copy.  rospy is not
needed: the caller passes the message's `data`, `point_step` and field offsets.
"""
import numpy as np

from .projector import LidarProjector
from . import synth


class StreamingFrontView:
    def __init__(self, device=0, max_points=1 << 18, point_step=synth.PC2_POINT_STEP, offsets=synth.PC2_OFFSETS,
                 dtype="f32"):
        import torch
        self.torch = torch
        self.pr = LidarProjector(device=device, enable_bev=False)
        self.point_step, self.offsets, self.dtype = int(point_step), tuple(offsets), dtype
        self.max_points = int(max_points)
        dev = torch.device("cuda", device)
        self.stream = torch.cuda.Stream(dev)
        self.h_blob = torch.empty(self.max_points * self.point_step, dtype=torch.uint8).pin_memory()
        self.d_blob = torch.empty_like(self.h_blob, device=dev)
        self.d_pts = torch.empty((self.max_points, 4), dtype=torch.float32, device=dev)
        self.d_out = self.pr.alloc_outputs(1, range_layout="hwc")
        self.h_out = torch.empty((self.pr.H, self.pr.W, 3), dtype=torch.float32).pin_memory()
        self.h_status = torch.zeros(2, dtype=torch.int32).pin_memory()

    def __call__(self, data, n_points=None):
        """data: bytes / uint8 array of the message's `data`.  Returns the (32,1801,3) float32
        numpy view of the pinned result buffer (valid until the next call).  Raises IndexError
        where the reference's projection would (train/predict.py:336 -> extract_rosbag_lidar.py:71)."""
        torch = self.torch
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.view(np.uint8).reshape(-1)
        n = len(buf) // self.point_step if n_points is None else int(n_points)
        if n > self.max_points:
            raise ValueError("message has more points than max_points")
        nb = n * self.point_step
        self.h_blob[:nb].numpy()[:] = buf[:nb]
        with torch.cuda.stream(self.stream):
            self.d_blob[:nb].copy_(self.h_blob[:nb], non_blocking=True)
            self.pr.decode_pointcloud2(self.d_blob, n, self.point_step, self.offsets, self.dtype, out=self.d_pts,
                                       stream=self.stream.cuda_stream)
            self.pr.project(self.d_pts, n_frames=1, uniform_n=n, out=self.d_out, stream=self.stream.cuda_stream)
            self.h_out.copy_(self.d_out["range"][0], non_blocking=True)
            self.h_status[0:1].copy_(self.d_out["status"], non_blocking=True)
        self.stream.synchronize()
        if int(self.h_status[0]):
            raise IndexError("index out of bounds: sweep has points outside the range image (status %d)" % int(self.h_status[0]))
        return self.h_out.numpy()

    def close(self):
        self.pr.close()
