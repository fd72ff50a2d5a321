"""Consumers of the range scan, kept on the device.

    back_project_2D_2_3D      train/predict.py:230-293      -> sfpe_back_project
    InvariantMask.update      process/create_lidar_invariant_mask.py:44-62 -> sfpe_invariant_mask_update
    locate_obstacles          train/predict.py:347-362 (the part of predict_point_cloud after model.predict)
    find_obstacle             train/predict.py:33-81 -> sfpe_find_obstacle
    find_bbox_3d              train/predict.py:94-199 (+ is_far :83-92) -> sfpe_find_bbox_3d
    roll_width                train/loader.py:47-50 (np.roll(image, rotate_by, axis=1)) -> sfpe_roll_width

The torch-tensor entries work on device-resident range scans as the projector writes them
(`(B, 32, 1801, 3)` float32, channels distance / height / intensity); the numpy entries keep the
reference's signatures (host arrays in, host arrays out, centroids / boxes updated in place).
There is no CPU path: every function goes through libsfpe.so.
"""
import ctypes as C

import numpy as np

from . import _lib, _session
from .process import globals as G


def _torch():
    import torch
    return torch


def _dtype_code(t):
    torch = _torch()
    if t.dtype == torch.float32:
        return _lib.SFPE_F32
    if t.dtype == torch.float64:
        return _lib.SFPE_F64
    raise TypeError("range images must be float32 or float64, got %s" % t.dtype)


def _stream_ptr(stream):
    torch = _torch()
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def _release_on(stream, *tensors):
    """Scratch tensors die when the caller returns.  torch's caching allocator hands a block back only to work queued
    later on the stream it was allocated on, so nothing is needed for the current stream; a kernel launched on
    another stream has to be recorded on the block."""
    if stream is not None:
        for t in tensors:
            t.record_stream(stream)


def _projector(device):
    pr = _session.projector(enable_bev=False)
    if device is not None and device.index is not None and device.index != pr.device:
        raise ValueError("tensors live on cuda:%d but the session projector is on cuda:%d (set SFPE_DEVICE)"
                         % (device.index, pr.device))
    return pr


class InvariantMask:
    """Static-pixel mask over a stream of range scans: a pixel stays 1 while it equals the reference frame in
    every frame seen (create_lidar_invariant_mask.py:59-62).  `reference`: device tensor of one frame, any shape;
    `update(frames)`: device tensor of shape (n,) + reference.shape."""

    def __init__(self, reference, stream=None):
        torch = _torch()
        if not reference.is_cuda:
            raise ValueError("reference must be a CUDA tensor")
        self.reference = reference.contiguous()
        self.mask = torch.ones(self.reference.shape, dtype=torch.uint8, device=reference.device)
        self._pr = _projector(reference.device)
        self._stream = stream

    def update(self, frames):
        ref = self.reference
        if frames.shape[1:] != ref.shape or frames.dtype != ref.dtype or frames.device != ref.device:
            raise ValueError("frames must be (n,)+%s %s on %s" % (tuple(ref.shape), ref.dtype, ref.device))
        frames = frames.contiguous()
        n_pix = ref.numel()
        rc = self._pr.lib.sfpe_invariant_mask_update(
            self._pr._ctx, C.c_void_p(ref.data_ptr()), C.c_void_p(frames.data_ptr()), _dtype_code(ref),
            frames.shape[0], n_pix, n_pix, 1, C.c_void_p(self.mask.data_ptr()), _stream_ptr(self._stream))
        _lib.check(rc, self._pr._ctx)
        return self.mask


def roll_width(images, rotate_by, out=None, stream=None):
    """np.roll(image, rotate_by[f], axis=1) for every (height, width, channels) item of a device batch
    (train/loader.py:47-50).  images: (B, H, W, C) or (B, H, W) tensor of a 4-byte-multiple pixel size;
    rotate_by: int, sequence of B ints, or int32 device tensor.  Returns a new tensor (or `out`)."""
    torch = _torch()
    if not images.is_cuda or images.dim() not in (3, 4):
        raise ValueError("images must be a CUDA tensor of shape (B,H,W) or (B,H,W,C)")
    images = images.contiguous()
    B, H, W = images.shape[:3]
    pix_bytes = images.element_size() * (images.shape[3] if images.dim() == 4 else 1)
    if torch.is_tensor(rotate_by):
        shift = rotate_by.to(device=images.device, dtype=torch.int32).contiguous()
    else:
        shift = torch.as_tensor(np.broadcast_to(np.asarray(rotate_by, dtype=np.int64) % max(W, 1), (B,)).astype(np.int32),
                                device=images.device)
    if shift.numel() != B:
        raise ValueError("need one shift per batch item")
    if out is None:
        out = torch.empty_like(images)
    elif out.shape != images.shape or out.dtype != images.dtype or not out.is_contiguous() or out.data_ptr() == images.data_ptr():
        raise ValueError("out must be a distinct contiguous tensor shaped like images")
    pr = _projector(images.device)
    rc = pr.lib.sfpe_roll_width(pr._ctx, C.c_void_p(images.data_ptr()), C.c_void_p(out.data_ptr()), B, H, W, pix_bytes,
                                C.c_void_p(shift.data_ptr()), _stream_ptr(stream))
    _lib.check(rc, pr._ctx)
    return out


def back_project_device(centroids, bboxes, range_scan=None, distance=None, height=None, stream=None):
    """Device form of back_project_2D_2_3D.  centroids (B,2) / bboxes (B,4): float64 CUDA tensors, updated in
    place.  Either `range_scan` (B,H,W,3) as the projector writes it (distance, height, intensity) or separate
    `distance` / `height` (B,H,W) tensors.  Returns (xyz (B,4) float64, status (B,) int32) device tensors;
    status 3 marks items whose centroid or box lies outside the image."""
    torch = _torch()
    if range_scan is not None:
        rs = range_scan.contiguous()
        B, H, W, Cn = rs.shape
        d_ptr, h_ptr = rs.data_ptr(), rs.data_ptr() + rs.element_size()
        frame_stride, pix_stride, code, dev = H * W * Cn, Cn, _dtype_code(rs), rs.device
        keep = (rs,)
    else:
        distance, height = distance.contiguous(), height.contiguous()
        if distance.shape != height.shape or distance.dtype != height.dtype:
            raise ValueError("distance and height must match")
        B, H, W = distance.shape
        d_ptr, h_ptr = distance.data_ptr(), height.data_ptr()
        frame_stride, pix_stride, code, dev = H * W, 1, _dtype_code(distance), distance.device
        keep = (distance, height)
    for t, cols in ((centroids, 2), (bboxes, 4)):
        if t.dtype != torch.float64 or tuple(t.shape) != (B, cols) or not t.is_contiguous() or t.device != dev:
            raise ValueError("centroids (B,2) and bboxes (B,4) must be contiguous float64 tensors on %s" % dev)
    xyz = torch.zeros((B, 4), dtype=torch.float64, device=dev)    # status-3 items keep zeros (the kernel writes nothing)
    status = torch.empty((B,), dtype=torch.int32, device=dev)
    pr = _projector(dev)
    rc = pr.lib.sfpe_back_project(
        pr._ctx, C.c_void_p(centroids.data_ptr()), C.c_void_p(bboxes.data_ptr()), B, C.c_void_p(d_ptr), C.c_void_p(h_ptr),
        code, frame_stride, pix_stride, H, W, float(G.LIDAR_MIN_HEIGHT), float(G.X_MIN), float(G.RES_RAD[1]),
        C.c_void_p(xyz.data_ptr()), C.c_void_p(status.data_ptr()), _stream_ptr(stream))
    _lib.check(rc, pr._ctx)
    del keep
    return xyz, status


def back_project_2D_2_3D(centroids, bboxes, distance_data, height_data):
    """Reference signature (train/predict.py:230): numpy arrays in, `xyz_coor` (B,4) out; `centroids` and
    `bboxes` are updated in place.  An item whose centroid or box lies outside the image raises IndexError
    (the reference indexes out of range, or silently wraps, there)."""
    torch = _torch()
    pr = _projector(None)
    dev = torch.device("cuda", pr.device)
    d = np.ascontiguousarray(distance_data)
    h = np.ascontiguousarray(height_data)
    if d.dtype not in (np.float32, np.float64):
        d = d.astype(np.float64)
    h = h.astype(d.dtype, copy=False)
    cen = torch.as_tensor(np.ascontiguousarray(centroids, dtype=np.float64)).to(dev)
    box = torch.as_tensor(np.ascontiguousarray(bboxes, dtype=np.float64)).to(dev)
    xyz, status = back_project_device(cen, box, distance=torch.as_tensor(d).to(dev), height=torch.as_tensor(h).to(dev))
    status = status.cpu().numpy()
    if (status == 3).any():
        raise IndexError("centroid or bounding box outside the image for item(s) %s" % np.nonzero(status == 3)[0].tolist())
    centroids[...] = cen.cpu().numpy()
    bboxes[...] = box.cpu().numpy()
    return xyz.cpu().numpy()


def find_bbox_3d_device(y_pred, bbox_2d, centroid_3d, range_scan=None, distance=None, height=None, cap=None, stream=None,
                        posbits=None):
    """Device form of find_bbox_3d for a batch.  y_pred (B, H*W, 26) float32/float64 CUDA tensor; bbox_2d (B,4)
    int32 [left, top, right, bottom]; centroid_3d (B,3) float64; images as in back_project_device.  Returns
    (pred (B,7) float64, box (B,8,3) float64, n_boxes (B,) int32) device tensors; n_boxes 0 = the reference's
    `(zeros(7), None)`."""
    torch = _torch()
    if range_scan is not None:
        rs = range_scan.contiguous()
        B, H, W, Cn = rs.shape
        d_ptr, h_ptr = rs.data_ptr(), rs.data_ptr() + rs.element_size()
        frame_stride, pix_stride, code, dev = H * W * Cn, Cn, _dtype_code(rs), rs.device
    else:
        distance, height = distance.contiguous(), height.contiguous()
        B, H, W = distance.shape
        d_ptr, h_ptr = distance.data_ptr(), height.data_ptr()
        frame_stride, pix_stride, code, dev = H * W, 1, _dtype_code(distance), distance.device
    y_pred = y_pred.contiguous()
    if tuple(y_pred.shape) != (B, H * W, 26) or y_pred.device != dev:
        raise ValueError("y_pred must be (B, H*W, 26) on %s" % dev)
    bbox_2d = bbox_2d.to(device=dev, dtype=torch.int32).contiguous()
    centroid_3d = centroid_3d.to(device=dev, dtype=torch.float64).contiguous()
    if tuple(bbox_2d.shape) != (B, 4) or tuple(centroid_3d.shape) != (B, 3):
        raise ValueError("bbox_2d must be (B,4) and centroid_3d (B,3)")
    if cap is None:     # the largest window: nothing can overflow
        bb = bbox_2d.cpu().numpy().astype(np.int64)
        wx = np.clip(np.minimum(bb[:, 2] + 100, W) - np.maximum(bb[:, 0] - 100, 0), 0, None)
        wy = np.clip(np.minimum(bb[:, 3] + 2, H) - np.maximum(bb[:, 1] - 2, 0), 0, None)
        cap = int((wx * wy).max()) if B else 0
    cap = max(int(cap), 1)
    boxes = torch.empty((B, cap, 24), dtype=torch.float64, device=dev)
    counts = torch.empty((B, cap), dtype=torch.int32, device=dev)
    pred = torch.empty((B, 7), dtype=torch.float64, device=dev)
    box = torch.empty((B, 8, 3), dtype=torch.float64, device=dev)
    n_boxes = torch.empty((B,), dtype=torch.int32, device=dev)
    pr = _projector(dev)
    rc = pr.lib.sfpe_find_bbox_3d(
        pr._ctx, C.c_void_p(d_ptr), C.c_void_p(h_ptr), code, frame_stride, pix_stride, C.c_void_p(y_pred.data_ptr()),
        _dtype_code(y_pred), B, H, W, C.c_void_p(bbox_2d.data_ptr()), C.c_void_p(centroid_3d.data_ptr()),
        float(G.X_MIN), float(G.Y_MIN), float(G.RES_RAD[1]), float(G.RES_RAD[0]), C.c_void_p(boxes.data_ptr()),
        C.c_void_p(counts.data_ptr()), cap, C.c_void_p(pred.data_ptr()), C.c_void_p(box.data_ptr()),
        C.c_void_p(n_boxes.data_ptr()), C.c_void_p(posbits.data_ptr()) if posbits is not None else None,
        _stream_ptr(stream))
    _lib.check(rc, pr._ctx)
    _release_on(stream, boxes, counts)
    return pred, box, n_boxes


def find_bbox_3d(distance_img, height_img, y_pred, bbox_2d, centroid_3d):
    """Reference signature (train/predict.py:94): one frame, numpy in, `(pred[7], candidate_bbox (8,3) or None)` out."""
    torch = _torch()
    pr = _projector(None)
    dev = torch.device("cuda", pr.device)
    d = np.ascontiguousarray(distance_img)
    if d.dtype not in (np.float32, np.float64):
        d = d.astype(np.float64)
    h = np.ascontiguousarray(height_img).astype(d.dtype, copy=False)
    H, W = d.shape
    y = np.ascontiguousarray(y_pred)
    if y.dtype not in (np.float32, np.float64):
        y = y.astype(np.float64)
    y = y.reshape(1, H * W, 26)
    (ulx, uly), (lrx, lry) = bbox_2d
    pred, box, n = find_bbox_3d_device(
        torch.as_tensor(y).to(dev), torch.tensor([[ulx, uly, lrx, lry]], dtype=torch.int32),
        torch.tensor([[float(v) for v in centroid_3d[:3]]], dtype=torch.float64),
        distance=torch.as_tensor(d[None]).to(dev), height=torch.as_tensor(h[None]).to(dev))
    if int(n[0]) == 0:
        return np.zeros(7), None
    return pred[0].cpu().numpy(), box[0].cpu().numpy()


def find_obstacle_device(y_pred, height, width, stream=None, posbits=None):
    """Device form of find_obstacle for a batch: y_pred (B, height*width, n_channels) float32/float64 CUDA tensor.
    Returns an int32 device tensor (B, 8): found, centroid x, y, box left, top, right, bottom, area."""
    torch = _torch()
    y_pred = y_pred.contiguous()
    if y_pred.dim() != 3 or y_pred.shape[1] != height * width or y_pred.shape[2] < 2 or not y_pred.is_cuda:
        raise ValueError("y_pred must be a CUDA tensor (B, height*width, n_channels >= 2)")
    B = y_pred.shape[0]
    scratch = torch.empty((B, 5, height * width), dtype=torch.int32, device=y_pred.device)
    out = torch.empty((B, 8), dtype=torch.int32, device=y_pred.device)
    pr = _projector(y_pred.device)
    rc = pr.lib.sfpe_find_obstacle(pr._ctx, C.c_void_p(y_pred.data_ptr()), _dtype_code(y_pred), y_pred.shape[2], B, height,
                                   width, C.c_void_p(scratch.data_ptr()), C.c_void_p(out.data_ptr()),
                                   C.c_void_p(posbits.data_ptr()) if posbits is not None else None, _stream_ptr(stream))
    _lib.check(rc, pr._ctx)
    _release_on(stream, scratch)
    return out


def find_obstacle(y_pred, input_shape):
    """Reference signature (train/predict.py:33): y_pred (height*width, n_classes) numpy, input_shape = (height,
    width, ...).  Returns ((centroid_x, centroid_y), ((left, top), (right, bottom)), area) or (None, None, None)."""
    torch = _torch()
    pr = _projector(None)
    y = np.ascontiguousarray(y_pred)
    if y.dtype not in (np.float32, np.float64):
        y = y.astype(np.float64)
    H, W = int(input_shape[0]), int(input_shape[1])
    o = find_obstacle_device(torch.as_tensor(y.reshape(1, H * W, -1)).to(torch.device("cuda", pr.device)), H, W)[0].cpu().numpy()
    if not o[0]:
        return None, None, None
    return (int(o[1]), int(o[2])), ((int(o[3]), int(o[4])), (int(o[5]), int(o[6]))), int(o[7])


def locate_obstacles(prediction, range_scan, stream=None):
    """Everything predict_point_cloud does after `model.predict` (train/predict.py:347-362, use_regression=True),
    for a device-resident batch and without leaving the device: prediction (B, H*W, 26) as the FCN returns it,
    range_scan (B, H, W, 3) as the projector wrote it.  find_obstacle -> back_project_2D_2_3D -> find_bbox_3d.
    Returns (pred (B,7) float64: x, y, z, yaw, length, width, height; found (B,) bool).  Frames without a cluster,
    without a valid pixel in the cluster's box, or without a corner set near the centroid give zeros / False
    (the reference returns [0,0,0,0] or zeros(7) there)."""
    torch = _torch()
    B, H, W, _ = range_scan.shape
    posbits = torch.empty((B, (H * W + 31) // 32), dtype=torch.int32, device=range_scan.device)   # one pass over the tensor
    obs = find_obstacle_device(prediction, H, W, stream=stream, posbits=posbits)
    centroids = obs[:, 1:3].to(torch.float64).contiguous()
    bboxes = obs[:, 3:7].to(torch.float64).contiguous()
    xyz, bp_status = back_project_device(centroids, bboxes, range_scan=range_scan, stream=stream)
    pred, _, n_boxes = find_bbox_3d_device(prediction, obs[:, 3:7].contiguous(), xyz[:, :3].contiguous(),
                                           range_scan=range_scan, stream=stream, posbits=posbits)
    _release_on(stream, posbits)
    # status 3: the cluster's box or centroid lies outside the image (the reference would index out of range)
    found = (obs[:, 0] != 0) & (bp_status != 3) & ~((xyz[:, 0] == 0) & (xyz[:, 1] == 0)) & (n_boxes > 0)      # :349, :358, :137
    return torch.where(found[:, None], pred, torch.zeros_like(pred)), found


def predict_point_cloud_tail(prediction, points_2d, use_regression=True):
    """One frame, reference types: `prediction` = model.predict(...)[0] (H*W, 26), `points_2d` = the dict returned by
    generate_lidar_2d_front_view.  Returns what predict_point_cloud returns (train/predict.py:347-362)."""
    d = np.asarray(points_2d["distance_float"])
    h = np.asarray(points_2d["height_float"])
    centroid_2d, bbox_2d, _ = find_obstacle(prediction, d.shape)
    pred = np.array([0, 0, 0, 0])
    if centroid_2d is not None:
        centroids = np.array(centroid_2d).reshape(1, 2)
        bboxes = np.array(bbox_2d).reshape(1, 4)
        centroid_3d = back_project_2D_2_3D(centroids, bboxes, d[None], h[None])[0]
        if use_regression:
            if not (centroid_3d[0] == 0. and centroid_3d[1] == 0.):
                pred, _ = find_bbox_3d(d, h, prediction, bbox_2d, centroid_3d)
        else:
            pred = centroid_3d[0]
    return pred
