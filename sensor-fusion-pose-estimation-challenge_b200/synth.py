"""Seeded synthetic Velodyne-shaped sweeps (SURVEY.md 8d "Synthetic inputs").

Host-side data generation only -- there is no dataset on the box.  A sweep is
`beams x n_az` returns in firing order (azimuth-major, beam-minor): each ray hits
either the ground plane z = -1.7 m or an "object" at U(2, 80) m, whichever is
nearer.  All coordinates are rounded to float32, which is what a PointCloud2 blob
carries (reference: team_sf_rosnode/scripts/lidar_predict.py:64-65 decodes float32
fields into a float64 array).  Elevations stay inside the reference's safe band
(process/globals.py:7 VFOV) so the reference neither wraps nor raises.
"""
import numpy as np

# name -> (beams, n_az, el_lo_deg, el_hi_deg)
SENSORS = {
    "hdl32": (32, 2048, -30.67, 10.67),      # 65 536 pts
    "hdl64": (64, 1875, -24.9, 2.0),         # 120 000 pts
    "vls128": (128, 1953, -25.0, 10.5),      # 249 984 pts
}


def make_sweep(sensor="hdl32", seed=0, dropout=0.0, shuffle=False, near_heavy=False,
               beams=None, n_az=None):
    """Returns a C-contiguous float32 (N, 4) array [x, y, z, intensity]."""
    b, na, el_lo, el_hi = SENSORS[sensor]
    if beams is not None:
        b = beams
    if n_az is not None:
        na = n_az
    rng = np.random.default_rng(seed)
    el = np.deg2rad(np.linspace(el_lo, el_hi, b))
    # keep the extreme beams strictly inside the band after float32 rounding
    el = np.clip(el, np.deg2rad(el_lo) + 1e-4, np.deg2rad(el_hi) - 1e-4)
    az = np.linspace(-np.pi, np.pi, na, endpoint=False)
    az = az + np.deg2rad(rng.uniform(-0.02, 0.02, size=na))
    az = np.clip(az, -np.pi + 1e-6, np.pi - 1e-6)
    A, E = np.meshgrid(az, el, indexing="ij")            # (n_az, beams): firing order
    A = A.ravel()
    E = E.ravel()
    n = A.size
    with np.errstate(divide="ignore"):
        ground = np.where(E < 0, 1.7 / np.sin(-E), np.inf)
    if near_heavy:                                        # C5: 60 % of object returns < 15 m
        near = rng.random(n) < 0.6
        obj = np.where(near, rng.uniform(2.0, 15.0, n), rng.uniform(15.0, 80.0, n))
    else:
        obj = rng.uniform(2.0, 80.0, n)
    has_obj = (rng.random(n) < 0.35) | (ground > 100.0)
    rho = np.where(has_obj, np.minimum(obj, ground), ground)
    ce = np.cos(E)
    pts = np.empty((n, 4), dtype=np.float32)
    pts[:, 0] = rho * ce * np.cos(A)
    pts[:, 1] = rho * ce * np.sin(A)
    pts[:, 2] = rho * np.sin(E)
    pts[:, 3] = rng.uniform(0.0, 255.0, n)
    if dropout > 0:
        pts = pts[rng.random(n) >= dropout]
    if shuffle:
        pts = pts[rng.permutation(len(pts))]
    return np.ascontiguousarray(pts)


def make_batch(n_sweeps, sensor="hdl32", seed0=0, **kw):
    """Concatenated points (sum N, 4) float32 and int64 frame offsets (n_sweeps + 1)."""
    sweeps = [make_sweep(sensor, seed0 + i, **kw) for i in range(n_sweeps)]
    offs = np.zeros(n_sweeps + 1, dtype=np.int64)
    offs[1:] = np.cumsum([len(s) for s in sweeps])
    return np.concatenate(sweeps, axis=0), offs


def as_reference_array(points_f32):
    """(N,5) float64 [x,y,z,intensity,ring] -- the layout the reference's callers hand to the
    projection (np.array(list(read_points(msg)))); ring is ignored by the hot path."""
    n = len(points_f32)
    out = np.zeros((n, 5), dtype=np.float64)
    out[:, :4] = points_f32
    return out


# Velodyne driver PointCloud2 layout: point_step 32 bytes,
# x,y,z float32 @0/4/8, intensity float32 @16, ring uint16 @20.
PC2_POINT_STEP = 32
PC2_OFFSETS = (0, 4, 8, 16)


def make_pointcloud2_blob(points_f32, beams=32):
    """uint8 array (N*32,) laid out like the `data` field of a /velodyne_points message."""
    n = len(points_f32)
    rec = np.zeros(n, dtype=np.dtype({"names": ["x", "y", "z", "intensity", "ring"],
                                      "formats": ["<f4", "<f4", "<f4", "<f4", "<u2"],
                                      "offsets": [0, 4, 8, 16, 20], "itemsize": PC2_POINT_STEP}))
    rec["x"], rec["y"], rec["z"], rec["intensity"] = points_f32[:, 0], points_f32[:, 1], points_f32[:, 2], points_f32[:, 3]
    rec["ring"] = np.arange(n) % beams
    return rec.view(np.uint8).reshape(-1)


def make_back_project_case(n_items=12, seed=0, height=32, width=1801):
    """Synthetic input of the back-projection (train/predict.py:230-293): per item a sparse (distance, height)
    range image pair plus a cluster centroid [column, row] and its box [left, top, right, bottom].  Covers a
    centroid on a valid pixel, on an invalid one (nearest valid pixel of the box wins, ties included), a box
    with nothing valid, a box whose left edge is 0 (the reference skips the search) and the (0,0) centroid.
    Returns float32 images (exactly representable in float64) and float64 centroids / boxes."""
    rng = np.random.default_rng(1000 + seed)
    dist = np.zeros((n_items, height, width), np.float32)
    hgt = np.full((n_items, height, width), -2.0, np.float32)
    cen = np.zeros((n_items, 2))
    box = np.zeros((n_items, 4))
    for i in range(n_items):
        fill = (0.01, 0.05, 0.3)[i % 3]
        hit = rng.random((height, width)) < fill
        dist[i][hit] = rng.uniform(1.0, 60.0, hit.sum()).astype(np.float32)
        hgt[i][hit] = rng.uniform(-2.6, 1.5, hit.sum()).astype(np.float32)     # some hits sit below the height floor
        cx, cy = rng.uniform(40, width - 40), rng.uniform(0, height - 0.01)
        half_w, half_h = rng.integers(1, 30), rng.integers(0, 6)
        l, r = int(cx) - half_w, int(cx) + half_w
        t, b = max(int(cy) - half_h, 0), min(int(cy) + half_h, height - 1)
        kind = i % 6
        if kind == 1:      # centroid on a valid pixel, fractional coordinates kept for theta
            dist[i, int(cy), int(cx)] = 12.5
            hgt[i, int(cy), int(cx)] = 0.25
        elif kind == 2:    # nothing valid inside the box
            dist[i, t:b + 1, l:r + 1] = 0
        elif kind == 3:    # left edge 0: no search, centroid used as it is
            l = 0
        elif kind == 4:    # symmetric pair at equal distance: the first in row-major order wins
            dist[i, t:b + 1, l:r + 1] = 0
            for dx in (-1, 1):
                dist[i, int(cy), int(cx) + dx] = 7.0 + dx
                hgt[i, int(cy), int(cx) + dx] = 0.5
        elif kind == 5 and i >= 6:
            cx, cy = 0.0, 0.0
            l, t, r, b = 0, 0, 5, 3
        cen[i] = (cx, cy)
        box[i] = (l, t, r, b)
    return dist, hgt, cen, box


def make_bbox_decode_case(seed=0, height=32, width=1801, noise=0.35, strays=6):
    """Synthetic input of the per-pixel box decode (train/predict.py:94-199, find_bbox_3d): an FCN output
    `y_pred` (height*width, 26) float32 -- [p(background), p(car), 8 corners x 3 in the pixel's rotated frame] --
    with one blob of positive pixels on a car, a few stray positives elsewhere (they widen the reference's
    row / column membership test), the matching float32 distance / height images, the blob's 2-D box
    ((left, top), (right, bottom)) and a 3-D centroid.  Regression noise makes the neighbour counts differ."""
    rng = np.random.default_rng(2000 + seed)
    res_v, res_h = 1.33 * np.pi / 180, 0.2 * np.pi / 180
    x_min, y_min = -900.0, -30.67 / 1.33
    cx0 = int(rng.integers(150, width - 150))
    cy0 = int(rng.integers(8, height - 8))
    d0 = float(rng.uniform(8.0, 30.0))
    half_w, half_h = int(rng.integers(6, 16)), int(rng.integers(2, 4))
    theta0 = (cx0 + x_min) * res_h
    centre = np.array([(d0 + 1.0) * np.cos(theta0), -(d0 + 1.0) * np.sin(theta0), -0.6])
    yaw = float(rng.uniform(-np.pi, np.pi))
    l, w, h = 4.5, 1.8, 1.5
    corners = np.zeros((8, 3))
    for k in range(8):   # bit 2: front/back, bit 1: left/right, bit 0: top/bottom (predict.py:176-190)
        sx, sy, sz = (1 if not k & 4 else -1), (1 if not k & 2 else -1), (1 if not k & 1 else -1)
        bx, by = sx * l / 2, sy * w / 2
        corners[k] = centre + (bx * np.cos(yaw) - by * np.sin(yaw), bx * np.sin(yaw) + by * np.cos(yaw), sz * h / 2)
    dist = rng.uniform(2.0, 60.0, (height, width)).astype(np.float32)
    hgt = rng.uniform(-1.9, 1.0, (height, width)).astype(np.float32)
    y = np.zeros((height, width, 26), np.float32)
    y[..., 1] = rng.uniform(0.0, 0.45, (height, width))
    positives = [(cy, cx) for cy in range(cy0 - half_h, cy0 + half_h + 1) for cx in range(cx0 - half_w, cx0 + half_w + 1)
                 if rng.random() < 0.85]
    for _ in range(strays):
        positives.append((int(rng.integers(0, height)), int(rng.integers(0, width))))
        positives.append((int(rng.integers(max(cy0 - 4, 0), min(cy0 + 4, height))), int(rng.integers(max(cx0 - 90, 0), min(cx0 + 90, width)))))
    for cy, cx in positives:
        y[cy, cx, 1] = rng.uniform(0.5, 1.0)
        dist[cy, cx] = d0 + rng.uniform(-0.5, 0.5)
        hgt[cy, cx] = rng.uniform(-1.2, 0.2)
    # regression targets for every pixel of the neighbourhood (the reference decodes non-positive pixels too when
    # their row and column each hold a positive somewhere): corners in the pixel's frame, plus noise
    for cy in range(max(cy0 - half_h - 6, 0), min(cy0 + half_h + 7, height)):
        for cx in range(max(cx0 - half_w - 110, 0), min(cx0 + half_w + 111, width)):
            th, ph = (cx + x_min) * res_h, (cy + y_min) * res_v
            p = np.array([dist[cy, cx] * np.cos(th), -dist[cy, cx] * np.sin(th), hgt[cy, cx]], dtype=np.float64)
            rz = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1]])
            ry = np.array([[np.cos(ph), 0, np.sin(ph)], [0, 1, 0], [-np.sin(ph), 0, np.cos(ph)]])
            rel = (corners + rng.normal(0, noise, (8, 3)) - p) @ (rz @ ry)      # rot^T (c - p), row-vector form
            y[cy, cx, 2:] = rel.reshape(24)
    bbox_2d = ((cx0 - half_w, cy0 - half_h), (cx0 + half_w, cy0 + half_h))
    centroid_3d = centre + rng.normal(0, 0.3, 3)
    return dist, hgt, y.reshape(height * width, 26), bbox_2d, centroid_3d


def make_obstacle_case(seed=0, height=32, width=1801, blobs=4, speckle=0.002):
    """Synthetic 2-class FCN output (height*width, 2) float32 for find_obstacle (train/predict.py:33-81): a few
    ragged blobs of positive pixels (some touching the image border, some joined by thin bridges, some of equal
    size), plus isolated speckle that stays under the heat threshold.  seed % 4 == 3 gives a frame whose largest
    cluster is too small (the reference returns (None, None, None))."""
    rng = np.random.default_rng(3000 + seed)
    p = rng.uniform(0.0, 0.49, (height, width)).astype(np.float32)
    small = seed % 4 == 3
    for b in range(blobs):
        bw, bh = (int(rng.integers(3, 8)), int(rng.integers(2, 5))) if small else (int(rng.integers(8, 40)), int(rng.integers(4, 14)))
        x0 = int(rng.integers(0, width - bw)) if b else 0                      # the first blob touches the left border
        y0 = int(rng.integers(0, height - bh)) if b != 1 else height - bh      # the second the bottom border
        keep = rng.random((bh, bw)) < 0.8
        p[y0:y0 + bh, x0:x0 + bw][keep] = rng.uniform(0.5, 1.0, keep.sum())
        if b == 2 and not small:                                               # a copy of equal size further right: tie on area
            x1 = min(x0 + bw + 60, width - bw)
            p[y0:y0 + bh, x1:x1 + bw] = p[y0:y0 + bh, x0:x0 + bw]
    spk = rng.random((height, width)) < speckle
    p[spk] = 0.9
    y = np.zeros((height * width, 2), np.float32)
    y[:, 1] = p.ravel()
    y[:, 0] = 1 - y[:, 1]
    return y


def make_detection_case(seed=0):
    """A frame for the whole post-network chain (find_obstacle -> back_project_2D_2_3D -> find_bbox_3d): the box-decode
    case with its blob filled densely enough to form one cluster above the area threshold, and the range scan
    hollowed out under the cluster's centre so that the back-projection has to search its box."""
    dist, hgt, y, bb, _ = make_bbox_decode_case(seed=seed, strays=0)
    rng = np.random.default_rng(4000 + seed)
    H, W = dist.shape
    (l, t), (r, b) = bb
    l, r, t, b = l - 6, r + 6, max(t - 2, 2), min(b + 2, H - 1)
    y3 = y.reshape(H, W, 26)
    y3[t:b + 1, l:r + 1, 1] = rng.uniform(0.55, 1.0, (b - t + 1, r - l + 1))
    cy, cx = (t + b) // 2, (l + r) // 2
    dist[cy - 1:cy + 2, cx - 2:cx + 3] = 0       # no return under the centroid (seed % 3 == 2: nowhere in the box)
    if seed % 3 == 2:
        dist[max(t - 3, 0):b + 4, l - 3:r + 4] = 0
    return dist, hgt, y3.reshape(H * W, 26)
