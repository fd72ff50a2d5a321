"""numpy restatement (fp64) of the reference's lidar rasterisation -- TEST INFRASTRUCTURE.

See oracle/__init__.py for who may import this.  Every function cites the reference
lines it restates (paths relative to /root/reference/modules/lidar).  The
restatement is structured differently from the reference on purpose: cells and the
winner are computed ONCE and made explicit (winner index map, tie mask, counts), so
parity tests can compare indices / winners / counts bit-exactly and not only the
rendered images.

Tie rule (the reference leaves it unspecified: unstable `argsort`,
process/extract_rosbag_lidar.py:65): within a cell the point with the smallest fp64
`l2_norm` wins; among points whose fp64 `l2_norm` is bit-identical the LOWEST
original index wins (what a stable sort would give the reference).  `tie_mask`
marks the cells where the reference's own choice is unspecified.
"""
import numpy as np

# ---- process/globals.py:3-13 ------------------------------------------------------
LIDAR_MAX_HEIGHT = 2
LIDAR_MIN_HEIGHT = -2
RES = (1.33, 0.2)                      # (vertical, horizontal) degrees
VFOV = (-30.67, 10.67)
RES_RAD = np.array(RES) * (np.pi / 180)
X_MIN = -360.0 / RES[1] / 2            # -900.0
Y_MIN = VFOV[0] / RES[0]               # -23.0601...
X_MAX = int(360.0 / RES[1])            # 1800
Y_MAX = int(abs(VFOV[0] - VFOV[1]) / RES[0])   # 31
H = Y_MAX + 1
W = X_MAX + 1

ERR_ROW_OOB = 1      # row outside [-H, H)   -> numpy IndexError in the reference
ERR_COL_OOB = 2      # col outside [-W, W)
ERR_NAN = 4          # NaN cell coordinate (NaN input or rho == 0)


class ReferenceIndexError(IndexError):
    """What the reference raises (numpy fancy-assignment IndexError) for the whole sweep."""


def front_view_cells(points):
    """Per-point cell coordinates, extract_rosbag_lidar.py:26-39,60-61.

    Returns (row, col, rho, err) where row/col are the PRE-flip integer indices after
    C truncation and numpy negative-index wrap, rho the fp64 l2 norm, err a per-point
    bitmask (non-zero = the reference would raise IndexError for the sweep)."""
    p = np.asarray(points, dtype=np.float64)
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    with np.errstate(all="ignore"):
        rho = np.sqrt(x * x + y * y + z * z)                       # :33
        xi = np.arctan2(-y, x) / RES_RAD[1] - X_MIN                # :34,38
        yi = np.arcsin(z / rho) / RES_RAD[0] - Y_MIN               # :35,39
    err = np.zeros(len(p), dtype=np.int32)
    nan = np.isnan(xi) | np.isnan(yi)
    err[nan] |= ERR_NAN
    xt = np.where(nan, 0.0, np.trunc(xi))                          # astype(int): toward zero :60-61
    yt = np.where(nan, 0.0, np.trunc(yi))
    col = xt.astype(np.int64)
    row = yt.astype(np.int64)
    err[(row < -H) | (row >= H)] |= ERR_ROW_OOB
    err[(col < -W) | (col >= W)] |= ERR_COL_OOB
    row = np.where(row < 0, row + H, row)                          # numpy negative index wrap
    col = np.where(col < 0, col + W, col)
    return row, col, rho, err


def front_view(points, strict=True):
    """generate_lidar_2d_front_view(points, cmap=None), extract_rosbag_lidar.py:18-86.

    strict=True  : reference semantics, raises ReferenceIndexError if any point would.
    strict=False : offending points are dropped (counted in `dropped`).
    Returns dict with float64 (32,1801) images `distance`, `height`, `intensity`
    (already flipped, :74), `winner` (int64, -1 = background), `tie_mask`, `dropped`,
    `err` (OR of the per-point masks)."""
    p = np.asarray(points, dtype=np.float64)
    n = len(p)
    row, col, rho, err = front_view_cells(p)
    bad = err != 0
    if strict and bad.any():
        raise ReferenceIndexError("reference would raise IndexError (err mask %d)" % int(np.bitwise_or.reduce(err)))
    idx = np.arange(n)[~bad]
    cell = (row * W + col)[~bad]
    r = rho[~bad]
    # nearest wins; equal rho -> lowest index.  Ascending by (rho, idx), assign in reverse
    # so that the first of the ascending order is written last (:65-71 last-write-wins).
    order = np.lexsort((idx, r))[::-1]
    winner = np.full(H * W, -1, dtype=np.int64)
    winner[cell[order]] = idx[order]
    # cells whose minimum rho is shared by >1 point: reference's pick is unspecified
    minr = np.full(H * W, np.inf)
    np.minimum.at(minr, cell, r)
    at_min = r == minr[cell]
    tie = np.bincount(cell[at_min], minlength=H * W) > 1
    has = winner >= 0
    w = winner[has]
    x, y, z, it = p[w, 0], p[w, 1], p[w, 2], p[w, 3]
    dist = np.zeros(H * W)
    hgt = np.full(H * W, float(LIDAR_MIN_HEIGHT))                  # :54,62
    inten = np.zeros(H * W)
    dist[has] = np.sqrt(x * x + y * y)                             # :32,56
    hgt[has] = z                                                   # :47
    inten[has] = it                                                # :45
    flip = lambda a: a.reshape(H, W)[::-1].copy()                  # :74
    return {
        "distance": flip(dist), "height": flip(hgt), "intensity": flip(inten),
        "winner": flip(winner), "tie_mask": flip(tie),
        "dropped": int(bad.sum()), "err": int(np.bitwise_or.reduce(err)) if n else 0,
    }


def bev_edges(max_range, res_step):
    """np.arange(-max_range, max_range, step) restated (birds_eye_view_generator.py:43):
    n = ceil((stop-start)/step); edge[i] = start + i*delta with
    delta = (start+step) - start evaluated in fp64 (numpy's arange fill rule)."""
    start = float(-max_range)
    stop = float(max_range)
    step = float(res_step)
    n = int(np.ceil((stop - start) / step))
    delta = (start + step) - start
    return start + np.arange(n, dtype=np.float64) * delta


def bev_bin(v, edges):
    """histogram2d explicit-edge binning (birds_eye_view_generator.py:44): left-closed bins,
    last bin right-closed, NaN / out-of-range -> -1."""
    v = np.asarray(v, dtype=np.float64)
    j = np.searchsorted(edges, v, side="right")
    j = np.where(v == edges[-1], j - 1, j)
    b = j - 1
    ok = (j >= 1) & (j <= len(edges) - 1)
    return np.where(ok, b, -1)


def bev_counts(points, res, max_range=120):
    """Un-normalised, un-flipped cell counts [x_bin, y_bin] (int64)."""
    p = np.asarray(points, dtype=np.float64)
    ex = bev_edges(max_range, res[1])
    ey = bev_edges(max_range, res[0])
    bx = bev_bin(p[:, 0], ex)
    by = bev_bin(p[:, 1], ey)
    nbx, nby = len(ex) - 1, len(ey) - 1
    ok = (bx >= 0) & (by >= 0)
    flat = bx[ok] * nby + by[ok]
    return np.bincount(flat, minlength=nbx * nby).reshape(nbx, nby).astype(np.int64)


def density_lut(nmax=64):
    """normalize(), birds_eye_view_generator.py:31-37, as a function of the count."""
    n = np.arange(nmax + 1, dtype=np.float64)
    with np.errstate(divide="ignore"):
        v = 255 * np.minimum(np.log(n + 1) / np.log(64), 1)
    v[0] = 0.0
    return v


def bev_density(points, res, max_range=120):
    """Pre-colormap float64 image of generate_birds_eye_view (:41-46): counts ->
    255*min(ln(N+1)/ln 64, 1) -> both axes reversed."""
    c = bev_counts(points, res, max_range)
    lut = density_lut(64)
    d = lut[np.minimum(c, 64)]
    return d[::-1, ::-1].copy()                                    # :46


# ---- MV3D extension channels -- PARITY UNPINNED -------------------
def bev_height_intensity(points, res, max_range, slice_edges):
    """Our own definition (no reference code exists): per height slice [lo,hi) the max z
    per BEV cell (background = slice lo), and the intensity of the highest point per cell
    over all slices (background 0).  Tie on z -> lowest index.  Flipped like the density."""
    p = np.asarray(points, dtype=np.float64)
    ex = bev_edges(max_range, res[1])
    ey = bev_edges(max_range, res[0])
    bx = bev_bin(p[:, 0], ex)
    by = bev_bin(p[:, 1], ey)
    nbx, nby = len(ex) - 1, len(ey) - 1
    z = p[:, 2]
    ok = (bx >= 0) & (by >= 0) & ~np.isnan(z)
    flat = bx * nby + by
    m = len(slice_edges) - 1
    hmaps = np.empty((m, nbx, nby))
    for s in range(m):
        lo, hi = slice_edges[s], slice_edges[s + 1]
        sel = ok & (z >= lo) & (z < hi)
        h = np.full(nbx * nby, float(lo))
        np.maximum.at(h, flat[sel], z[sel])
        hmaps[s] = h.reshape(nbx, nby)
    sel = ok & (z >= slice_edges[0]) & (z < slice_edges[-1])
    idx = np.arange(len(p))[sel]
    order = np.lexsort((-idx, z[sel]))            # ascending z, descending idx; last write = max z, lowest idx
    top = np.full(nbx * nby, -1, dtype=np.int64)
    top[flat[sel][order]] = idx[order]
    inten = np.zeros(nbx * nby)
    has = top >= 0
    inten[has] = p[top[has], 3]
    return hmaps[:, ::-1, ::-1].copy(), inten.reshape(nbx, nby)[::-1, ::-1].copy()


# ---- op-for-op port used ONLY as the timed CPU baseline --------------------------------
def port_front_view_once(points, kind):
    """One call of lidar_2d_front_view(cmap=None), same numpy call sequence and the same
    redundant work as the reference (:26-39,60-74), for timing."""
    x = points[:, 0]; y = points[:, 1]; z = points[:, 2]; r = points[:, 3]
    d = np.sqrt(x ** 2 + y ** 2)
    l2 = np.sqrt(x ** 2 + y ** 2 + z ** 2)
    xi = np.arctan2(-y, x) / RES_RAD[1]
    yi = np.arcsin(z / l2) / RES_RAD[0]
    xi -= X_MIN
    yi -= Y_MIN
    vals, bg = {"intensity": (r, 0), "height": (z, LIDAR_MIN_HEIGHT), "distance": (d, 0)}[kind]
    yy = yi.astype(int)
    xx = xi.astype(int)
    img = np.ones((H, W)) * bg
    s = np.flipud(l2.argsort())
    img[yy[s], xx[s]] = vals[s]
    return np.flipud(img)


def port_generate_lidar_2d_front_view(points):
    """generate_lidar_2d_front_view: three full passes (:81-83)."""
    return {k: port_front_view_once(points, k) for k in ("intensity", "distance", "height")}


def port_generate_birds_eye_view(points, res, max_range=120, python_loop=True):
    """generate_birds_eye_view up to the colormap (:42-46).  python_loop=True keeps the
    reference's per-cell Python `normalize` loop (:31-37); False uses the vectorised
    equivalent so the cost of that artefact is visible separately."""
    pts = np.array(points)
    bins = (np.arange(-max_range, max_range, res[1]), np.arange(-max_range, max_range, res[0]))
    ch, _, _ = np.histogram2d(pts[:, 0], pts[:, 1], bins=bins)
    if python_loop:
        l64 = np.log(64)
        for i, rowv in enumerate(ch):
            for j, c in enumerate(rowv):
                if c > 0:
                    ch[i][j] = 255 * min(np.log(c + 1) / l64, 1)
    else:
        nz = ch > 0
        ch[nz] = 255 * np.minimum(np.log(ch[nz] + 1) / np.log(64), 1)
    return np.flipud(np.fliplr(ch))


# ---------------------------------------------------------------------------------------------------------
# Consumers of the range scan.  numpy restatements; test infrastructure only.
# ---------------------------------------------------------------------------------------------------------

def invariant_mask_update(mask, reference, float_pixels):
    """process/create_lidar_invariant_mask.py:59-62: the mask keeps only pixels equal to the reference
    frame in every frame seen so far (elementwise ==, so NaN never matches).  In place, returns mask."""
    same = reference == float_pixels                                   # :61
    np.logical_and(mask, same, mask)                                   # :62
    return mask


def roll_width(image, rotate_by):
    """train/loader.py:47-50: np.roll of an (height, width, channels) array along its width."""
    return np.roll(image, rotate_by, axis=1)


def back_project(centroids, bboxes, distance_data, height_data):
    """train/predict.py:230-293 (back_project_2D_2_3D).  centroids (B,2) [column, row], bboxes (B,4)
    [left, top, right, bottom], both float and updated IN PLACE like the reference; distance_data /
    height_data (B,h,w).  Returns xyz (B,4) float64 (last column stays 0).  Boxes are expected inside the
    image (the reference's slices would silently clip and then unravel with the unclipped shape)."""
    from math import sin, cos
    B, h, w = distance_data.shape
    xyz = np.zeros((B, 4))                                             # :231
    valid = np.logical_and(distance_data > 0, height_data > LIDAR_MIN_HEIGHT)   # :236
    for i in range(B):
        cx, cy = int(centroids[i, 0]), int(centroids[i, 1])
        if (not valid[i, cy, cx]) and bboxes[i, 0] != 0 and bboxes[i, 2] != 0:  # :243
            left, right = int(bboxes[i, 0]), int(bboxes[i, 2]) + 1     # :245-248
            top, bottom = int(bboxes[i, 1]), int(bboxes[i, 3]) + 1
            rows, cols = np.mgrid[top:bottom, left:right]
            d = np.sqrt((cols - cx) * (cols - cx) + (rows - cy) * (rows - cy)).astype(np.float64)   # :251-253
            d = np.where(valid[i, top:bottom, left:right], d, 10e7)    # :255
            k = int(np.argmin(d))                                      # :256 first minimum in row-major order
            if d.flat[k] == 10e7:                                      # :262-269 nothing valid in the box
                centroids[i, :] = 0
                bboxes[i, :] = 0
            else:
                r, c = divmod(k, right - left)                         # :271
                centroids[i, 1], centroids[i, 0] = r + top, c + left   # :272-275
        if not (centroids[i, 0] == 0 and centroids[i, 1] == 0):        # :277
            dist = float(distance_data[i, int(centroids[i, 1]), int(centroids[i, 0])])   # :278
            hgt = float(height_data[i, int(centroids[i, 1]), int(centroids[i, 0])])      # :279
            theta = (centroids[i, 0] + X_MIN) * RES_RAD[1]             # :280
            dist += 0.75                                               # :283
            xyz[i, 0] = dist * cos(theta)                              # :285
            xyz[i, 1] = -dist * sin(theta)                             # :286
            xyz[i, 2] = hgt                                            # :287
    return xyz


MIN_PROB = 0.5          # train/predict.py:28
MAX_BBX_DIST = 5.0      # train/predict.py:31
FAR_DELTA = (9.0, 3.0, 3.0)   # train/predict.py:85


def decode_boxes(distance_img, height_img, y_pred, bbox_2d, centroid_3d):
    """First half of find_bbox_3d (train/predict.py:94-135): the decoded corner sets (K,8,3), in the reference's
    order (column-major over the window), that are not far from the centroid.  A window pixel takes part when
    its column holds a positive pixel SOMEWHERE and its row holds one SOMEWHERE (:105), not only when it is
    positive itself."""
    from math import sin, cos
    h, w = distance_img.shape
    y3 = np.reshape(y_pred, (h, w, 26))                                         # :95
    pos_rows, pos_cols = np.where(y3[:, :, 1] >= MIN_PROB)                      # :96
    row_has, col_has = np.zeros(h, bool), np.zeros(w, bool)
    row_has[pos_rows] = True
    col_has[pos_cols] = True
    (ulx, uly), (lrx, lry) = bbox_2d                                            # :99
    c3 = np.asarray(centroid_3d, dtype=np.float64)[:3]                          # is_far reads [0], [1], [2] (:87-89)
    boxes = []
    for x in range(max(ulx - 100, 0), min(lrx + 100, w)):                       # :103 (outside the image nothing matches)
        if not col_has[x]:
            continue
        for y in range(max(uly - 2, 0), min(lry + 2, h)):                       # :104
            if not row_has[y]:
                continue
            d, z = float(distance_img[y, x]), float(height_img[y, x])           # :108-109
            theta, phi = (x + X_MIN) * RES_RAD[1], (y + Y_MIN) * RES_RAD[0]     # :110-111
            p = np.array([d * cos(theta), -d * sin(theta), z])                  # :112-115
            rot_z = np.array([[cos(theta), -sin(theta), 0.0], [sin(theta), cos(theta), 0.0], [0.0, 0.0, 1.0]])
            rot_y = np.array([[cos(phi), 0.0, sin(phi)], [0.0, 1.0, 0.0], [-sin(phi), 0.0, cos(phi)]])
            rot = np.matmul(rot_z, rot_y)                                       # :128
            c = np.matmul(rot, np.reshape(y3[y, x, 2:], (8, 3)).transpose()).transpose() + p   # :116,129
            if (np.abs(c - c3) <= np.array(FAR_DELTA)).all():                   # :83-92, :132
                boxes.append(c)
    return np.array(boxes).reshape(len(boxes), 8, 3)


def find_bbox_3d(distance_img, height_img, y_pred, bbox_2d, centroid_3d):
    """train/predict.py:94-199.  Returns (pred[7] = x, y, z, yaw, length, width, height; candidate box (8,3) or None)."""
    from math import atan2, cos, pi
    boxes = decode_boxes(distance_img, height_img, y_pred, bbox_2d, centroid_3d)
    if len(boxes) == 0:
        return np.zeros(7), None                                                # :137-138
    diff = boxes[:, None] - boxes[None]                                         # :146-151 all pairs
    d = np.sqrt((diff * diff).sum(axis=(2, 3)))
    counts = np.logical_and(d > 0, d < MAX_BBX_DIST).sum(axis=1)
    cand = boxes[counts == counts.max()].mean(axis=0)                           # :159-164 ties: all of them
    pred = np.zeros(7)
    pred[:3] = np.mean(cand, 0)                                                 # :167
    yaws, ls, ws, hs = [], [], [], []
    for i in range(4):                                                          # :174-192
        dx, dy = cand[i, 0] - cand[i + 4, 0], cand[i, 1] - cand[i + 4, 1]
        yaw = atan2(dy, dx)
        yaws.append(yaw)
        ls.append(abs(dx / cos(yaw) if yaw != pi / 2.0 else dy))
        dx, dy = cand[i, 0] - cand[i + 2, 0], cand[i, 1] - cand[i + 2, 1]
        ws.append(abs(dy / cos(yaw) if yaw != pi / 2.0 else dx))
        hs.append(abs(cand[i, 2] - cand[i + 1, 2]))
    pred[3], pred[4], pred[5], pred[6] = np.mean(yaws), np.mean(ls), np.mean(ws), np.mean(hs)   # :194-197
    return pred, cand


MIN_HEAT = 2            # train/predict.py:30
MIN_BBOX_AREA = 100     # train/predict.py:29


def obstacle_heatmap(y_pred, input_shape):
    """train/predict.py:34-52: 4x4 blocks of heat around positive pixels, thresholded.  heatmap[y-2:y+2, x-2:x+2]
    with y < 2 or x < 2 is an EMPTY python slice (negative start), so those positives add nothing."""
    h, w = input_shape[:2]
    pred = np.reshape(y_pred[:, 1], (h, w))                                     # :34-35
    heat = np.zeros((h, w))
    for y, x in zip(*np.where(pred >= MIN_PROB)):                               # :36-48
        if y >= 2 and x >= 2:
            heat[y - 2:y + 2, x - 2:x + 2] += 1
    heat[heat <= MIN_HEAT] = 0                                                  # :51
    return heat


def find_obstacle(y_pred, input_shape):
    """train/predict.py:33-81.  Clusters are the 4-connected components of the thresholded heat map
    (scipy.ndimage.label, default structure -- the reference's own dependency, :16,52), visited in label order."""
    from scipy.ndimage import label
    labels, n = label(obstacle_heatmap(y_pred, input_shape))                    # :52
    best, max_area = None, 0
    for k in range(1, n + 1):                                                   # :57-70
        ys, xs = np.nonzero(labels == k)
        area = int(xs.max() - xs.min()) * int(ys.max() - ys.min())
        if area > max_area:
            best, max_area = ((int(xs.min()), int(ys.min())), (int(xs.max()), int(ys.max()))), area
    if max_area <= MIN_BBOX_AREA:                                               # :73-74
        return None, None, None
    box = ((best[0][0] + 2, best[0][1] + 2), (best[1][0] - 2, best[1][1] - 2))  # :76
    centroid = (int((box[0][0] + box[1][0]) / 2.0), int((box[0][1] + box[1][1]) / 2.0))   # :77-78
    return centroid, box, max_area


def predict_tail(prediction, distance, height, fns=None):
    """What predict_point_cloud does after model.predict (train/predict.py:347-362, use_regression=True), composed
    from the three restatements above (or, for pinning, from the reference's own three functions in `fns`)."""
    f_obstacle, f_back, f_bbox = fns or (find_obstacle, back_project, find_bbox_3d)
    centroid_2d, bbox_2d, _ = f_obstacle(prediction, distance.shape + (3,))     # :347
    pred = np.array([0, 0, 0, 0])                                               # :348
    if centroid_2d is not None:
        centroids = np.array(centroid_2d).reshape(1, 2)                         # :351 (integer arrays, as there)
        bboxes = np.array(bbox_2d).reshape(1, 4)
        centroid_3d = f_back(centroids, bboxes, distance[None], height[None])[0]   # :355
        if not (centroid_3d[0] == 0. and centroid_3d[1] == 0.):                 # :358
            pred, _ = f_bbox(distance, height, prediction, bbox_2d, centroid_3d)   # :359
    return pred
