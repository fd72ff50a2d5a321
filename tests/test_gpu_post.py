"""-m gpu: the consumers of the range scan through libsfpe.so vs the oracle."""
import numpy as np
import pytest

from conftest import golden, load_pkg, load_sub
from oracle import lidar_oracle as lo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def post():
    load_pkg()
    return load_sub("post")


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


def test_back_project_vs_reference_golden(post):
    g = golden("back_project_seed0")
    for dt in (np.float64, np.float32):
        cen, box = g["centroids"].copy(), g["bboxes"].copy()
        xyz = post.back_project_2D_2_3D(cen, box, g["distance"].astype(dt), g["height"].astype(dt))
        assert np.array_equal(cen, g["centroids_out"]) and np.array_equal(box, g["bboxes_out"])   # pixel choice: exact
        assert np.array_equal(xyz[:, 2:], g["xyz"][:, 2:])                                        # height, pad: exact
        # x, y go through the device's fp64 cos/sin (<= 2 ulp) against libm's: 1e-13 relative
        np.testing.assert_allclose(xyz[:, :2], g["xyz"][:, :2], rtol=1e-13, atol=1e-13)


def test_back_project_device_on_projector_output(post, torch):
    """Range scans as the projector writes them ((B,H,W,3) float32) feed the back-projection without a copy."""
    synth = load_sub("synth")
    dist, hgt, cen, box = synth.make_back_project_case(n_items=30, seed=3)
    rs = np.zeros(dist.shape + (3,), np.float32)
    rs[..., 0], rs[..., 1], rs[..., 2] = dist, hgt, 7.0
    c_want, b_want = cen.copy(), box.copy()
    want = lo.back_project(c_want, b_want, dist.astype(np.float64), hgt.astype(np.float64))
    d_cen, d_box = torch.as_tensor(cen).cuda(), torch.as_tensor(box).cuda()
    xyz, status = post.back_project_device(d_cen, d_box, range_scan=torch.as_tensor(rs).cuda())
    assert np.array_equal(d_cen.cpu().numpy(), c_want) and np.array_equal(d_box.cpu().numpy(), b_want)
    np.testing.assert_allclose(xyz.cpu().numpy(), want, rtol=1e-13, atol=1e-13)
    st = status.cpu().numpy()
    moved = (cen != c_want).any(axis=1)
    zeroed = (b_want == 0).all(axis=1) & (box != 0).any(axis=1)
    assert np.array_equal(st == 2, zeroed) and np.array_equal(st == 1, moved & ~zeroed) and not (st == 3).any()


def test_back_project_flags_items_outside_the_image(post):
    dist = np.ones((3, 32, 1801)); hgt = np.zeros((3, 32, 1801))
    dist[1] = 0
    cen = np.array([[5.0, 40.0], [10.0, 3.0], [12.0, 4.0]])          # row 40 is outside; box of item 1 too wide
    box = np.array([[1.0, 1, 9, 9], [1.0, 1, 1801, 5], [1.0, 1, 30, 5]])
    with pytest.raises(IndexError):
        post.back_project_2D_2_3D(cen, box, dist, hgt)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_invariant_mask_matches_numpy(post, torch, dtype):
    rng = np.random.default_rng(5)
    shape = (32, 1801, 3)
    ref = rng.integers(0, 4, shape).astype(dtype)
    ref[0, 0, 0] = np.nan
    frames = np.repeat(ref[None], 37, axis=0)
    flips = rng.random(frames.shape) < 0.01
    frames[flips] += 1
    frames[:, 3, 7, 1] = ref[3, 7, 1]                # a pixel that stays
    want = np.ones(shape, dtype=bool)
    for f in frames:
        lo.invariant_mask_update(want, ref, f)
    im = post.InvariantMask(torch.as_tensor(ref).cuda())
    im.update(torch.as_tensor(frames[:20]).cuda())
    got = im.update(torch.as_tensor(frames[20:]).cuda()).cpu().numpy().astype(bool)
    assert np.array_equal(got, want)
    assert not got[0, 0, 0] and got[3, 7, 1] and 0 < got.sum() < got.size


def test_invariant_mask_single_channel_stride(post, torch):
    """pix_stride 3: one channel of an HWC frame (the reference keeps one mask per channel)."""
    import ctypes as C
    rng = np.random.default_rng(6)
    ref = rng.integers(0, 3, (32, 1801, 3)).astype(np.float32)
    frames = np.repeat(ref[None], 9, axis=0)
    frames[rng.random(frames.shape) < 0.05] += 1
    want = (frames[..., 1] == ref[..., 1]).all(axis=0)
    d_ref, d_fr = torch.as_tensor(ref).cuda(), torch.as_tensor(frames).cuda()
    mask = torch.ones((32, 1801), dtype=torch.uint8, device="cuda")
    pr = load_sub("_session").projector(enable_bev=False)
    rc = pr.lib.sfpe_invariant_mask_update(pr._ctx, C.c_void_p(d_ref.data_ptr() + 4), C.c_void_p(d_fr.data_ptr() + 4), 0,
                                           9, ref.size, 32 * 1801, 3, C.c_void_p(mask.data_ptr()), None)
    assert rc == 0
    torch.cuda.synchronize()
    assert np.array_equal(mask.cpu().numpy().astype(bool), want)


@pytest.mark.parametrize("shape", [(5, 32, 1801, 3), (3, 32, 1801), (4, 7, 13, 5), (2, 1, 1, 1)])
def test_roll_width_matches_numpy(post, torch, shape):
    rng = np.random.default_rng(8)
    img = rng.standard_normal(shape).astype(np.float32)
    shifts = rng.integers(0, shape[2], shape[0])
    shifts[0] = 0
    got = post.roll_width(torch.as_tensor(img).cuda(), shifts).cpu().numpy()
    for f in range(shape[0]):
        assert np.array_equal(got[f], lo.roll_width(img[f], int(shifts[f])))
    neg = post.roll_width(torch.as_tensor(img).cuda(), -3 - 2 * shape[2]).cpu().numpy()     # any integer, like np.roll
    assert np.array_equal(neg[-1], lo.roll_width(img[-1], -3 - 2 * shape[2]))


def test_roll_width_labels_float64(post, torch):
    """Labels roll with the image (loader.py:47-49): (H, W, label_dim) float64 works the same way."""
    rng = np.random.default_rng(9)
    lab = rng.standard_normal((3, 32, 1801, 2))
    got = post.roll_width(torch.as_tensor(lab).cuda(), [1, 900, 1800]).cpu().numpy()
    for f, s in enumerate([1, 900, 1800]):
        assert np.array_equal(got[f], np.roll(lab[f], s, axis=1))


def _bbox_cases():
    synth = load_sub("synth")
    g = golden("bbox_decode")
    for k in range(int(g["n"])):
        yield synth.make_bbox_decode_case(seed=int(g["seed%d" % k])) + (g["pred%d" % k], g["box%d" % k])


def test_find_bbox_3d_vs_reference_golden(post):
    """Decisions (which corner sets take part, neighbour counts, ties) follow the reference; the fp64 trigonometry
    and 3x3 products are the device's: 1e-9 absolute on metres / radians."""
    for dist, hgt, y, bb, c3, pred, box in _bbox_cases():
        for dt in (np.float64, np.float32):
            got_pred, got_box = post.find_bbox_3d(dist.astype(dt), hgt.astype(dt), y, bb, c3)
            np.testing.assert_allclose(got_box, box, rtol=0, atol=1e-9)
            np.testing.assert_allclose(got_pred, pred, rtol=0, atol=1e-9)


def test_find_bbox_3d_none_when_everything_is_far(post):
    dist, hgt, y, bb, c3 = load_sub("synth").make_bbox_decode_case(seed=5)
    pred, box = post.find_bbox_3d(dist, hgt, y, bb, c3 + 100.0)
    assert box is None and pred.shape == (7,) and not pred.any()


def test_find_bbox_3d_device_batch_and_overflow(post, torch):
    synth = load_sub("synth")
    cases = [synth.make_bbox_decode_case(seed=s, strays=st) for s, st in ((21, 0), (22, 6), (23, 20))]
    scan = np.zeros((3, 32, 1801, 3), np.float32)
    for i, c in enumerate(cases):
        scan[i, ..., 0], scan[i, ..., 1] = c[0], c[1]
    y = torch.as_tensor(np.stack([c[2] for c in cases])).cuda()
    bb = torch.tensor([[c[3][0][0], c[3][0][1], c[3][1][0], c[3][1][1]] for c in cases], dtype=torch.int32)
    c3 = torch.as_tensor(np.stack([c[4] for c in cases]))
    pred, box, n = post.find_bbox_3d_device(y, bb, c3, range_scan=torch.as_tensor(scan).cuda())
    for i, c in enumerate(cases):
        d64, h64 = c[0].astype(np.float64), c[1].astype(np.float64)
        want_pred, want_box = lo.find_bbox_3d(d64, h64, c[2], c[3], c[4])
        assert int(n[i]) == len(lo.decode_boxes(d64, h64, c[2], c[3], c[4]))          # same sets kept
        np.testing.assert_allclose(box[i].cpu().numpy(), want_box, rtol=0, atol=1e-9)
        np.testing.assert_allclose(pred[i].cpu().numpy(), want_pred, rtol=0, atol=1e-9)
    _, _, n_small = post.find_bbox_3d_device(y, bb, c3, range_scan=torch.as_tensor(scan).cuda(), cap=8)
    assert (n_small.cpu().numpy() == -1).all()                                          # scratch too small: flagged


def _obstacle_tuple(res):
    c, b, a = res
    return [0] * 8 if c is None else [int(v) for v in (1, c[0], c[1], b[0][0], b[0][1], b[1][0], b[1][1], a)]


def test_find_obstacle_vs_reference_golden(post):
    synth = load_sub("synth")
    for row in golden("find_obstacle")["rows"]:
        y = synth.make_obstacle_case(seed=int(row[0]))
        assert _obstacle_tuple(post.find_obstacle(y, (32, 1801, 3))) == [int(v) for v in row[1:]]
        assert _obstacle_tuple(post.find_obstacle(y.astype(np.float64), (32, 1801, 3))) == [int(v) for v in row[1:]]


def test_find_obstacle_batch_shapes_ties_and_snakes(post, torch):
    """Equal-area clusters (first in raster order wins), a one-pixel-wide spiral (worst case for label
    propagation), U shapes that merge late, 3-channel output, everything positive, nothing positive."""
    H, W = 32, 1801
    frames = []
    p = np.zeros((H, W), np.float32)
    p[10:20, 1000:1030] = 1; p[4:14, 200:230] = 1; p[4:14, 600:630] = 1          # three equal boxes: the one at row 4, col 200
    frames.append(p)
    p = np.zeros((H, W), np.float32)
    for r in range(2, 30, 6):                                                     # a snake: long rows joined at alternating ends
        p[r:r + 2, 2:1790] = 1
        col = 1786 if (r // 6) % 2 == 0 else 2
        p[r:r + 8, col:col + 3] = 1
    frames.append(p)
    p = np.zeros((H, W), np.float32)
    p[3:25, 50:54] = 1; p[3:25, 90:94] = 1; p[22:25, 50:94] = 1                    # a U: two arms that meet at the bottom
    p[5:9, 1500:1530] = 1
    frames.append(p)
    frames.append(np.ones((H, W), np.float32))
    frames.append(np.zeros((H, W), np.float32))
    rng = np.random.default_rng(3)
    frames.append((rng.random((H, W)) < 0.25).astype(np.float32))                  # dense noise: many ragged clusters
    y = np.zeros((len(frames), H * W, 3), np.float32)
    for i, f in enumerate(frames):
        y[i, :, 1] = f.ravel()
    out = post.find_obstacle_device(torch.as_tensor(y).cuda(), H, W).cpu().numpy()
    for i in range(len(frames)):
        assert out[i].tolist() == _obstacle_tuple(lo.find_obstacle(y[i], (H, W, 3))), i
    assert out[0, 0] == 1 and out[4, 0] == 0


def test_post_network_chain_on_device(post, torch):
    """locate_obstacles (batch, nothing leaves the device) and predict_point_cloud_tail (one frame, reference types)
    against the oracle's composition of the three restatements."""
    synth = load_sub("synth")
    cases = [synth.make_detection_case(seed) for seed in range(6)]
    want = [lo.predict_tail(y, d.astype(np.float64), h.astype(np.float64)) for d, h, y in cases]
    assert {len(w) for w in want} == {4, 7}
    scan = np.zeros((len(cases), 32, 1801, 3), np.float32)
    for i, (d, h, _) in enumerate(cases):
        scan[i, ..., 0], scan[i, ..., 1] = d, h
    pred, found = post.locate_obstacles(torch.as_tensor(np.stack([c[2] for c in cases])).cuda(), torch.as_tensor(scan).cuda())
    pred, found = pred.cpu().numpy(), found.cpu().numpy()
    for i, w in enumerate(want):
        assert bool(found[i]) == (len(w) == 7)
        np.testing.assert_allclose(pred[i], w if len(w) == 7 else np.zeros(7), rtol=0, atol=1e-9)
    for i in (0, 2):
        d, h, y = cases[i]
        got = post.predict_point_cloud_tail(y, {"distance_float": d.astype(np.float64), "height_float": h.astype(np.float64)})
        assert got.shape == want[i].shape
        np.testing.assert_allclose(got, want[i], rtol=0, atol=1e-9)


@pytest.mark.parametrize("cmap,pickle", [("jet", True), ("gray", False), (None, True)])
def test_velodyne_message_renderer_equals_the_drop_ins(cmap, pickle):
    """handle_msg's PointCloud2 branch (extract_rosbag.py:162-187) in one device pass == the two drop-in calls the
    reference makes there, on the points its per-point decode would produce."""
    pkg = load_pkg()
    synth = load_sub("synth")
    er = load_sub("process.extract_rosbag")
    pts = synth.make_sweep("hdl32", seed=12, n_az=600)
    P = synth.as_reference_array(pts)
    r = er.VelodyneMessageRenderer(topdown_res=2, topdown_max_range=120, cmap=cmap, pickle=pickle)
    density, images = r(synth.make_pointcloud2_blob(pts))
    want_density = pkg.generate_birds_eye_view(P, 0, res=(2, 2), max_range=120, cmap=cmap or "gray").astype(np.uint8)
    want = pkg.generate_lidar_2d_front_view(P, cmap=cmap)
    assert density.dtype == np.uint8 and np.array_equal(density, want_density)
    kinds = ["intensity", "distance", "height"]
    assert sorted(images) == sorted(kinds + ([k + "_float" for k in kinds] if pickle else []))
    for k in images:
        assert images[k].dtype == want[k].dtype and np.array_equal(images[k], want[k]), k
    if pickle:
        ref = lo.front_view(P)
        for k in kinds:
            assert np.array_equal(images[k + "_float"], ref[k])
    bad = np.vstack([pts, np.zeros((1, 4), np.float32)])                       # a point at the origin
    with pytest.raises(IndexError):
        r(synth.make_pointcloud2_blob(bad))
