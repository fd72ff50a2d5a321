"""The numpy restatement (oracle/lidar_oracle.py) against the committed outputs of the
reference's own functions (tests/golden/, made by tests/golden/make_golden.py).
Bit-exact: the oracle computes in fp64 with the same numpy primitives."""
import numpy as np
import pytest

from conftest import golden
from oracle import lidar_oracle as lo
from oracle import reference_loader as rl

FRONT = ["hdl32_seed0", "hdl32_seed1_shuffled_dropout", "hdl64_seed2", "vls128_seed3_near",
         "small_seed4", "edges"]


def _P(g):
    p = np.zeros((len(g["points"]), 5))
    p[:, :4] = g["points"]
    return p


@pytest.mark.parametrize("name", FRONT)
def test_front_view_matches_reference_golden(name):
    g = golden(name)
    o = lo.front_view(_P(g))
    for k in ("distance", "height", "intensity"):
        a, b = g[k], o[k]
        # cells where the reference's own winner is unspecified (exact fp64 range ties) are excluded
        ok = ~o["tie_mask"]
        assert np.array_equal(a[ok], b[ok]), k
    assert o["winner"].shape == (32, 1801)


@pytest.mark.parametrize("name", FRONT + ["bev_edges"])
def test_bev_density_matches_reference_golden(name):
    g = golden(name)
    for key in [k for k in g.files if k.startswith("bev_")]:
        r = float(key[4:])
        d = lo.bev_density(_P(g), (r, r), 120)
        assert d.shape == g[key].shape
        assert np.array_equal(d, g[key]), key


@pytest.mark.parametrize("name", ["raises_high_elevation", "raises_origin", "raises_nan", "raises_low_elevation"])
def test_reference_index_error_cases(name):
    g = golden(name)
    with pytest.raises(IndexError):
        lo.front_view(_P(g), strict=True)
    o = lo.front_view(_P(g), strict=False)
    assert o["dropped"] == 1 and o["err"] != 0


def test_edges_rule():
    for R, res in [(120, 0.2), (120, 2), (120, 0.05), (100, 0.25), (120, 0.1)]:
        assert np.array_equal(lo.bev_edges(R, res), np.arange(-R, R, res).astype(np.float64))


def test_cell_convention_landmarks():
    # col 900 straight ahead, y>0 -> smaller column, el=0 -> output row 8
    pts = np.array([[10, 0, 0, 1], [0, 10, 0, 2], [0, -10, 0, 3], [-10, 1e-9, 0, 4], [-10, -0.0, 0, 5]], float)
    w = lo.front_view(pts)["winner"]
    rows, cols = np.nonzero(w >= 0)
    got = {int(w[r, c]): (int(r), int(c)) for r, c in zip(rows, cols)}
    assert got[0] == (8, 900) and got[1] == (8, 450) and got[2] == (8, 1350)
    assert got[3] == (8, 0) and got[4] == (8, 1800)


def test_density_lut_saturates():
    lut = lo.density_lut(64)
    assert lut[0] == 0 and lut[63] == 255.0 and lut[64] == 255.0 and 0 < lut[1] < lut[2]


@pytest.mark.skipif(not rl.available(), reason="reference tree not mounted (GPU box)")
def test_oracle_equals_live_reference(synth):
    p = synth.make_sweep("hdl32", seed=11, n_az=300, shuffle=True)
    P = synth.as_reference_array(p)
    ref = rl.front_view(P)
    o = lo.front_view(P)
    for k in ("distance", "height", "intensity"):
        assert np.array_equal(ref[k + "_float"], o[k])
    assert np.array_equal(rl.bev_density(P, (0.5, 0.5)), lo.bev_density(P, (0.5, 0.5)))
    port = lo.port_generate_lidar_2d_front_view(P)
    for k in port:
        assert np.array_equal(port[k], ref[k + "_float"])
    assert np.array_equal(lo.port_generate_birds_eye_view(P, (2, 2)), rl.bev_density(P, (2, 2)))


def test_extension_definition_hand_case():
    """MV3D extension (parity unpinned): the definition the GPU tests compare against."""
    from oracle import lidar_oracle as lo
    # 2 m cells: x in [0,2) -> bin 60, y in [0,2) -> bin 60 of 119; output is reversed on both axes
    P = np.array([[0.5, 0.5, -0.5, 10.0, 0], [1.0, 1.5, 0.7, 20.0, 0], [1.2, 0.2, 0.7, 30.0, 0],   # same cell
                  [0.5, 0.5, 5.0, 40.0, 0],                                                           # above the last edge
                  [-3.0, 0.5, 0.2, 50.0, 0]])                                                         # another cell
    h, it = lo.bev_height_intensity(P, (2, 2), 120, [-1.0, 0.0, 1.0])
    assert h.shape == (2, 119, 119) and it.shape == (119, 119)
    c = (118 - 60, 118 - 60)
    assert h[0][c] == -0.5 and h[1][c] == np.float64(0.7)
    assert it[c] == 20.0                      # highest point, tie on z -> lowest index
    c2 = (118 - 58, 118 - 60)
    assert h[0][c2] == -1.0 and h[1][c2] == 0.2 and it[c2] == 50.0
    assert (h[0] == -1.0).sum() == 119 * 119 - 1 and (it != 0).sum() == 2


def test_back_project_matches_reference_golden():
    """oracle.back_project vs outputs of the reference's own back_project_2D_2_3D (train/predict.py:230-293)."""
    g = golden("back_project_seed0")
    cen, box = g["centroids"].copy(), g["bboxes"].copy()
    xyz = lo.back_project(cen, box, g["distance"].astype(np.float64), g["height"].astype(np.float64))
    assert np.array_equal(xyz, g["xyz"])
    assert np.array_equal(cen, g["centroids_out"]) and np.array_equal(box, g["bboxes_out"])
    moved = (g["centroids"] != g["centroids_out"]).any(axis=1)
    zeroed = (g["bboxes_out"] == 0).all(axis=1)
    assert moved.sum() >= 4 and zeroed.sum() >= 2 and (~moved).sum() >= 3     # the fixture exercises every branch


@pytest.mark.skipif(not rl.full_tree_available(), reason="reference tree not mounted")
def test_back_project_equals_live_reference(synth):
    dist, hgt, cen, box = synth.make_back_project_case(n_items=18, seed=7)
    c1, b1, c2, b2 = cen.copy(), box.copy(), cen.copy(), box.copy()
    want = rl.back_project_fn()(c1, b1, dist.astype(np.float64), hgt.astype(np.float64))
    got = lo.back_project(c2, b2, dist.astype(np.float64), hgt.astype(np.float64))
    assert np.array_equal(want, got) and np.array_equal(c1, c2) and np.array_equal(b1, b2)


def test_invariant_mask_and_roll_definitions():
    """create_lidar_invariant_mask.py:59-62 and loader.py:47-50 are two numpy calls each; pin the restatement
    on a hand case (a NaN never matches, the mask never comes back)."""
    ref = np.array([1.0, 2.0, np.nan, 4.0])
    mask = np.ones(4, dtype=bool)
    lo.invariant_mask_update(mask, ref, np.array([1.0, 2.5, np.nan, 4.0]))
    assert mask.tolist() == [True, False, False, True]
    lo.invariant_mask_update(mask, ref, np.array([0.0, 2.0, 3.0, 4.0]))
    assert mask.tolist() == [False, False, False, True]
    img = np.arange(2 * 5 * 3).reshape(2, 5, 3)
    out = lo.roll_width(img, 2)
    assert np.array_equal(out[:, 2], img[:, 0]) and np.array_equal(out[:, 0], img[:, 3])


def _bbox_cases(synth):
    g = golden("bbox_decode")
    for k in range(int(g["n"])):
        dist, hgt, y, bb, c3 = synth.make_bbox_decode_case(seed=int(g["seed%d" % k]))
        assert np.array_equal(np.array(bb), g["bbox%d" % k]) and np.array_equal(c3, g["centroid%d" % k])   # same inputs
        yield dist, hgt, y, bb, c3, g["pred%d" % k], (None if bool(g["none%d" % k]) else g["box%d" % k])


def test_find_bbox_3d_matches_reference_golden(synth):
    """oracle.find_bbox_3d vs outputs of the reference's own find_bbox_3d (train/predict.py:94-199)."""
    for dist, hgt, y, bb, c3, pred, box in _bbox_cases(synth):
        got_pred, got_box = lo.find_bbox_3d(dist.astype(np.float64), hgt.astype(np.float64), y, bb, c3)
        assert np.array_equal(got_pred, pred) and np.array_equal(got_box, box)
        assert len(lo.decode_boxes(dist.astype(np.float64), hgt.astype(np.float64), y, bb, c3)) > 100


@pytest.mark.skipif(not rl.full_tree_available(), reason="reference tree not mounted")
def test_find_bbox_3d_equals_live_reference(synth):
    for seed, strays in ((11, 0), (12, 12)):
        dist, hgt, y, bb, c3 = synth.make_bbox_decode_case(seed=seed, strays=strays)
        d64, h64 = dist.astype(np.float64), hgt.astype(np.float64)
        want = rl.find_bbox_3d_fn()(d64, h64, y, bb, c3)
        got = lo.find_bbox_3d(d64, h64, y, bb, c3)
        assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1])
    far = rl.find_bbox_3d_fn()(d64, h64, y, bb, c3 + 100.0)          # every set is far: zeros(7), None
    assert far[1] is None and not far[0].any()
    assert lo.find_bbox_3d(d64, h64, y, bb, c3 + 100.0)[1] is None


def _obstacle_tuple(res):
    c, b, a = res
    return [0] * 8 if c is None else [1, c[0], c[1], b[0][0], b[0][1], b[1][0], b[1][1], a]


def test_find_obstacle_matches_reference_golden(synth):
    """oracle.find_obstacle vs outputs of the reference's own find_obstacle (train/predict.py:33-81)."""
    rows = golden("find_obstacle")["rows"]
    assert (rows[:, 1] == 0).sum() >= 2 and (rows[:, 1] == 1).sum() >= 6
    for row in rows:
        got = _obstacle_tuple(lo.find_obstacle(synth.make_obstacle_case(seed=int(row[0])), (32, 1801, 3)))
        assert [int(v) for v in got] == [int(v) for v in row[1:]]


@pytest.mark.skipif(not rl.full_tree_available(), reason="reference tree not mounted")
def test_find_obstacle_equals_live_reference(synth):
    for seed in (40, 41, 42):
        y = synth.make_obstacle_case(seed=seed, blobs=7, speckle=0.01)
        assert _obstacle_tuple(rl.find_obstacle_fn()(y, (32, 1801, 3))) == _obstacle_tuple(lo.find_obstacle(y, (32, 1801, 3)))


@pytest.mark.skipif(not rl.full_tree_available(), reason="reference tree not mounted")
def test_post_network_chain_equals_live_reference(synth):
    """find_obstacle -> back_project_2D_2_3D -> find_bbox_3d as predict_point_cloud chains them (:347-362)."""
    fns = (rl.find_obstacle_fn(), rl.back_project_fn(), rl.find_bbox_3d_fn())
    kinds = set()
    for seed in range(4):
        dist, hgt, y = synth.make_detection_case(seed)
        d, h = dist.astype(np.float64), hgt.astype(np.float64)
        want, got = lo.predict_tail(y, d, h, fns), lo.predict_tail(y, d, h)
        assert want.shape == got.shape and np.array_equal(want, got)
        kinds.add(len(got))
    assert kinds == {4, 7}
