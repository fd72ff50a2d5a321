"""Parity of the CUDA path (through the C ABI) against the reference goldens and the oracle.

Bit-exact for cell assignment, winners, counts and the range images.  The log-normalised density
takes 64 values 255*min(ln(N+1)/ln 64, 1): the float32 image must equal the oracle's float64 image
rounded to float32 bit for bit, the float64 image within 4 ulp (device log vs libm log; far inside
BASELINE.json's north_star tolerance of 1e-6 relative)."""
import numpy as np
import pytest

from conftest import golden, load_pkg
from oracle import lidar_oracle as lo

pytestmark = pytest.mark.gpu

FRONT = ["hdl32_seed0", "hdl32_seed1_shuffled_dropout", "hdl64_seed2", "vls128_seed3_near",
         "small_seed4", "edges"]
DENSITY_RTOL = 9e-16   # float64 images: 4 ulp


@pytest.fixture(scope="module")
def pkg():
    import torch
    assert torch.cuda.is_available()
    return load_pkg()


def _P(pts32):
    p = np.zeros((len(pts32), 5))
    p[:, :4] = pts32
    return p


# ---------------------------------------------------------------- drop-in entries vs reference goldens
@pytest.mark.parametrize("name", FRONT)
def test_dropin_front_view_vs_reference_golden(pkg, name):
    g = golden(name)
    P = _P(g["points"])
    out = pkg.generate_lidar_2d_front_view(P)
    tie = lo.front_view(P)["tie_mask"]
    for k in ("distance", "height", "intensity"):
        a = out[k + "_float"]
        assert a.dtype == np.float64 and a.shape == (32, 1801)
        assert np.array_equal(a[~tie], g[k][~tie]), k
        assert out[k] is not None
    img, flt = pkg.lidar_2d_front_view(P, res=(1.33, 0.2), fov=(-30.67, 10.67), type="height")
    assert np.array_equal(flt[~tie], g["height"][~tie]) and img is flt


@pytest.mark.parametrize("name", FRONT + ["bev_edges"])
def test_dropin_bev_density_vs_reference_golden(pkg, name):
    g = golden(name)
    P = _P(g["points"])
    for key in [k for k in g.files if k.startswith("bev_")]:
        r = float(key[4:])
        d = pkg.birds_eye_view_density(P, (r, r), 120)
        assert d.shape == g[key].shape and d.dtype == np.float64
        _check_density(d, g[key])                                 # occupancy exact, values within 4 ulp
        rgb = pkg.generate_birds_eye_view(P, 0, (r, r), 120)
        assert rgb.dtype == np.uint8 and rgb.shape == d.shape + (3,)


@pytest.mark.parametrize("name", ["raises_high_elevation", "raises_origin", "raises_nan", "raises_low_elevation"])
def test_dropin_raises_like_reference(pkg, name):
    g = golden(name)
    with pytest.raises(IndexError):
        pkg.generate_lidar_2d_front_view(_P(g["points"]))


def test_dropin_asserts_like_reference(pkg):
    P = _P(golden("small_seed4")["points"])
    with pytest.raises(AssertionError):
        pkg.lidar_2d_front_view(P, res=(1.33,), fov=(-30.67, 10.67), type="height")
    with pytest.raises(AssertionError):
        pkg.lidar_2d_front_view(P, res=(1.33, 0.2), fov=(1.0, 10.67), type="height")
    with pytest.raises(AssertionError):
        pkg.lidar_2d_front_view(P, res=(1.33, 0.2), fov=(-30.67, 10.67), type="depth")


# ---------------------------------------------------------------- batched device API vs oracle
def _run_batch(pkg, sweeps, ragged=True, **kw):
    import torch
    allp = np.concatenate(sweeps) if sweeps else np.zeros((0, 4), np.float32)
    offs = np.zeros(len(sweeps) + 1, np.int64)
    offs[1:] = np.cumsum([len(s) for s in sweeps])
    counts = kw.pop("counts", True)
    proj_kw = {k: kw.pop(k) for k in list(kw) if k in ("bev_res", "max_range", "key_prefix_bits", "path",
                                                       "extrinsic", "slice_edges", "firing_stride")}
    pr = pkg.LidarProjector(device=0, **proj_kw)
    d = torch.from_numpy(allp).cuda() if len(allp) else torch.zeros((1, 4), dtype=torch.float32, device="cuda")
    if ragged:
        out = pr.project(d, n_frames=len(sweeps), frame_offsets=torch.from_numpy(offs).cuda(),
                         max_n=max([len(s) for s in sweeps] + [0]), winner=True, counts=counts, **kw)
    else:
        out = pr.project(d, n_frames=len(sweeps), winner=True, counts=counts, **kw)
    torch.cuda.synchronize()
    res = {k: v.cpu().numpy() for k, v in out.items() if hasattr(v, "cpu")}
    res["_path"] = pr.last_path
    pr.close()
    return res


def _check_frame(res, f, pts32, bev_res=(0.2, 0.2), layout="hwc", strict_ok=True):
    P = _P(pts32)
    o = lo.front_view(P, strict=False)
    assert np.array_equal(res["winner"][f], o["winner"].astype(np.int32))
    rng = res["range"][f]
    if layout == "hwc":
        rng = np.moveaxis(rng, -1, 0)
    for c, k in enumerate(("distance", "height", "intensity")):
        assert np.array_equal(rng[c], o[k].astype(rng.dtype)), k
    assert res["status"][f] == o["err"] and res["dropped"][f] == o["dropped"]
    if "counts" in res:
        cnt = lo.bev_counts(P, bev_res, 120)
        assert np.array_equal(res["counts"][f], cnt.astype(np.int32))
    dens = lo.bev_density(P, bev_res, 120)
    _check_density(res["bev"][f], dens)


def _check_density(got, dens64):
    if got.dtype == np.float32:
        want = dens64.astype(np.float32)
        if not np.array_equal(got, want):
            bad = np.argwhere(got != want)
            raise AssertionError("BEV density: %d cells differ, rows %d..%d; first (row, col, got, want): %s" % (
                len(bad), bad[:, 0].min(), bad[:, 0].max(),
                [(int(r), int(c), float(got[r, c]), float(want[r, c])) for r, c in bad[:8]]))
    else:
        assert np.array_equal(got == 0, dens64 == 0)
        np.testing.assert_allclose(got, dens64, rtol=DENSITY_RTOL, atol=0)


def test_batch_ragged_f32_hwc(pkg, synth):
    sweeps = [synth.make_sweep("hdl32", seed=20), synth.make_sweep("hdl32", seed=21, dropout=0.3, shuffle=True),
              np.zeros((0, 4), np.float32), synth.make_sweep("hdl64", seed=22, n_az=200),
              synth.make_sweep("hdl32", seed=23, n_az=1)]
    res = _run_batch(pkg, sweeps)
    for f, s in enumerate(sweeps):
        _check_frame(res, f, s)
    # an empty frame is pure background
    assert (res["winner"][2] == -1).all() and (res["bev"][2] == 0).all()
    assert (res["range"][2][..., 1] == -2.0).all() and (res["range"][2][..., 0] == 0).all()


def test_batch_uniform_f64_chw(pkg, synth):
    sweeps = [synth.make_sweep("hdl32", seed=30 + i, n_az=300) for i in range(40)]   # > one scratch chunk
    res = _run_batch(pkg, sweeps, ragged=False, range_dtype="f64", range_layout="chw", bev_dtype="f64")
    for f in (0, 1, 17, 39):
        _check_frame(res, f, sweeps[f], layout="chw")


def test_batch_other_bev_resolutions(pkg, synth):
    sweeps = [synth.make_sweep("vls128", seed=40, n_az=400, near_heavy=True), synth.make_sweep("hdl32", seed=41)]
    for res_m in ((2, 2), (0.5, 0.25), (0.1, 0.1)):
        res = _run_batch(pkg, sweeps, bev_res=res_m)
        for f, s in enumerate(sweeps):
            _check_frame(res, f, s, bev_res=res_m)


def test_invalid_points_are_dropped_and_flagged(pkg, synth):
    base = synth.make_sweep("hdl32", seed=50, n_az=64)
    bad = np.array([[0, 0, 0, 1], [np.nan, 1, 1, 1], [1, 0, 5, 1], [1, 0, -30, 2], [np.inf, 1, 1, 1]], np.float32)
    sweeps = [np.vstack([base, bad]), base]
    res = _run_batch(pkg, sweeps)
    for f, s in enumerate(sweeps):
        _check_frame(res, f, s)
    assert res["status"][0] != 0 and res["status"][1] == 0 and res["dropped"][0] >= 3


@pytest.mark.parametrize("prefix_bits", [8, 12, 16, 24])
def test_forced_key_collisions_settle_exactly(pkg, synth, prefix_bits):
    """With only a few range bits in the packed key nearly every multi-point cell takes the
    exact-compare CAS path; the winners must still equal the oracle's."""
    sweeps = [synth.make_sweep("hdl32", seed=60), synth.make_sweep("vls128", seed=61, n_az=600, shuffle=True)]
    res = _run_batch(pkg, sweeps, key_prefix_bits=prefix_bits)
    for f, s in enumerate(sweeps):
        _check_frame(res, f, s)


def test_near_equal_and_duplicate_ranges(pkg):
    rng = np.random.default_rng(3)
    base = np.array([8.0, 1.0, -0.25], np.float32)
    pts = []
    for k in range(200):                      # same cell, ranges differing in the last float32 bits
        p = base.copy()
        p[rng.integers(0, 3)] = np.nextafter(p[rng.integers(0, 3)], np.float32(100), dtype=np.float32)
        for _ in range(int(rng.integers(0, 4))):
            j = rng.integers(0, 3)
            p[j] = np.nextafter(p[j], np.float32(100 if rng.random() < 0.5 else -100), dtype=np.float32)
        pts.append([p[0], p[1], p[2], float(k)])
    pts += pts[:50]                           # exact duplicates: tie -> lowest index
    pts = np.array(pts, np.float32)
    res = _run_batch(pkg, [pts, pts[::-1].copy()])
    _check_frame(res, 0, pts)
    _check_frame(res, 1, pts[::-1].copy())


def test_float64_inputs_not_float32_representable(pkg):
    """The drop-in takes the reference's float64 (N,5) arrays as they are."""
    rng = np.random.default_rng(5)
    n = 5000
    az = rng.uniform(-np.pi, np.pi, n); el = np.deg2rad(rng.uniform(-30, 10, n)); r = rng.uniform(1, 90, n)
    P = np.stack([r * np.cos(el) * np.cos(az), r * np.cos(el) * np.sin(az), r * np.sin(el),
                  rng.uniform(0, 255, n), rng.integers(0, 32, n).astype(float)], axis=1)
    o = lo.front_view(P)
    out = pkg.generate_lidar_2d_front_view(P)
    for k in ("distance", "height", "intensity"):
        assert np.array_equal(out[k + "_float"][~o["tie_mask"]], o[k][~o["tie_mask"]]), k
    d = pkg.birds_eye_view_density(P, (0.2, 0.2), 120)
    _check_density(d, lo.bev_density(P, (0.2, 0.2), 120))


def test_dropin_float32_input_is_float64_semantics_and_thread_safe(pkg, synth):
    """dtype contract of the drop-ins: a float32 array gives the images of the float64 array of the same
    values (the reference's callers always build float64 arrays).  And the free functions may be called
    from several threads at once (they share one projector; the reference's functions are pure)."""
    import threading
    sweeps = [synth.make_sweep("hdl32", seed=8200 + i, n_az=700) for i in range(4)]
    want = [pkg.generate_lidar_2d_front_view(_P(s)) for s in sweeps]
    dens = [pkg.birds_eye_view_density(_P(s), (0.2, 0.2), 120) for s in sweeps]
    a32 = pkg.generate_lidar_2d_front_view(sweeps[0])                  # float32 (N,4) as it is
    for k in ("distance_float", "height_float", "intensity_float"):
        assert a32[k].dtype == np.float64 and np.array_equal(a32[k], want[0][k])
    errs = []

    def work(i):
        try:
            for _ in range(10):
                got = pkg.generate_lidar_2d_front_view(_P(sweeps[i]))
                for k in ("distance_float", "height_float", "intensity_float"):
                    assert np.array_equal(got[k], want[i][k]), (i, k)
                assert np.array_equal(pkg.birds_eye_view_density(_P(sweeps[i]), (0.2, 0.2), 120), dens[i]), i
        except Exception as e:
            errs.append(repr(e))
    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs


def test_order_invariance_and_conservation(pkg, synth):
    """Size-independent properties at the bench's sweep size: shuffling the points changes
    nothing but the winner indices; every in-range point is counted exactly once."""
    s = synth.make_sweep("hdl32", seed=70)
    perm = np.random.default_rng(0).permutation(len(s))
    res = _run_batch(pkg, [s, s[perm]])
    assert np.array_equal(res["range"][0], res["range"][1])
    assert np.array_equal(res["counts"][0], res["counts"][1])
    assert np.array_equal(perm[res["winner"][1][res["winner"][1] >= 0]], res["winner"][0][res["winner"][0] >= 0])
    inside = (np.abs(s[:, 0].astype(np.float64)) < 119.7) & (np.abs(s[:, 1].astype(np.float64)) < 119.7)
    assert res["counts"][0].sum() >= inside.sum() and res["counts"][0].sum() <= len(s)


# ---------------------------------------------------------------- fused one-sweep-per-SM path
def test_fused_path_ragged(pkg, synth):
    sweeps = [synth.make_sweep("hdl32", seed=80), synth.make_sweep("hdl32", seed=81, dropout=0.3, shuffle=True),
              np.zeros((0, 4), np.float32), synth.make_sweep("hdl64", seed=82), synth.make_sweep("hdl32", seed=83, n_az=1),
              synth.make_sweep("vls128", seed=84, near_heavy=True)]
    res = _run_batch(pkg, sweeps, path=2, counts=False)
    assert res["_path"] == "fused"
    for f, s in enumerate(sweeps):
        _check_frame(res, f, s)


def test_fused_path_many_frames_uniform_chw(pkg, synth):
    """More frames than SMs: every CTA takes several sweeps; scratch and tiles must self-clean."""
    sweeps = [synth.make_sweep("hdl32", seed=100 + i, n_az=256, shuffle=(i % 3 == 0)) for i in range(320)]
    res = _run_batch(pkg, sweeps, ragged=False, path=2, counts=False, range_layout="chw")
    assert res["_path"] == "fused"
    for f in (0, 1, 147, 148, 149, 295, 296, 319):
        _check_frame(res, f, sweeps[f], layout="chw")


@pytest.mark.parametrize("res_m", [(2, 2), (0.5, 0.25), (0.1, 0.1), (0.05, 0.05)])   # 0.05: 176 tiles, 4 per bucket (C5)
def test_fused_path_other_bev_resolutions(pkg, synth, res_m):
    sweeps = [synth.make_sweep("vls128", seed=40, n_az=400, near_heavy=True), synth.make_sweep("hdl32", seed=41),
              synth.make_sweep("hdl32", seed=42, n_az=3)]
    res = _run_batch(pkg, sweeps, bev_res=res_m, path=2, counts=False)
    assert res["_path"] == "fused"
    for f, s in enumerate(sweeps):
        _check_frame(res, f, s, bev_res=res_m)


def test_fused_path_invalid_points_and_edges(pkg, synth):
    base = synth.make_sweep("hdl32", seed=50, n_az=64)
    bad = np.array([[0, 0, 0, 1], [np.nan, 1, 1, 1], [1, 0, 5, 1], [1, 0, -30, 2], [np.inf, 1, 1, 1]], np.float32)
    sweeps = [np.vstack([base, bad]), base, golden("edges")["points"], golden("bev_edges")["points"][:0]]
    res = _run_batch(pkg, sweeps, path=2, counts=False)
    for f, s in enumerate(sweeps):
        _check_frame(res, f, s)
    assert res["status"][0] != 0 and res["status"][1] == 0


def test_fused_path_saturating_counts_and_range_ties(pkg):
    """>255 points in one BEV cell (u8 tile saturation) and hundreds of near-equal / equal ranges
    in one range cell (quantised-key collisions settled on the fp64 range)."""
    rng = np.random.default_rng(9)
    n = 3000
    pts = np.empty((n, 4), np.float32)
    pts[:, 0] = 10.05 + rng.uniform(0, 0.05, n)      # all inside BEV cell x in [10.0, 10.2)
    pts[:, 1] = 4.05 + rng.uniform(0, 0.05, n)
    pts[:, 2] = -1.0 + rng.uniform(0, 0.002, n)      # ranges within a few mm: same 1/64 m bucket
    pts[:, 3] = np.arange(n)
    pts[100:200] = pts[0:100]                        # exact duplicates
    pts[100:200, 3] = 1000 + np.arange(100)
    res = _run_batch(pkg, [pts, pts[::-1].copy()], path=2, counts=False)
    _check_frame(res, 0, pts)
    _check_frame(res, 1, pts[::-1].copy())
    assert res["bev"][0].max() == 255.0


def test_fused_equals_generic_on_bench_shape(pkg, synth):
    sweeps = [synth.make_sweep("hdl32", seed=200 + i) for i in range(4)]
    a = _run_batch(pkg, sweeps, ragged=False, path=2, counts=False)
    b = _run_batch(pkg, sweeps, ragged=False, path=1, counts=False)
    for k in ("range", "winner", "bev", "status", "dropped"):
        assert np.array_equal(a[k], b[k]), k


# ---------------------------------------------------------------- fp32 fast path: guard bands hold
def _adversarial_points(n, seed):
    """Directions spread over the whole sphere band the fast path accepts, ranges from 0.3 m to
    250 m, plus points placed right on range-scan cell boundaries and BEV bin edges."""
    rng = np.random.default_rng(seed)
    az = rng.uniform(-np.pi, np.pi, n)
    el = np.deg2rad(rng.uniform(-33.0, 33.0, n))
    r = np.exp(rng.uniform(np.log(0.3), np.log(250.0), n))
    k = n // 4
    az[:k] = (np.round(rng.uniform(0, 1800, k)) - 900.0) * lo.RES_RAD[1] + rng.normal(0, 2e-7, k)   # column boundaries
    el[k:2 * k] = (np.round(rng.uniform(0, 31, k)) + lo.Y_MIN) * lo.RES_RAD[0] + rng.normal(0, 2e-7, k)  # row boundaries
    p = np.empty((n, 4), np.float32)
    p[:, 0] = r * np.cos(el) * np.cos(az)
    p[:, 1] = -r * np.cos(el) * np.sin(az)
    p[:, 2] = r * np.sin(el)
    p[:, 3] = rng.uniform(0, 255, n)
    e = np.arange(-120, 120, 0.2)
    p[2 * k:3 * k, 0] = (e[rng.integers(0, len(e), k)] + rng.normal(0, 1e-5, k)).astype(np.float32)  # BEV x edges
    p[2 * k:3 * k, 1] = (e[rng.integers(0, len(e), k)] + rng.normal(0, 1e-5, k)).astype(np.float32)  # BEV y edges
    return p


def test_fastpath_guard_bands_hold(pkg, synth):
    import torch
    pr = pkg.LidarProjector(device=0)
    worst = np.zeros(4)
    sets = [_adversarial_points(1 << 20, s) for s in range(4)] + [synth.make_sweep("vls128", seed=5, near_heavy=True)]
    for p in sets:
        worst = np.maximum(worst, pr.selftest_fastpath(torch.from_numpy(p).cuda()))
    pr.close()
    # deviation / guard band: < 1 is required for exactness, < 0.5 leaves a 2x margin
    assert (worst < 0.5).all(), worst
    assert worst.max() > 0.0


def test_fused_path_adversarial_boundaries(pkg):
    """Points sitting on cell boundaries: the guarded fp32 decisions and the fp64 re-dos together
    must still give the oracle's cells and winners (in-FOV subset, so the reference would not raise)."""
    p = _adversarial_points(200000, 11)
    P = np.zeros((len(p), 5)); P[:, :4] = p
    _, _, _, err = lo.front_view_cells(P)
    sweeps = [p[err == 0][:150000], p[:100000]]
    res = _run_batch(pkg, sweeps, path=2, counts=False)
    assert res["_path"] == "fused"
    for f, s in enumerate(sweeps):
        _check_frame(res, f, s)


# ---------------------------------------------------------------- next rows: PointCloud2 decode, streaming entry
def test_decode_pointcloud2_matches_struct_unpack(pkg, synth):
    import torch
    pts = synth.make_sweep("hdl32", seed=90, n_az=500)
    blob = synth.make_pointcloud2_blob(pts)
    pr = pkg.LidarProjector(device=0)
    got = pr.decode_pointcloud2(torch.from_numpy(blob).cuda(), len(pts), synth.PC2_POINT_STEP, synth.PC2_OFFSETS)
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy(), pts)
    # float64 fields (datatype 8) are rounded to float32 like np.float32(value)
    rec = np.zeros(len(pts), dtype=np.dtype({"names": list("xyzi"), "formats": ["<f8"] * 4, "offsets": [0, 8, 16, 24], "itemsize": 40}))
    for k, c in zip("xyzi", range(4)):
        rec[k] = pts[:, c].astype(np.float64) * (1 + 1e-9)
    got64 = pr.decode_pointcloud2(torch.from_numpy(rec.view(np.uint8).reshape(-1)).cuda(), len(pts), 40, (0, 8, 16, 24), dtype="f64")
    torch.cuda.synchronize()
    want = np.stack([rec[k] for k in "xyzi"], axis=1).astype(np.float32)
    assert np.array_equal(got64.cpu().numpy(), want)
    pr.close()


def test_streaming_entry_matches_oracle(pkg, synth):
    sv = pkg.StreamingFrontView(device=0, max_points=70000)
    for seed in (91, 92):
        pts = synth.make_sweep("hdl32", seed=seed, dropout=0.02)
        out = sv(synth.make_pointcloud2_blob(pts).tobytes())
        o = lo.front_view(_P(pts))
        assert out.shape == (32, 1801, 3) and out.dtype == np.float32
        for c, k in enumerate(("distance", "height", "intensity")):          # train/predict.py:339-341 order
            assert np.array_equal(out[..., c], o[k].astype(np.float32)), k
    bad = np.vstack([synth.make_sweep("hdl32", seed=93, n_az=8), np.zeros((1, 4), np.float32)])
    with pytest.raises(IndexError):
        sv(synth.make_pointcloud2_blob(bad).tobytes())
    sv.close()


# ---------------------------------------------------------------- extension: MV3D height slices + intensity (parity unpinned)
@pytest.mark.parametrize("with_extrinsic", [False, True])
def test_extension_height_slices_and_intensity(pkg, synth, with_extrinsic):
    """no reference code exists; the oracle's definition is the contract
    (per slice max z, empty = slice floor; intensity of the highest point, tie -> lowest index).
    Default-off: the drop-in outputs (range scan, density) are unchanged when it is on."""
    import torch
    edges = [-2.0, -1.0, 0.0, 1.0, 3.0]
    sweeps = [synth.make_sweep("hdl32", seed=70, n_az=700), synth.make_sweep("hdl64", seed=71, n_az=300, shuffle=True)]
    # duplicates of the highest points: ties on z must go to the lowest index
    dup = sweeps[0][np.argsort(-sweeps[0][:, 2])[:50]].copy()
    dup[:, 3] += 1.0
    sweeps[0] = np.concatenate([sweeps[0], dup]).astype(np.float32)
    M = None
    if with_extrinsic:
        M = np.array([[0.999, -0.04, 0.0, 0.3], [0.04, 0.999, 0.0, -0.2], [0.0, 0.0, 1.0, 0.5]])
    pr = pkg.LidarProjector(device=0, bev_res=(0.5, 0.5), slice_edges=edges, extrinsic=M, enable_range=False)
    offs = np.cumsum([0] + [len(s) for s in sweeps]).astype(np.int64)
    out = pr.project(torch.from_numpy(np.concatenate(sweeps)).cuda(), n_frames=len(sweeps),
                     frame_offsets=torch.from_numpy(offs).cuda(), max_n=max(len(s) for s in sweeps))
    torch.cuda.synchronize()
    assert pr.last_path == "fused+ext"
    for f, s in enumerate(sweeps):
        P = _P(s)
        if with_extrinsic:   # the transformed cloud is again a float32 cloud (sfpe_common.cuh apply_extrinsic)
            xyz = (M[:, 0] * P[:, 0:1] + M[:, 1] * P[:, 1:2]) + M[:, 2] * P[:, 2:3] + M[:, 3]
            P[:, :3] = xyz.astype(np.float32).astype(np.float64)
        h, it = lo.bev_height_intensity(P, (0.5, 0.5), 120, edges)
        assert np.array_equal(out["bev_height"][f].cpu().numpy(), h.astype(np.float32))
        assert np.array_equal(out["bev_intensity"][f].cpu().numpy(), it.astype(np.float32))
        _check_density(out["bev"][f].cpu().numpy(), lo.bev_density(P, (0.5, 0.5), 120))
    pr.close()


@pytest.mark.parametrize("with_extrinsic", [False, True])
def test_extension_next_to_the_fused_kernel_equals_generic(pkg, synth, with_extrinsic):
    """Extension E on the bench's shape (HDL-64E-shaped sweeps, 0.2 m, four slices): the fused kernel followed by
    the fill / project / finish kernels of sfpe_ext.cu gives the generic path's planes bit for bit, on
    back-to-back calls with different batch sizes (the top-key scratch resets itself) and into NaN-filled outputs.
    With the rigid extrinsic the fused kernel's XF instantiation transforms every point it loads (fp64, rounded
    to float32, as the generic path does): range scan, winners, density and planes all equal the generic path's."""
    import torch
    edges = [-2.0, -1.0, 0.0, 1.0, 3.0]
    M = np.array([[0.999, -0.04, 0.01, 0.3], [0.04, 0.999, 0.0, -0.2], [-0.01, 0.0, 1.0, 0.15]]) if with_extrinsic else None
    sweeps = [synth.make_sweep("hdl64", seed=7300 + i, shuffle=(i % 5 == 2), dropout=0.1 * (i % 3)) for i in range(27)]
    sweeps[4] = sweeps[4][:0]
    top = sweeps[0][np.argsort(-sweeps[0][:, 2])[:200]].copy()   # ties on z: the lowest index wins
    top[:, 3] += 1.0
    sweeps[0] = np.concatenate([sweeps[0], top]).astype(np.float32)
    fused = pkg.LidarProjector(device=0, slice_edges=edges, extrinsic=M)   # one ctx for all calls: its scratch is reused
    for n_frames in (27, 3, 27):
        part = sweeps[:n_frames]
        offs = torch.from_numpy(np.cumsum([0] + [len(s) for s in part]).astype(np.int64)).cuda()
        d = torch.from_numpy(np.concatenate(part)).cuda()
        res = {}
        for path in (0, 1):
            pr = pkg.LidarProjector(device=0, slice_edges=edges, extrinsic=M, path=1) if path else fused
            out = pr.alloc_outputs(n_frames, winner=True)
            for k in ("range", "bev", "bev_height", "bev_intensity"):
                out[k].fill_(float("nan"))
            pr.project(d, n_frames=n_frames, frame_offsets=offs, max_n=max(len(s) for s in part), out=out)
            torch.cuda.synchronize()
            assert pr.last_path == ("generic" if path else "fused+ext")
            res[path] = out
            if path:
                pr.close()
        for k in ("range", "winner", "bev", "bev_height", "bev_intensity", "status", "dropped"):
            assert torch.equal(res[0][k], res[1][k]), (n_frames, k)
    fused.close()
    P = _P(sweeps[0])
    if with_extrinsic:
        xyz = (M[:, 0] * P[:, 0:1] + M[:, 1] * P[:, 1:2]) + M[:, 2] * P[:, 2:3] + M[:, 3]
        P[:, :3] = xyz.astype(np.float32).astype(np.float64)
        o = lo.front_view(P, strict=False)
        assert np.array_equal(res[0]["winner"][0].cpu().numpy(), o["winner"].astype(np.int32))
    h, it = lo.bev_height_intensity(P, (0.2, 0.2), 120, edges)
    assert np.array_equal(res[0]["bev_height"][0].cpu().numpy(), h.astype(np.float32))
    assert np.array_equal(res[0]["bev_intensity"][0].cpu().numpy(), it.astype(np.float32))


# ---------------------------------------------------------------- randomized: fused path == generic path
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_fused_equals_generic_randomized(pkg, synth, seed):
    """Many frames per CTA (dynamic frame claiming, zero fill one frame ahead, sector overwrite),
    ragged sizes, mixed sensors / orders / drop-outs, several BEV resolutions and both range
    layouts: every output of the fused kernel equals the generic (global-atomics, fp64) path."""
    rng = np.random.default_rng(seed)
    kinds = ["hdl32", "hdl64", "vls128"]
    sweeps = []
    for i in range(330):        # > 2 frames per CTA on 148 SMs
        k = kinds[int(rng.integers(0, 3))]
        n_az = int(rng.integers(1, 180))
        sweeps.append(synth.make_sweep(k, seed=1000 * seed + i, n_az=n_az, dropout=float(rng.uniform(0, 0.5)),
                                       shuffle=bool(rng.integers(0, 2)), near_heavy=bool(rng.integers(0, 2))))
    sweeps[7] = np.zeros((0, 4), np.float32)
    sweeps[100] = synth.make_sweep("hdl32", seed=77)          # one full-size sweep among the small ones
    res_m = [(0.2, 0.2), (0.5, 0.25), (1.0, 1.0)][seed % 3]
    layout = "hwc" if seed % 2 else "chw"
    a = _run_batch(pkg, sweeps, path=2, counts=False, bev_res=res_m, range_layout=layout)
    b = _run_batch(pkg, sweeps, path=1, counts=False, bev_res=res_m, range_layout=layout)
    assert a["_path"] == "fused" and b["_path"] == "generic"
    for k in ("range", "winner", "bev", "status", "dropped"):
        assert np.array_equal(a[k], b[k]), k



# ---------------------------------------------------------------- the benchmarked launch shape
def _device_batch(pkg, pts_all, n_frames, n_pts, path, bev_res, firing_stride=0):
    """One project() call over n_frames uniform sweeps already on the device; outputs stay there."""
    import torch
    pr = pkg.LidarProjector(device=0, bev_res=bev_res, path=path, firing_stride=firing_stride)
    out = pr.project(pts_all, n_frames=n_frames, uniform_n=n_pts, winner=True)
    torch.cuda.synchronize()
    assert pr.last_path == ("fused" if path == 2 else "generic")
    pr.close()
    return out


@pytest.mark.parametrize("sensor,res,n_frames,n_oracle", [("hdl32", 0.2, 320, 16), ("hdl64", 0.2, 300, 16),
                                                          ("vls128", 0.05, 300, 3)])
def test_fused_full_size_batches_like_the_bench(pkg, synth, sensor, res, n_frames, n_oracle):
    """The shape bench.py times (C2 / C3 / C5): >= 300 FULL-SIZE sweeps in one project() call on the
    fused path, so that every CTA rasterises two or three full sweeps back to back while its store
    warp writes the previous sweep's BEV image.  Every output equals the generic path's (compared on
    the device), a sample of frames equals the oracle, and the BEV image is written completely
    (the output buffers start as NaN)."""
    import torch
    sweeps = [synth.make_sweep(sensor, seed=5000 + i, near_heavy=(sensor == "vls128"), shuffle=(i % 7 == 3))
              for i in range(n_frames)]
    n_pts = len(sweeps[0])
    d = torch.from_numpy(np.concatenate(sweeps)).cuda()
    a = _device_batch(pkg, d, n_frames, n_pts, 2, (res, res), firing_stride=synth.SENSORS[sensor][0])   # as bench.py does
    b = _device_batch(pkg, d, n_frames, n_pts, 1, (res, res))
    for k in ("range", "winner", "bev", "status", "dropped"):
        assert torch.equal(a[k], b[k]), k
    assert int((a["status"] != 0).sum()) == 0
    pick = np.linspace(0, n_frames - 1, n_oracle).astype(int)
    for f in pick:
        P = _P(sweeps[f])
        o = lo.front_view(P, strict=False)
        assert np.array_equal(a["winner"][f].cpu().numpy(), o["winner"].astype(np.int32))
        rng = np.moveaxis(a["range"][f].cpu().numpy(), -1, 0)
        for c, kk in enumerate(("distance", "height", "intensity")):
            assert np.array_equal(rng[c], o[kk].astype(np.float32)), kk
        _check_density(a["bev"][f].cpu().numpy(), lo.bev_density(P, (res, res), 120))


def test_fused_rewrites_every_bev_byte(pkg, synth):
    """The store warp writes each 32-byte sector of a frame's image exactly once: poison the output
    (NaN) and the neighbouring guard cells, run, and look for survivors / overruns.  An unaligned
    BEV base (the tile coordinates follow the absolute address) goes through the same check."""
    import torch
    sweeps = [synth.make_sweep("hdl32", seed=6000 + i, n_az=512 if i % 2 else 2048) for i in range(9)]
    n = [len(s) for s in sweeps]
    offs = torch.from_numpy(np.concatenate([[0], np.cumsum(n)]).astype(np.int64)).cuda()
    d = torch.from_numpy(np.concatenate(sweeps)).cuda()
    pr = pkg.LidarProjector(device=0, path=2)
    cells = pr.nbx * pr.nby
    for shift in (0, 1, 37, 250):
        buf = torch.full((len(sweeps) * cells + 1024,), float("nan"), dtype=torch.float32, device="cuda")
        out = pr.alloc_outputs(len(sweeps))
        out["bev"] = buf[256 + shift: 256 + shift + len(sweeps) * cells].view(len(sweeps), pr.nbx, pr.nby)
        pr.project(d, n_frames=len(sweeps), frame_offsets=offs, max_n=max(n), out=out)
        torch.cuda.synchronize()
        assert pr.last_path == "fused"
        assert not torch.isnan(out["bev"]).any()
        assert torch.isnan(buf[:256 + shift]).all() and torch.isnan(buf[256 + shift + len(sweeps) * cells:]).all()
        for f in (0, 4, 8):
            _check_density(out["bev"][f].cpu().numpy(), lo.bev_density(_P(sweeps[f]), (0.2, 0.2), 120))
    pr.close()


@pytest.mark.parametrize("sensor,res,frames_per_sm", [("vls128", 0.2, 6), ("hdl64", 0.2, 7), ("hdl32", 0.1, 6)])
def test_fused_store_warp_never_skips_a_frame(pkg, synth, sensor, res, frames_per_sm):
    """Six or seven sweeps per CTA with a long point pass and the small (0.2 m) image: the store warp is
    done with a frame's zero lines long before the compute warps reach the next frame's BEV phase.  It
    must then wait for that frame's block cursor to be reset (a stale cursor reads "no blocks left" and
    the frame's zero lines would never be written; with fewer than five frames per CTA the CTA's last two
    frames, written by all warps at the end, hide it).  The third case is the opposite regime: a short point
    pass and a four times larger image, so the compute warps wait for the store warp instead.  Outputs start
    as NaN; every frame is compared on the device with the generic path's result for the same sweep."""
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    distinct = 8
    n_frames = (sms * frames_per_sm + distinct - 1) // distinct * distinct
    sweeps = [synth.make_sweep(sensor, seed=7000 + i, near_heavy=(sensor == "vls128")) for i in range(distinct)]
    n_pts = len(sweeps[0])
    base = torch.from_numpy(np.concatenate(sweeps)).cuda()
    ref = _device_batch(pkg, base, distinct, n_pts, 1, (res, res))
    d = base.repeat(n_frames // distinct, 1)
    pr = pkg.LidarProjector(device=0, path=2, bev_res=(res, res), firing_stride=synth.SENSORS[sensor][0])
    for rep in range(2):
        out = pr.alloc_outputs(n_frames, winner=True)
        out["bev"].fill_(float("nan"))
        out["range"].fill_(float("nan"))
        pr.project(d, n_frames=n_frames, uniform_n=n_pts, out=out)
        torch.cuda.synchronize()
        assert pr.last_path == "fused"
        for k in ("bev", "range", "winner"):
            got = out[k].view((n_frames // distinct, distinct) + tuple(out[k].shape[1:]))
            same = torch.stack([(got[:, j] == ref[k][j]).flatten(1).all(dim=1) for j in range(distinct)], dim=1)
            assert bool(same.all()), "%s: frames %s differ" % (k, torch.nonzero(~same.flatten())[:16].flatten().tolist())
        del out
        torch.cuda.empty_cache()
    pr.close()


@pytest.mark.parametrize("stride", [0, 32, 64, 128, 256, 48])
def test_fused_firing_stride_is_only_a_hint(pkg, synth, stride):
    """sfpe_params.firing_stride tells the fused kernel's point pass how the sweep is laid out; right, wrong
    or absurd, every output must stay the same (ragged frames, firing-ordered and shuffled, three sensors)."""
    sweeps = [synth.make_sweep("hdl32", seed=7100), synth.make_sweep("hdl64", seed=7101, shuffle=True),
              synth.make_sweep("vls128", seed=7102, n_az=500, near_heavy=True), synth.make_sweep("hdl64", seed=7103, dropout=0.2),
              synth.make_sweep("hdl32", seed=7104, n_az=3), np.zeros((0, 4), np.float32)] * 26        # 156 frames > SM count
    a = _run_batch(pkg, sweeps, path=2, counts=False, firing_stride=stride)
    assert a["_path"] == "fused"
    for f in (0, 1, 2, 3, 4, 5, 150, 151, 152, 153):
        _check_frame(a, f, sweeps[f])
    for f in range(6, len(sweeps)):
        for k in ("range", "winner", "bev", "status", "dropped"):
            assert np.array_equal(a[k][f], a[k][f % 6]), (f, k)


@pytest.mark.parametrize("name", ["hdl32_seed0", "vls128_seed3_near", "edges", "hdl32_seed1_shuffled_dropout"])
def test_fused_path_vs_reference_golden(pkg, name):
    """Reference goldens straight through the fused kernel (float32 outputs): the images of the
    reference's own functions, rounded to float32, outside the oracle's tie mask."""
    g = golden(name)
    pts = np.ascontiguousarray(g["points"], dtype=np.float32)
    res = _run_batch(pkg, [pts, pts], path=2, counts=False, range_layout="chw")
    assert res["_path"] == "fused"
    tie = lo.front_view(_P(pts), strict=False)["tie_mask"]
    for f in (0, 1):
        for c, k in enumerate(("distance", "height", "intensity")):
            assert np.array_equal(res["range"][f][c][~tie], g[k].astype(np.float32)[~tie]), k
        if "bev_0.2" in g.files:
            _check_density(res["bev"][f], g["bev_0.2"])

@pytest.mark.parametrize("bev_dtype", ["f32", "f64"])
def test_bev_u8_transport_round_trip(pkg, synth, bev_dtype):
    """sfpe_bev_pack_u8 + sfpe_bev_unpack_host rebuild the density image bit for bit (saturated cells
    included), the class IS min(count, 63) of the oracle, and a foreign value raises the flag."""
    import importlib
    import torch
    proj = importlib.import_module(pkg.__name__ + ".projector")
    frames = [synth.make_sweep("hdl32", seed=21), synth.make_sweep("hdl64", seed=22, n_az=300)]
    dense = np.zeros((4000, 4), np.float32)          # 4000 points in a handful of cells: counts >= 63
    dense[:, 0] = 10.0 + 0.3 * np.random.default_rng(3).random(4000)
    dense[:, 1] = -5.0
    dense[:, 2] = -1.0
    frames.append(dense)
    pr = pkg.LidarProjector(device=0, bev_res=(0.2, 0.2))
    pts = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = np.zeros(len(frames) + 1, np.int64)
    offs[1:] = np.cumsum([len(f) for f in frames])
    out = pr.alloc_outputs(len(frames), bev_dtype=bev_dtype)
    pr.project(pts, n_frames=len(frames), frame_offsets=torch.from_numpy(offs).cuda(),
               max_n=max(len(f) for f in frames), out=out)
    codes, flag = pr.pack_bev_u8(out["bev"])
    torch.cuda.synchronize()
    assert int(flag.item()) == 0
    lut = pr.density_lut()
    np_dt = np.float32 if bev_dtype == "f32" else np.float64
    back = proj.unpack_bev_host(codes.cpu().numpy(), lut, dtype=np_dt)
    bev = out["bev"].cpu().numpy()
    assert back.dtype == bev.dtype and np.array_equal(back, bev)
    assert codes.max().item() == 63
    for f, p in enumerate(frames):
        counts = lo.bev_counts(p.astype(np.float64), (0.2, 0.2))
        assert np.array_equal(codes[f].cpu().numpy(), np.minimum(counts, 63)[::-1, ::-1].astype(np.uint8))
    bad = out["bev"].clone()
    bad.view(-1)[12345] = 1.0                         # not a density value
    _, flag = pr.pack_bev_u8(bad)
    assert int(flag.item()) == 1
    pr.close()


@pytest.mark.parametrize("transport", ["dense", "u8"])
def test_host_batch_pipeline_matches_device_batch(pkg, synth, transport):
    """HostBatchPipeline (pinned host in -> pinned host out, ragged last chunk, run twice) delivers the
    images of one device-resident project() call, whichever way the BEV crosses PCIe."""
    import torch
    n_frames, n = 7, 65536
    host = np.concatenate([synth.make_sweep("hdl32", seed=40 + i) for i in range(n_frames)])
    assert host.shape == (n_frames * n, 4)
    pin = torch.from_numpy(host).pin_memory()
    pr = pkg.LidarProjector(device=0, bev_res=(0.2, 0.2))
    ref = pr.alloc_outputs(n_frames)
    pr.project(pin.cuda(), n_frames=n_frames, uniform_n=n, out=ref)
    hb = pkg.HostBatchPipeline(pr, n_frames, n, chunk=3, bev_transport=transport)
    hb.h_bev.fill_(-1.0)
    hb.run(pin)
    hb.run(pin)
    hb.synchronize()
    assert np.array_equal(hb.h_range.numpy(), ref["range"].cpu().numpy())
    assert np.array_equal(hb.h_bev.numpy(), ref["bev"].cpu().numpy())
    assert np.array_equal(hb.h_status.numpy(), ref["status"].cpu().numpy())
    assert hb.d2h_bytes == n_frames * (32 * 1801 * 3 * 4 + pr.nbx * pr.nby * (1 if transport == "u8" else 4) + 4)
    hb.close()
    pr.close()


def test_fused_is_repeatable(pkg, synth):
    """The fused kernel's warps talk through flags in shared memory (claims, parked tiles, pacing window) and
    share scratch lists filled through atomics: ten launches over the same batch must give the same bits
    (winners are decided by (range, index), never by arrival order), also while another stream keeps the
    GPU busy."""
    import torch
    sweeps = [synth.make_sweep(("hdl32", "hdl64")[i % 2], seed=9800 + i, shuffle=(i % 5 == 0)) for i in range(160)]
    n = [len(s) for s in sweeps]
    offs = torch.from_numpy(np.concatenate([[0], np.cumsum(n)]).astype(np.int64)).cuda()
    d = torch.from_numpy(np.concatenate(sweeps)).cuda()
    pr = pkg.LidarProjector(device=0, path=2, firing_stride=32)
    first = None
    side = torch.cuda.Stream()
    junk = torch.empty(64 << 20, dtype=torch.float32, device="cuda")
    for rep in range(10):
        if rep % 2:
            with torch.cuda.stream(side):
                for _ in range(20):
                    junk.mul_(1.0001)
        out = pr.project(d, n_frames=len(sweeps), frame_offsets=offs, max_n=max(n), winner=True)
        torch.cuda.synchronize()
        got = {k: out[k].clone() for k in ("range", "winner", "bev", "status", "dropped")}
        if first is None:
            first = got
        else:
            for k in got:
                assert torch.equal(got[k], first[k]), (rep, k)
    pr.close()


def test_two_devices_in_one_process(pkg, synth):
    """One sfpe_ctx per device, driven from two threads of ONE process (one host thread per
    device): the fused kernel's opt-in to 227 KB of dynamic shared memory is a per-device attribute."""
    import threading
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    sweeps = [synth.make_sweep("hdl32", seed=9900 + i) for i in range(6)]
    allp = np.concatenate(sweeps)
    res, err = {}, []

    def work(dev):
        try:
            pr = pkg.LidarProjector(device=dev, path=2)
            d = torch.from_numpy(allp).to(torch.device("cuda", dev))
            with torch.cuda.device(dev):
                out = pr.project(d, n_frames=len(sweeps), uniform_n=65536, winner=True)
                torch.cuda.synchronize(dev)
            assert pr.last_path == "fused"
            res[dev] = {k: v.cpu().numpy() for k, v in out.items() if hasattr(v, "cpu")}
            pr.close()
        except Exception as e:      # surfaced in the main thread
            err.append((dev, repr(e)))

    th = [threading.Thread(target=work, args=(d,)) for d in (0, 1)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not err, err
    for dev in (0, 1):
        for f in (0, 5):
            _check_frame(res[dev], f, sweeps[f])


@pytest.mark.parametrize("transport", ["dense", "u8"])
def test_host_batch_pipeline_back_to_back_runs(pkg, synth, transport):
    """Two DIFFERENT batches issued back to back without a synchronize in between (what bench.py's e2e loop
    does): the device slots are reused across run() calls, so the second run's uploads and kernels must
    wait for the first run's kernels and downloads.  Each run has its own host outputs; both must be right."""
    import torch
    n_frames, n = 9, 65536
    pins, refs = [], []
    pr = pkg.LidarProjector(device=0, bev_res=(0.2, 0.2))
    for b in range(3):
        host = np.concatenate([synth.make_sweep("hdl32", seed=900 + 20 * b + i, shuffle=(i % 2 == b % 2)) for i in range(n_frames)])
        pins.append(torch.from_numpy(host).pin_memory())
        r = pr.alloc_outputs(n_frames)
        pr.project(pins[-1].cuda(), n_frames=n_frames, uniform_n=n, out=r)
        refs.append({k: r[k].cpu().numpy() for k in ("range", "bev", "status")})
    hb = pkg.HostBatchPipeline(pr, n_frames, n, chunk=2, bev_transport=transport)
    outs = [hb.host_outputs(), hb.alloc_host_outputs(), hb.alloc_host_outputs()]
    for rep in range(3):                      # several rounds: any missing dependency gets its chance
        for b in range(3):
            hb.run(pins[b], out=outs[b])
        hb.synchronize()
        for b in range(3):
            assert np.array_equal(outs[b]["range"].numpy(), refs[b]["range"]), (rep, b)
            assert np.array_equal(outs[b]["bev"].numpy(), refs[b]["bev"]), (rep, b)
            assert np.array_equal(outs[b]["status"].numpy(), refs[b]["status"]), (rep, b)
    hb.close()
    pr.close()


@pytest.mark.parametrize("cmap", ["gray", "jet"])
def test_host_batch_pipeline_rgb_is_the_reference_result_type(pkg, synth, cmap):
    """bev_transport="rgb": the host receives what the reference's generate_birds_eye_view returns -- the
    uint8 (nb, nb, 3) picture -- equal to the colormap stage applied to the float density image (the
    drop-in's own path), and the (32,1801,3) float32 range tensor; no float BEV image on the host."""
    import torch
    n_frames, n = 5, 65536
    host = np.concatenate([synth.make_sweep("hdl32", seed=950 + i) for i in range(n_frames)])
    pin = torch.from_numpy(host).pin_memory()
    pr = pkg.LidarProjector(device=0, bev_res=(0.2, 0.2))
    ref = pr.alloc_outputs(n_frames)
    pr.project(pin.cuda(), n_frames=n_frames, uniform_n=n, out=ref)
    hb = pkg.HostBatchPipeline(pr, n_frames, n, chunk=2, bev_transport="rgb", cmap=cmap)
    assert hb.h_bev is None and hb.h_bev_rgb.shape == (n_frames, pr.nbx, pr.nby, 3)
    hb.run(pin)
    hb.synchronize()
    colormap = load_pkg().colormap
    want = colormap.device_rgb(pr, ref["bev"], cmap, 0, 1500 * 0.2).cpu().numpy()
    assert np.array_equal(hb.h_bev_rgb.numpy(), want)
    assert np.array_equal(hb.h_range.numpy(), ref["range"].cpu().numpy())
    assert hb.d2h_bytes == n_frames * (32 * 1801 * 3 * 4 + pr.nbx * pr.nby + 4)
    hb.close()
    pr.close()


@pytest.mark.parametrize("cmap", ["gray", "jet"])
def test_generate_birds_eye_view_single_device_pass(pkg, synth, cmap):
    """generate_birds_eye_view (birds_eye_view_generator.py:41-54) keeps the float image on the device;
    its bytes are those of the colormap applied to the pinned density image, for every input form the
    reference's np.array(arrPoints) accepts (float32, float64, two columns, lists, no points)."""
    import importlib
    cm = importlib.import_module(pkg.__name__ + ".colormap")
    p32 = synth.make_sweep("hdl32", seed=77)
    cases = [p32, p32.astype(np.float64) * 1.0000001, p32[:5000, :2].astype(np.float64), p32[:300, :3].tolist(),
             np.zeros((0, 4), np.float32)]
    for res in [(0.2, 0.2), (0.5, 0.25)]:
        for pts in cases:
            img = pkg.generate_birds_eye_view(pts, None, res, cmap=cmap)
            dens = pkg.birds_eye_view_density(pts, res)
            want = cm.to_rgb_bytes(dens, cmap, vmin=0, vmax=1500 * res[0])
            assert img.dtype == np.uint8 and img.shape == dens.shape + (3,)
            assert np.array_equal(img, want)
    assert pkg.generate_birds_eye_view(p32, None, (0.2, 0.2), cmap="gray").max() > 0
    with pytest.raises(ValueError):
        pkg.generate_birds_eye_view(p32, None, (0.2, 0.2), cmap="viridis")
