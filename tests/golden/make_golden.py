"""Generate tests/golden/*.npz by running the REFERENCE's own functions.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
Each fixture stores the float32 input points and what the unmodified reference
returned for them: `generate_lidar_2d_front_view(points, cmap=None)`'s three
`*_float` images (modules/lidar/process/extract_rosbag_lidar.py:80-86) and the
pre-colormap density image of `generate_birds_eye_view`
(modules/lidar/common/birds_eye_view_generator.py:41-46) at res 0.2 m and 2 m.
`raises_*.npz` hold point sets for which the reference raises IndexError.
"""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import reference_loader as rl  # noqa: E402

_spec = importlib.util.spec_from_file_location(
    "synth", os.path.join(ROOT, "sensor-fusion-pose-estimation-challenge_b200", "synth.py"))
synth = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(synth)


def polar(rho, az_deg, el_deg, inten):
    az, el = np.deg2rad(az_deg), np.deg2rad(el_deg)
    return [rho * np.cos(el) * np.cos(az), rho * np.cos(el) * np.sin(az), rho * np.sin(el), inten]


def edge_points():
    """Hand-made points on the semantics SURVEY.md 8(a) lists."""
    p = []
    # axes / diagonals: exact atan2 special values
    for x, y in [(5, 0), (-5, 0.0), (-5, -0.0), (0, 5), (0, -5), (3, 3), (3, -3), (-3, 3), (-3, -3),
                 (7, 1e-30), (-7, 1e-30), (-7, -1e-30)]:
        p.append([x, y, -0.5, 10 + len(p)])
    # el = 0 (row 23 pre-flip), el near the band edges, truncation band (-32, -30.67),
    # negative-index wrap (el < -32 deg -> top rows)
    for i, el in enumerate([0.0, 10.6, 11.8, -30.6, -30.9, -31.9, -32.5, -40.0, -60.0, -73.0]):
        p.append(polar(12.0 + i, 17.0 + 3 * i, el, 100 + i))
    # same cell, different ranges -> nearest wins; exact duplicates -> tie (lowest index)
    for rho in (30.0, 10.0, 20.0):
        p.append(polar(rho, 45.05, -5.0, rho))
    p.append(polar(10.0, 45.05, -5.0, 77.0))
    # near-equal ranges in one cell (differ below float32 resolution of rho)
    p.append([8.0, 1.0, -0.25, 1.0])
    p.append([8.0, np.nextafter(np.float32(1.0), np.float32(2.0)), -0.25, 2.0])
    # BEV: points on bin edges, on the last edge, outside, and exactly 0.0
    for x, y in [(0.0, 1.0), (0.2, 0.2), (-120.0, -120.0), (119.8, 119.8), (119.80000305175781, 1.0),
                 (119.9, 0.1), (120.0, 0.0), (-120.1, 0.0), (60.0, -60.0), (-0.2, 0.4)]:
        p.append([x, y, -1.0, 5.0])
    # dense cell: 70 points in one BEV cell (density saturates at 63) spread over range cells
    for k in range(70):
        p.append([10.05 + 1e-3 * k, 4.05, -1.7 + 0.02 * k, k])
    return np.array(p, dtype=np.float32)


def bev_edge_points():
    """BEV-only set (x = y = 0 has no defined elevation, so it cannot go through the range scan)."""
    xs = [0.0, -0.0, 0.2, 0.4, 0.6000000238418579, -120.0, 119.8, 119.80000305175781, 119.9, 120.0,
          -120.00001, 60.0, -60.0, 1e-12, -1e-12, np.nan, np.inf, -np.inf, 2.0, 118.0, 117.99999]
    p = [[x, y, 0.0, 1.0] for x in xs for y in xs]
    return np.array(p, dtype=np.float32)


def run_case(name, pts32, bev_res=((0.2, 0.2), (2, 2)), front=True):
    P = synth.as_reference_array(pts32)
    out = {"points": pts32}
    if front:
        fv = rl.front_view(P)
        out.update(distance=fv["distance_float"], height=fv["height_float"], intensity=fv["intensity_float"])
    for res in bev_res:
        out["bev_%g" % res[0]] = rl.bev_density(P, res, max_range=120)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(name, len(pts32), "->", os.path.getsize(path) // 1024, "KiB")


def raises_case(name, pts32):
    P = synth.as_reference_array(pts32)
    try:
        with np.errstate(all="ignore"):
            rl.front_view(P)
    except IndexError:
        np.savez_compressed(os.path.join(HERE, name + ".npz"), points=pts32)
        print(name, "reference raises IndexError (recorded)")
        return
    raise SystemExit("%s: reference did NOT raise" % name)


def back_project_case(name, seed):
    """Reference back_project_2D_2_3D (train/predict.py:230-293) on a synthetic case; inputs kept as float32."""
    dist, hgt, cen, box = synth.make_back_project_case(seed=seed)
    cen_out, box_out = cen.copy(), box.copy()
    xyz = rl.back_project_fn()(cen_out, box_out, dist.astype(np.float64), hgt.astype(np.float64))
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, distance=dist, height=hgt, centroids=cen, bboxes=box, centroids_out=cen_out,
                        bboxes_out=box_out, xyz=xyz)
    print(name, len(cen), "items ->", os.path.getsize(path) // 1024, "KiB")


def bbox_decode_case(name, seeds):
    """Reference find_bbox_3d (train/predict.py:94-199) on synthetic cases.  Only the window neighbourhood of
    y_pred is kept (float16-free: float32 as generated); the test rebuilds the full array around it."""
    out = {}
    for k, seed in enumerate(seeds):
        dist, hgt, y, bb, c3 = synth.make_bbox_decode_case(seed=seed)
        pred, box = rl.find_bbox_3d_fn()(dist.astype(np.float64), hgt.astype(np.float64), y, bb, c3)
        out.update({"seed%d" % k: seed, "bbox%d" % k: np.array(bb), "centroid%d" % k: c3, "pred%d" % k: pred,
                    "box%d" % k: np.zeros((8, 3)) if box is None else box, "none%d" % k: box is None})
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, n=len(seeds), **out)
    print(name, len(seeds), "cases ->", os.path.getsize(path) // 1024, "KiB")


def obstacle_case(name, seeds):
    """Reference find_obstacle (train/predict.py:33-81) on synthetic 2-class outputs: found, centroid, box, area."""
    rows = []
    for seed in seeds:
        c, b, a = rl.find_obstacle_fn()(synth.make_obstacle_case(seed=seed), (32, 1801, 3))
        rows.append([seed, 0, 0, 0, 0, 0, 0, 0, 0] if c is None else
                    [seed, 1, c[0], c[1], b[0][0], b[0][1], b[1][0], b[1][1], a])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), rows=np.array(rows, dtype=np.int64))
    print(name, len(seeds), "cases")


def main():
    back_project_case("back_project_seed0", 0)
    obstacle_case("find_obstacle", list(range(12)))
    bbox_decode_case("bbox_decode", [0, 1, 2, 3])
    run_case("hdl32_seed0", synth.make_sweep("hdl32", seed=0))
    run_case("hdl32_seed1_shuffled_dropout", synth.make_sweep("hdl32", seed=1, shuffle=True, dropout=0.05))
    run_case("hdl64_seed2", synth.make_sweep("hdl64", seed=2), bev_res=((0.2, 0.2),))
    run_case("vls128_seed3_near", synth.make_sweep("vls128", seed=3, near_heavy=True), bev_res=((2, 2),))
    run_case("small_seed4", synth.make_sweep("hdl32", seed=4, n_az=97, dropout=0.1), bev_res=((0.2, 0.2), (2, 2), (0.5, 0.5)))
    run_case("edges", edge_points(), bev_res=((0.2, 0.2), (2, 2)))
    run_case("bev_edges", bev_edge_points(), bev_res=((0.2, 0.2), (2, 2), (0.1, 0.1)), front=False)
    base = synth.make_sweep("hdl32", seed=5, n_az=16)
    raises_case("raises_high_elevation", np.vstack([base, np.array([polar(9.0, 10.0, 12.5, 1.0)], np.float32)]))
    raises_case("raises_origin", np.vstack([base, np.zeros((1, 4), np.float32)]))
    raises_case("raises_nan", np.vstack([base, np.array([[np.nan, 1, 1, 1]], np.float32)]))
    raises_case("raises_low_elevation", np.vstack([base, np.array([polar(9.0, 10.0, -80.0, 1.0)], np.float32)]))


if __name__ == "__main__":
    main()
