#!/usr/bin/env python
"""C4: per-sweep latency of the streaming entry on one B200.

One PointCloud2-shaped blob (point_step 32, Velodyne field layout) per call: pinned host buffer
-> H2D -> device decode -> range scan -> D2H of the (32,1801,3) float32 tensor.  Prints one JSON
line with p50 / p99 / mean latency over N sweeps, next to the reference's CPU path for the same
message (per-point python decode as in lidar_predict.py:64-65 + the numpy front view port).
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "sensor-fusion-pose-estimation-challenge_b200"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweeps", type=int, default=1000)
    ap.add_argument("--distinct", type=int, default=32)
    ap.add_argument("--sensor", default="hdl32")
    ap.add_argument("--cpu-sweeps", type=int, default=3)
    a = ap.parse_args()
    pkg = importlib.import_module(PKG)
    synth = pkg.synth
    blobs = [synth.make_pointcloud2_blob(synth.make_sweep(a.sensor, seed=s)) for s in range(a.distinct)]
    sv = pkg.StreamingFrontView(device=0)
    for b in blobs[:8]:
        sv(b)
    lat = []
    for i in range(a.sweeps):
        b = blobs[i % len(blobs)]
        t0 = time.perf_counter()
        sv(b)
        lat.append(time.perf_counter() - t0)
    lat = np.array(lat) * 1e3
    line = {"metric": "streaming per-sweep latency (PointCloud2 blob -> (32,1801,3) tensor on host)", "unit": "ms",
            "sweeps": a.sweeps, "points_per_sweep": int(len(blobs[0]) // synth.PC2_POINT_STEP),
            "p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)), "mean": float(lat.mean()),
            "h2d_bytes": int(len(blobs[0])), "d2h_bytes": int(32 * 1801 * 3 * 4), "path": sv.pr.last_path}
    sv.close()
    if a.cpu_sweeps:
        import struct
        from oracle import lidar_oracle as lo
        t = []
        for b in blobs[: a.cpu_sweeps]:
            t0 = time.perf_counter()
            raw = bytes(b)
            step = synth.PC2_POINT_STEP
            ox, oy, oz, oi = synth.PC2_OFFSETS
            pts = [struct.unpack_from("<fff", raw, k * step + ox) + struct.unpack_from("<f", raw, k * step + oi)
                   for k in range(len(raw) // step)]          # read_points + list(): one python tuple per point
            P = np.array(pts)
            lo.port_generate_lidar_2d_front_view(P)
            t.append(time.perf_counter() - t0)
        line["cpu_reference_ms"] = float(np.median(t) * 1e3)
        line["cpu_reference_note"] = "per-point python decode (as read_points + list) + numpy front view port, 1 core"
    print(json.dumps(line))


if __name__ == "__main__":
    main()
