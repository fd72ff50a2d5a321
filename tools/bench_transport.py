#!/usr/bin/env python
"""Host side of the one-byte BEV transport: sfpe_bev_unpack_host throughput by thread count (no GPU
needed), and -- with a GPU -- HostBatchPipeline end to end by transport, chunk and thread count.
One JSON line per measurement."""
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
    ap.add_argument("--sweeps", type=int, default=1024)
    ap.add_argument("--chunk-frames", type=int, default=32)
    ap.add_argument("--no-gpu", action="store_true")
    a = ap.parse_args()
    pkg = importlib.import_module(PKG)
    proj = importlib.import_module(PKG + ".projector")
    cells = 1199 * 1199
    n = a.chunk_frames * cells
    rng = np.random.default_rng(0)
    codes = np.zeros(n, np.uint8)
    idx = rng.integers(0, n, a.chunk_frames * 20000)       # ~20 k occupied cells per sweep, like an HDL-32E sweep
    codes[idx] = rng.integers(1, 64, idx.size)
    lut = np.array([0.0] + [255.0 * min(np.log(k + 1.0) / np.log(64.0), 1.0) for k in range(1, 65)])
    out = np.empty(n, np.float32)
    for threads in (1, 2, 4, 8, 12, 16, 24, 32):
        if threads > 2 * (os.cpu_count() or 1):
            break
        proj.unpack_bev_host(codes, lut, out=out, threads=threads)
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            proj.unpack_bev_host(codes, lut, out=out, threads=threads)
        dt = (time.perf_counter() - t0) / reps
        print(json.dumps({"what": "sfpe_bev_unpack_host", "frames": a.chunk_frames, "threads": threads, "ms": round(dt * 1e3, 3),
                          "GB/s_written": round(n * 4 / dt / 1e9, 1), "sweeps/s": round(a.chunk_frames / dt)}), flush=True)
    if a.no_gpu:
        return
    import torch
    n_pts = 65536
    pool = 64
    host = np.concatenate([pkg.synth.make_sweep("hdl32", seed=i) for i in range(pool)])
    host = np.concatenate([host] * (a.sweeps // pool))
    pin = torch.from_numpy(host).pin_memory()
    pr = pkg.LidarProjector(device=0)
    for transport, chunk, threads, workers in [("dense", 32, 0, 1), ("dense", 64, 0, 1), ("u8", 32, 0, 1), ("u8", 64, 0, 1),
                                               ("u8", 128, 0, 1), ("u8", 32, 8, 2), ("u8", 64, 8, 2), ("u8", 64, 16, 2),
                                               ("u8", 16, 8, 2), ("u8", 32, 6, 3)]:
        hb = pkg.HostBatchPipeline(pr, a.sweeps, n_pts, chunk=chunk, bev_transport=transport, unpack_threads=threads,
                                   unpack_workers=workers)
        hb.run(pin)
        hb.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            hb.run(pin)
        hb.synchronize()
        dt = (time.perf_counter() - t0) / 3
        print(json.dumps({"what": "HostBatchPipeline", "transport": transport, "chunk": chunk, "threads": threads, "workers": workers,
                          "sweeps/s": round(a.sweeps / dt), "d2h_GB/s": round(hb.d2h_bytes / dt / 1e9, 1)}), flush=True)
        del hb
    pr.close()


if __name__ == "__main__":
    main()
