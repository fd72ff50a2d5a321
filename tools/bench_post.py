#!/usr/bin/env python
"""Throughput of the range-scan consumers on one B200, inputs resident in HBM.

Prints one JSON line per kernel: algorithmic bytes (read + written once) / CUDA-event time, next to the
HBM peak from MEASURED_PEAKS.json when present.  Inputs are larger than L2 (1024 frames = 708 MB)."""
import argparse
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "sensor-fusion-pose-estimation-challenge_b200"


def timed(torch, fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import torch
    importlib.import_module(PKG)
    post = importlib.import_module(PKG + ".post")
    peak = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mp = json.load(f)
        peak = float(mp.get("hbm_gbs") or 0) or None
    except Exception:
        pass
    B, H, W, Cn = a.frames, 32, 1801, 3
    rs = torch.rand((B, H, W, Cn), device="cuda", dtype=torch.float32)
    frame_bytes = H * W * Cn * 4

    def line(name, ms, nbytes, extra=None):
        gbps = nbytes / ms / 1e6
        out = {"kernel": name, "frames": B, "ms": round(ms, 4), "algorithmic_GB": round(nbytes / 1e9, 4),
               "GB/s": round(gbps, 1), "frames/s": round(B / ms * 1e3)}
        if peak:
            out["frac_of_hbm_peak"] = round(gbps / peak, 3)
        out.update(extra or {})
        print(json.dumps(out))

    im = post.InvariantMask(rs[0].clone())
    ms = timed(torch, lambda: im.update(rs), a.reps)
    line("k_mask_update<float,4>", ms, B * frame_bytes + frame_bytes)

    out = torch.empty_like(rs)
    shift = torch.randint(0, W, (B,), device="cuda", dtype=torch.int32)
    ms = timed(torch, lambda: post.roll_width(rs, shift, out=out), a.reps)
    line("k_roll_width", ms, 2 * B * frame_bytes)

    synth = importlib.import_module(PKG + ".synth")
    dist, hgt, cen, box = synth.make_back_project_case(n_items=48, seed=1)
    rep = (B + 47) // 48
    scan = np.zeros(dist.shape + (3,), np.float32)
    scan[..., 0], scan[..., 1] = dist, hgt
    d_scan = torch.as_tensor(np.tile(scan, (rep, 1, 1, 1))[:B]).cuda()
    cen0 = torch.as_tensor(np.tile(cen, (rep, 1))[:B]).cuda()
    box0 = torch.as_tensor(np.tile(box, (rep, 1))[:B]).cuda()
    cen1, box1 = cen0.clone(), box0.clone()

    def bp():
        cen1.copy_(cen0); box1.copy_(box0)
        post.back_project_device(cen1, box1, range_scan=d_scan)
    ms = timed(torch, bp, a.reps)
    line("k_back_project<float> (+2 small copies, 2 allocations)", ms, 0, {"items/s": round(B / ms * 1e3)})

    # the post-network chain on a batch of FCN outputs (B, H*W, 26) float32: 6 MB per frame
    del rs, out, d_scan
    torch.cuda.empty_cache()
    Bn = min(B, 296)
    cases = [synth.make_detection_case(s) for s in range(8)]
    y = torch.as_tensor(np.stack([c[2] for c in cases])).cuda().repeat((Bn + 7) // 8, 1, 1)[:Bn].contiguous()
    sc = np.zeros((8, H, W, 3), np.float32)
    for i, c in enumerate(cases):
        sc[i, ..., 0], sc[i, ..., 1] = c[0], c[1]
    scan2 = torch.as_tensor(sc).cuda().repeat((Bn + 7) // 8, 1, 1, 1)[:Bn].contiguous()
    pred_bytes = Bn * H * W * 26 * 4

    def as_frames(name, ms, nbytes):
        gb = nbytes / ms / 1e6
        print(json.dumps({"kernel": name, "frames": Bn, "ms": round(ms, 4), "frames/s": round(Bn / ms * 1e3),
                          "network_output_GB/s": round(gb, 1), "frac_of_hbm_peak": round(gb / peak, 3) if peak else None}))
    ms = timed(torch, lambda: post.find_obstacle_device(y, H, W), a.reps)
    as_frames("k_find_obstacle<float> (+ scratch allocation)", ms, pred_bytes)
    obs = post.find_obstacle_device(y, H, W)
    cen = obs[:, 1:3].double().contiguous()
    box = obs[:, 3:7].double().contiguous()
    xyz, _ = post.back_project_device(cen.clone(), box.clone(), range_scan=scan2)
    bb2, c3 = obs[:, 3:7].contiguous(), xyz[:, :3].contiguous()
    ms = timed(torch, lambda: post.find_bbox_3d_device(y, bb2, c3, range_scan=scan2, cap=4096), a.reps)
    as_frames("k_find_bbox_3d<float,float> (+ scratch allocation)", ms, pred_bytes)
    ms = timed(torch, lambda: post.locate_obstacles(y, scan2), a.reps)
    as_frames("locate_obstacles: find_obstacle -> back_project -> find_bbox_3d (+ torch glue)", ms, 2 * pred_bytes)


if __name__ == "__main__":
    main()
