"""Repeats the ragged fused-path scenario of tests/test_gpu_parity.py::test_fused_path_ragged and reports
where the BEV density / range scan differ from the oracle (frame, number of cells, rows)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np
from conftest import load_pkg
from oracle import lidar_oracle as lo
import torch
pkg = load_pkg(); synth = pkg.synth
sweeps = [synth.make_sweep("hdl32", seed=80), synth.make_sweep("hdl32", seed=81, dropout=0.3, shuffle=True),
          np.zeros((0, 4), np.float32), synth.make_sweep("hdl64", seed=82), synth.make_sweep("hdl32", seed=83, n_az=1),
          synth.make_sweep("vls128", seed=84, near_heavy=True)]
want = []
for s in sweeps:
    P = np.zeros((len(s), 5)); P[:, :4] = s
    want.append((lo.bev_density(P, (0.2, 0.2), 120).astype(np.float32), lo.front_view(P, strict=False)["winner"].astype(np.int32)))
allp = np.concatenate(sweeps); offs = np.zeros(len(sweeps) + 1, np.int64); offs[1:] = np.cumsum([len(s) for s in sweeps])
d = torch.from_numpy(allp).cuda(); o = torch.from_numpy(offs).cuda()
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 10):
    pr = pkg.LidarProjector(device=0, path=2)
    out = pr.project(d, n_frames=len(sweeps), frame_offsets=o, max_n=max(len(s) for s in sweeps), winner=True)
    torch.cuda.synchronize()
    bev = out["bev"].cpu().numpy(); win = out["winner"].cpu().numpy()
    pr.close()
    for f in range(len(sweeps)):
        bad = np.argwhere(bev[f] != want[f][0])
        wbad = int((win[f] != want[f][1]).sum())
        if len(bad) or wbad:
            rows = np.unique(bad[:, 0]) if len(bad) else []
            print("rep %d frame %d: %d BEV cells differ (rows %s..%s, %d rows), got nonzero %d want nonzero %d; winner diffs %d" % (
                rep, f, len(bad), rows[0] if len(rows) else "-", rows[-1] if len(rows) else "-", len(rows),
                int((bev[f][tuple(bad.T)] != 0).sum()) if len(bad) else 0, int((want[f][0][tuple(bad.T)] != 0).sum()) if len(bad) else 0, wbad))
            if len(bad): print("   first:", [(int(r), int(c), float(bev[f][r, c]), float(want[f][0][r, c])) for r, c in bad[:6]])
print("done")
