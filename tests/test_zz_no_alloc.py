"""sfpe_range_scan_static: the entry point that allocates nothing (include/sfpe.h).

Collected last on purpose (the file name sorts after the others): its scratch is one static device
array shared by every call in the process, so it must pass alone in a fresh process and after the
other tests.  Called through the C ABI with ctypes; checked against the existing entry point,
sfpe_project_batch, bit for bit (integer winners and the three range images: the tolerance of the
parity tests for this work is exact), on calls in a row with different inputs."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_pkg, load_sub

pytestmark = pytest.mark.gpu


def _static_call(lib_mod, params, pts, n_frames, offs, max_n, uniform_n, out, layout, stream):
    import torch
    lib = lib_mod.load()
    desc = lib_mod.Points()
    desc.data = pts.data_ptr()
    desc.stride_bytes = pts.stride(0) * pts.element_size()
    desc.dtype = lib_mod.SFPE_F32 if pts.dtype == torch.float32 else lib_mod.SFPE_F64
    e = pts.element_size()
    desc.off_x, desc.off_y, desc.off_z, desc.off_i = 0, e, 2 * e, 3 * e
    o = lib_mod.Outputs()
    o.range = out["range"].data_ptr()
    o.range_dtype = lib_mod.SFPE_F32 if out["range"].dtype == torch.float32 else lib_mod.SFPE_F64
    o.range_layout = lib_mod.RANGE_HWC if layout == "hwc" else lib_mod.RANGE_CHW
    o.range_winner = out["winner"].data_ptr()
    o.status = out["status"].data_ptr()
    o.dropped = out["dropped"].data_ptr()
    return lib.sfpe_range_scan_static(0, C.byref(params), C.byref(desc),
                                      offs.data_ptr() if offs is not None else None,
                                      int(uniform_n), int(max_n), int(n_frames), C.byref(o), C.c_void_p(int(stream)))


def _outputs(torch, n_frames, H, W, layout, dtype):
    shape = (n_frames, H, W, 3) if layout == "hwc" else (n_frames, 3, H, W)
    return {"range": torch.full(shape, -7.0, dtype=dtype, device="cuda"),
            "winner": torch.full((n_frames, H, W), -9, dtype=torch.int32, device="cuda"),
            "status": torch.full((n_frames,), 77, dtype=torch.int32, device="cuda"),
            "dropped": torch.full((n_frames,), 77, dtype=torch.int32, device="cuda")}


def _same(a, b):
    for k in ("range", "winner", "status", "dropped"):
        assert np.array_equal(a[k].cpu().numpy(), b[k].cpu().numpy()), k


def test_static_range_scan_matches_project_batch_on_calls_in_a_row():
    import torch
    assert torch.cuda.is_available()
    pkg, lib_mod, synth = load_pkg(), load_sub("_lib"), load_sub("synth")
    params = lib_mod.default_params()
    H, W, _, _ = lib_mod.grid_dims(params)
    assert H * W <= lib_mod.STATIC_MAX_CELLS
    stream = torch.cuda.current_stream(0).cuda_stream

    # call 1: a ragged batch with more frames than the array holds at once (4 per pass on this grid),
    #         dropout and two points the reference would refuse; call 2: one short uniform sweep whose
    #         empty cells would show the winners of call 1 if the scratch were not cleared; call 3:
    #         float64 points, planar layout, forced key collisions
    sweeps = [synth.make_sweep("hdl32", seed=40 + i, dropout=0.05 * (i % 3), shuffle=bool(i & 1)) for i in range(11)]
    sweeps[3] = sweeps[3].copy()
    sweeps[3][5] = (0.0, 0.0, 0.0, 1.0)
    sweeps[3][9] = (np.nan, 1.0, 0.0, 1.0)
    sweeps[7] = sweeps[7][:0]
    cases = []
    allp = np.concatenate(sweeps)
    offs = np.zeros(len(sweeps) + 1, np.int64)
    offs[1:] = np.cumsum([len(s) for s in sweeps])
    cases.append(dict(pts=torch.from_numpy(allp).cuda(), n=len(sweeps), offs=torch.from_numpy(offs).cuda(),
                      max_n=max(len(s) for s in sweeps), layout="hwc", dtype=torch.float32, prefix=0))
    short = synth.make_sweep("hdl32", seed=99, n_az=300)
    cases.append(dict(pts=torch.from_numpy(short).cuda(), n=1, offs=None, max_n=len(short), layout="hwc",
                      dtype=torch.float32, prefix=0))
    two = np.concatenate([synth.make_sweep("hdl32", seed=7), synth.make_sweep("hdl32", seed=8)]).astype(np.float64)
    two[:, :3] *= 1.0 + 1e-11     # not float32 representable
    cases.append(dict(pts=torch.from_numpy(two).cuda(), n=2, offs=None, max_n=len(two) // 2, layout="chw",
                      dtype=torch.float64, prefix=12))

    for rep in range(2):                      # the whole sequence twice: every call follows another one
        for c in cases:
            params.key_prefix_bits = c["prefix"]
            got = _outputs(torch, c["n"], H, W, c["layout"], c["dtype"])
            rc = _static_call(lib_mod, params, c["pts"], c["n"], c["offs"], c["max_n"],
                              0 if c["offs"] is not None else c["max_n"], got, c["layout"], stream)
            assert rc == 0, lib_mod.load().sfpe_strerror(rc)
            torch.cuda.synchronize()
            for path in (1, 0):               # the existing entry point: generic kernels; its own choice (fused)
                pr = pkg.LidarProjector(device=0, path=path, key_prefix_bits=c["prefix"])
                want = _outputs(torch, c["n"], H, W, c["layout"], c["dtype"])
                want["_layout"] = c["layout"]
                if c["offs"] is not None:
                    pr.project(c["pts"], n_frames=c["n"], frame_offsets=c["offs"], max_n=c["max_n"], out=want)
                else:
                    pr.project(c["pts"], n_frames=c["n"], uniform_n=c["max_n"], out=want)
                torch.cuda.synchronize()
                pr.close()
                _same(got, want)
            assert (int(got["status"][3]) != 0) if c["n"] > 3 else (int(got["status"][0]) == 0)

    # no device memory was taken by the calls themselves: same inputs, outputs allocated beforehand
    c = cases[0]
    got = _outputs(torch, c["n"], H, W, c["layout"], c["dtype"])
    params.key_prefix_bits = 0
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info(0)[0]
    for _ in range(3):
        assert _static_call(lib_mod, params, c["pts"], c["n"], c["offs"], c["max_n"], 0, got, c["layout"], stream) == 0
    torch.cuda.synchronize()
    assert abs(torch.cuda.mem_get_info(0)[0] - before) < (2 << 20)   # device allocations come in 2 MiB steps


def test_static_range_scan_refuses_grids_beyond_its_array():
    import torch
    lib_mod = load_sub("_lib")
    lib = lib_mod.load()
    params = lib_mod.default_params()
    params.res_h_deg = 0.02                   # W = 18001, H*W = 576 032 > 262 144
    H, W, _, _ = lib_mod.grid_dims(params)
    assert H * W > lib_mod.STATIC_MAX_CELLS
    pts = torch.zeros((4, 4), dtype=torch.float32, device="cuda")
    out = _outputs(torch, 1, 1, 1, "hwc", torch.float32)
    rc = _static_call(lib_mod, params, pts, 1, None, 4, 4, out, "hwc", 0)
    assert rc == lib_mod.ERR_TOO_LARGE
    assert b"SFPE_STATIC_MAX_CELLS" in lib.sfpe_strerror(rc)
    # BEV outputs have no scratch in this entry point
    params = lib_mod.default_params()
    desc, o = lib_mod.Points(), lib_mod.Outputs()
    desc.data, desc.stride_bytes, desc.off_y, desc.off_z, desc.off_i = pts.data_ptr(), 16, 4, 8, 12
    o.bev = pts.data_ptr()
    assert lib.sfpe_range_scan_static(0, C.byref(params), C.byref(desc), None, 4, 4, 1, C.byref(o), None) == -1
