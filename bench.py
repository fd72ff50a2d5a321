#!/usr/bin/env python
"""bench.py -- lidar sweeps/s of the BEV + range-scan projection (BASELINE.json metric).

Workload (BASELINE.json configs[1]): a batch of 1024 synthetic HDL-32E sweeps (32 x 2048 =
65 536 points each, float32 x,y,z,intensity), reference grid (32 x 1801 range scan, +-120 m BEV
at 0.2 m -> 1199 x 1199), one B200.  One "step" = one pass of the hot path over the whole batch.
Each rank owns its own batch (frame sharding, no collective): weak scaling.

  python bench.py [--gpus N --steps K --warmup W]            our CUDA path
  python bench.py --impl reference [...]                      the reference's numpy algorithm on host cores
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "sensor-fusion-pose-estimation-challenge_b200"

H, W = 32, 1801


def algorithmic_bytes(n_pts, nbx, nby, c_bev=1):
    """float4 read once + fp32 range tensor written once + fp32 BEV written once."""
    return 16 * n_pts + 4 * 3 * H * W + 4 * c_bev * nbx * nby


def measured_traffic(path_name, sweeps):
    """dram__bytes_read+write per launch of the dominant kernel from the committed ncu capture
    (profiles/traffic.json, bytes per sweep measured on a 296-sweep launch), scaled to this launch."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        t = json.load(open(p)).get(path_name)
        return None if not t else float(t["dram_bytes_per_sweep"]) * sweeps
    except (OSError, ValueError, KeyError):
        return None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------ CPU baseline
def _cpu_impl():
    """(kind, front_view(P), bev(P, res, vectorised)).  kind "reference": the reference's own files
    (process/extract_rosbag_lidar.py, common/birds_eye_view_generator.py, process/globals.py), loaded unmodified
    through oracle/reference_loader.py from /root/reference or from oracle/_ref (the copy oracle/make_ref.py made
    at build time, which travels to the GPU box).  kind "port": the op-for-op numpy port, only when neither
    exists.  The vectorised-normalize variant is always the port (the reference has no such form)."""
    from oracle import lidar_oracle as lo
    from oracle import reference_loader as rl
    vec = lambda P, res: lo.port_generate_birds_eye_view(P, res, 120, python_loop=False)
    if rl.available():
        return "reference", rl.front_view, (lambda P, res: rl.bev_density(P, res, 120)), vec
    return "port", lo.port_generate_lidar_2d_front_view, (lambda P, res: lo.port_generate_birds_eye_view(P, res, 120, python_loop=True)), vec


def _cpu_worker(args):
    """Rasterises its sweeps with the reference's CPU code on one host core; returns (kind, seconds per sweep)."""
    seeds, sensor, bev_res, vectorised = args
    import importlib.util
    spec = importlib.util.spec_from_file_location("synth", os.path.join(ROOT, PKG, "synth.py"))   # the package itself needs the .so
    synth = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synth)
    kind, front, bev, bev_vec = _cpu_impl()
    pts = [synth.as_reference_array(synth.make_sweep(sensor, s)) for s in seeds]
    out = []
    for P in pts:
        t0 = time.perf_counter()
        front(P)
        (bev_vec if vectorised else bev)(P, bev_res)
        out.append(time.perf_counter() - t0)
    return kind, out


class CpuReference:
    """The reference's CPU code on all host cores: one persistent pool of worker processes (spawn), each
    rasterising its own disjoint sweeps.  pass_(n) rasterises n sweeps and returns wall seconds."""

    def __init__(self, sensor, bev_res, cores=None):
        import multiprocessing as mp
        self.cores = cores or os.cpu_count() or 1
        self.sensor, self.bev_res = sensor, bev_res
        self.pool = mp.get_context("spawn").Pool(self.cores)
        r = self.pool.map(_cpu_worker, [([i], sensor, bev_res, True) for i in range(self.cores)])     # imports + warm-up
        self.kind = r[0][0]

    def pass_(self, n_sweeps, seed0=0, vectorised=False):
        chunks = [list(range(seed0 + i, seed0 + n_sweeps, self.cores)) for i in range(self.cores)]
        chunks = [c for c in chunks if c]
        t0 = time.perf_counter()
        self.pool.map(_cpu_worker, [(c, self.sensor, self.bev_res, vectorised) for c in chunks])
        return time.perf_counter() - t0

    def single_core(self, n=5, seed0=500, vectorised=False):
        """n sweeps one after the other in ONE worker process: per-sweep milliseconds (best, median)."""
        _, t = self.pool.apply(_cpu_worker, ((list(range(seed0, seed0 + n)), self.sensor, self.bev_res, vectorised),))
        return {"best_of_%d" % n: 1e3 * min(t), "median": 1e3 * float(np.median(t))}

    def describe(self):
        return ("the reference's own files (extract_rosbag_lidar.py, birds_eye_view_generator.py, globals.py; "
                "oracle/_ref)" if self.kind == "reference" else "numpy port of the reference (oracle/_ref absent)") + \
               ": 3x lidar_2d_front_view + histogram2d + per-cell python normalize loop, colormap stubbed"

    def close(self):
        self.pool.close()
        self.pool.join()


def cpu_baseline_block(a, target_s=15.0):
    """(i) one core, best of 5 and median per sweep; (ii) all host cores over disjoint sweeps
    (the headline `value`); and the same two with the python per-cell `normalize` loop replaced by its vectorised
    equivalent, which shows how much of the CPU time is that loop."""
    ref = CpuReference(a.sensor, (a.bev_res, a.bev_res))
    one = ref.single_core(5)
    one_vec = ref.single_core(5, vectorised=True)
    n = ref.cores * 2
    dt = ref.pass_(n)
    n = int(max(ref.cores * 2, min(4096, n * target_s / max(dt, 1e-3))))
    dt = ref.pass_(n, seed0=1000)
    nv = ref.cores * 16
    dtv = ref.pass_(nv, seed0=3000, vectorised=True)
    ref.close()
    return {"value": n / dt, "unit": "sweeps/s", "cores": ref.cores, "kind": ref.kind,
            "sample": "%d sweeps of the same workload in %d processes (%.1f s); %s" % (n, ref.cores, dt, ref.describe()),
            "single_core_ms_per_sweep": one,
            "vectorised_normalize": {"single_core_ms_per_sweep": one_vec, "all_cores_sweeps_per_s": nv / dtv,
                                     "note": "numpy port with the python normalize loop "
                                             "(birds_eye_view_generator.py:31-37) replaced by array operations"}}


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref: its three files,
    unmodified) on all host cores, same metric / config.  Each step is a bounded sample of the workload
    (4 sweeps per core)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ref = CpuReference(a.sensor, (a.bev_res, a.bev_res))
    cores = ref.cores
    per_step = cores * 4
    times = []
    for i in range(a.warmup + a.steps):
        dt = ref.pass_(per_step, seed0=i * per_step)
        if i >= a.warmup:
            times.append(dt)
    desc = ref.describe()
    ref.close()
    total = sum(times)
    val = per_step * len(times) / total
    cfg = workload_config(a)
    line = {
        "impl": "reference", "metric": "lidar sweeps/s (BEV+range projection)", "value": val, "unit": "sweeps/s",
        "points_per_s": val * a.n_pts,
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": val, "unit": "sweeps/s", "cores": cores, "kind": ref.kind,
                         "sample": "%d sweeps per step x %d steps of the same workload in %d processes; %s"
                                   % (per_step, len(times), cores, desc)},
        "e2e": {"value": val, "unit": "sweeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def workload_config(a):
    return {"workload": "batch of %d synthetic %s sweeps (%d pts each), range scan 32x1801x3 + BEV density "
                        "+-120 m @ %g m" % (a.sweeps, a.sensor, a.n_pts, a.bev_res),
            "sweeps_per_gpu": a.sweeps, "points_per_sweep": a.n_pts, "bev_res_m": a.bev_res,
            "l2_policy": "inputs (%.2f GB) and outputs larger than the 126 MB L2; no flush needed" % (a.sweeps * a.n_pts * 16 / 1e9),
            "parallelism": "frame sharding, %d rank(s), no collective" % a.gpus}


# ------------------------------------------------------------------------------ GPU arm
def run_ours(a):
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module(PKG)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    # synthetic batch: distinct seeds per rank; generated on the host, resident in HBM before timing
    pool = min(a.sweeps, a.distinct)
    host = np.concatenate([pkg.synth.make_sweep(a.sensor, seed=rank * a.sweeps + i) for i in range(pool)])
    reps = (a.sweeps + pool - 1) // pool
    host_all = np.concatenate([host] * reps)[: a.sweeps * a.n_pts] if reps > 1 else host
    pin_in = torch.from_numpy(host_all).pin_memory()
    d_pts = pin_in.to(dev, non_blocking=True)
    pr = pkg.LidarProjector(device=local, bev_res=(a.bev_res, a.bev_res), path=a.path, debug_flags=a.debug_flags,
                            firing_stride=pkg.synth.SENSORS[a.sensor][0])   # sweeps are stored firing by firing
    out = pr.alloc_outputs(a.sweeps)
    torch.cuda.synchronize()

    def step():
        pr.project(d_pts, n_frames=a.sweeps, uniform_n=a.n_pts, out=out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    l0 = pr.launch_count
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
    barrier()
    ev[0].record()
    for i in range(a.steps):
        step()
        ev[i + 1].record()
    barrier()
    launches = pr.launch_count - l0
    path_timed = pr.last_path
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(a.steps)]
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())

    # ---- e2e: same metric through the host-buffer API, H2D + D2H inside the timed region
    e2e = None
    if not a.no_e2e:
        e2e = run_e2e(a, pkg, pr, pin_in, dev, world)

    # ---- optional: per-phase SM cycles of the fused kernel (separate untimed steps)
    phases = None
    if a.phases and rank == 0:
        pr.phase_profile(True)
        for _ in range(2):
            step()
        torch.cuda.synchronize()
        cyc = pr.phase_cycles()
        pr.phase_profile(False)
        tot = float(sum(cyc.values())) or 1.0
        phases = {k: round(v / tot, 4) for k, v in cyc.items() if v}
        phases["sm_cycles_per_sweep"] = tot / (2 * a.sweeps)

    # ---- the timed launch is checked: K random frames of the batch are rasterised again by the generic
    # (global-atomics, fp64) path and must equal the slices of the timed output bit for bit (every rank)
    verified = verify_frames(a, pkg, path_timed, d_pts, out, local, rank)
    if world > 1:
        v = torch.tensor([verified], device=dev, dtype=torch.int64)
        dist.all_reduce(v, op=dist.ReduceOp.MIN)
        verified = int(v.item())

    status_bad = int((out["status"] != 0).sum().item())
    dropped_mean = float(out["dropped"].float().mean().item()) if "dropped" in out else None
    if rank == 0:
        peak, peak_src = measured_peak()
        bytes_sweep = algorithmic_bytes(a.n_pts, pr.nbx, pr.nby)
        med_ms = float(np.median(step_ms))
        achieved = a.sweeps * bytes_sweep / (med_ms * 1e-3) / 1e9
        value = world * a.sweeps * a.steps / (total_ms * 1e-3)
        line = {
            "metric": "lidar sweeps/s (BEV+range projection)", "value": value, "unit": "sweeps/s",
            "points_per_s": value * a.n_pts,
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": total_ms / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 decisions proven by guard bands, f64 where unproven (bit-exact with the reference's f64); f32 outputs",
            "data": "synthetic", "config": workload_config(a),
            "path": path_timed, "gpu_launches": int(launches), "frames_with_status": status_bad, "dropped_mean": dropped_mean,
            "verified_frames": verified,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": measured_traffic(path_timed, a.sweeps), "peak_source": peak_src,
                         "algorithmic_bytes_per_sweep": bytes_sweep, "launch_ms": med_ms,
                         "kernel": "fused k_fused_sweep (one launch per step)" if path_timed == "fused"
                         else "generic path: whole step (k_project + k_resolve_range + k_resolve_bev per chunk)"},
            "clocks": clocks,
        }
        if phases is not None:
            line["phase_share"] = phases
        if e2e is not None:
            line["e2e"] = e2e
        if not a.no_cpu and world == 1:
            line["cpu_baseline"] = cpu_baseline_block(a)
        emit(line)
    pr.close()
    if world > 1:
        dist.destroy_process_group()


def verify_frames(a, pkg, path_timed, d_pts, out, local, rank, k=8):
    """Returns k when k random frames of the timed output equal the generic path's rasterisation of the same
    points (range tensor and BEV density, bit for bit); raises otherwise.  0 when the timed path was the
    generic one itself."""
    import torch
    if path_timed != "fused" or a.debug_flags:
        return 0
    rng = np.random.default_rng(1234 + rank)
    frames = sorted(int(f) for f in rng.choice(a.sweeps, size=min(k, a.sweeps), replace=False))
    chk = pkg.LidarProjector(device=local, bev_res=(a.bev_res, a.bev_res), path=1)
    for f in frames:
        o = chk.project(d_pts[f * a.n_pts:(f + 1) * a.n_pts], n_frames=1, uniform_n=a.n_pts)
        torch.cuda.synchronize()
        for key in ("range", "bev", "status", "dropped"):
            if not torch.equal(o[key][0], out[key][f]):
                raise RuntimeError("bench: frame %d of the timed output differs from the generic path in %r" % (f, key))
    chk.close()
    return len(frames)


def run_e2e(a, pkg, pr, pin_in, dev, world):
    """Host pinned in -> H2D -> kernels -> D2H of every output image, pipelined over chunks on 3 streams.
    Both BEV transports are timed (the host ends with the same float images either way); the line's value
    is the faster one and the note names it."""
    import gc
    import torch
    import torch.distributed as dist
    steps = max(2, min(a.steps, 5))
    res = {}
    for transport in ("dense", "u8"):
        hb = pkg.HostBatchPipeline(pr, a.sweeps, a.n_pts, chunk=a.e2e_chunk, bev_transport=transport)
        hb.run(pin_in)            # warm-up (also faults the pinned output pages in)
        hb.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            hb.run(pin_in)
        hb.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        res[transport] = {"value": world * a.sweeps * steps / dt, "h2d": int(hb.h2d_bytes), "d2h": int(hb.d2h_bytes),
                          "chunk": hb.chunk}
        hb.close()
        del hb
        gc.collect()
    best = max(res, key=lambda k: res[k]["value"])
    r = res[best]
    return {"value": r["value"], "unit": "sweeps/s", "steps": steps,
            "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"], "bev_transport": best,
            "by_transport": {k: round(v["value"], 1) for k, v in res.items()},
            "note": "pinned host in/out, %d-sweep chunks pipelined on 3 streams; dense: float BEV over PCIe; "
                    "u8: one byte per BEV cell over PCIe, expanded into the same pinned float image by host "
                    "threads inside the timed region (bit-identical result)" % r["chunk"]}


_JSON_FD = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout.  Native libraries print there too (NCCL's version banner under
    torchrun, for one), so file descriptor 1 is pointed at stderr for the whole run and the JSON line goes to a
    private duplicate of the original stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sweeps", type=int, default=1024)
    ap.add_argument("--sensor", default="hdl32")
    ap.add_argument("--bev-res", type=float, default=0.2)
    ap.add_argument("--distinct", type=int, default=1024, help="distinct synthetic sweeps generated (rest repeat)")
    ap.add_argument("--path", type=int, default=0, help="0 auto, 1 generic, 2 fused")
    ap.add_argument("--e2e-chunk", type=int, default=32)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--debug-flags", type=int, default=0, help="kernel timing experiments (results invalid when non-zero)")
    ap.add_argument("--phases", action="store_true", help="also report the fused kernel's per-phase cycle shares")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup
    sens = {"hdl32": 65536, "hdl64": 120000, "vls128": 249984}
    a.n_pts = sens[a.sensor]
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
