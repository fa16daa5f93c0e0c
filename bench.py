#!/usr/bin/env python
"""bench.py -- site-steps/s of the time-stepping hot path on B200, with its HBM roofline and
the CPU baseline beside it (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference]

A "step" is one time step of the workload (Lanczos runs + tridiagonal eigensolves +
recombination/pointwise/boundary).  Default workload = BASELINE.json configs[1]:
sine-Gordon 2D 1024x1024, Gautschi integrator, Krylov m=10 (`c2_gautschi`).
N>1: one process per GPU (torchrun), every rank advances its own trajectory of the same
workload (ensemble sharding, no data-path collective) -> "scaling": "weak".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import nls_b200  # noqa: E402
from nls_b200 import Desc, Engine, EQ, INT, LAP, FIELD  # noqa: E402
import ics  # noqa: E402

METRIC = "site-steps/sec 2D/3D NLSE & SGE at 1/2/4/8 B200; % of HBM roofline"


def lz_full(m):
    """vector passes of one Lanczos run with full re-orthogonalisation."""
    return (m - 1) * (m + 8) + (m - 2)


def lz_3t(m):
    return 8 * m - 9


WORKLOADS = {
    # name: (description, builder kwargs)
    "c2_gautschi": dict(desc="sine-Gordon 2D 1024x1024 Gautschi, Lanczos m=10 (BASELINE configs[1])",
                        n=1024, dim=2, L=10.0, eq=EQ.SGE, integ=INT.GAUTSCHI, m=10, dt=1e-3, seg=50),
    "c2_sv": dict(desc="sine-Gordon 2D 1024x1024 Stoermer-Verlet (BASELINE configs[1])",
                  n=1024, dim=2, L=10.0, eq=EQ.SGE, integ=INT.SV, m=10, dt=1e-3, seg=50),
    "c1": dict(desc="NLSE 2D cubic focusing 256x256, dt=1e-3, m=10 (BASELINE configs[0])",
               n=256, dim=2, L=10.0, eq=EQ.NLSE_CUBIC, integ=INT.SS2, m=10, dt=1e-3, seg=10),
    "c3": dict(desc="NLSE 3D cubic-quintic 256^3, Strang, m=12 (BASELINE configs[2])",
               n=256, dim=3, L=6.0, eq=EQ.NLSE_CQ, integ=INT.SS2, m=12, dt=5e-4, seg=10, lap=LAP.ANISO,
               s1=1.0, s2=-0.1),
    "c4": dict(desc="ensemble of NLSE 2D cubic 512x512 trajectories, m=10 (BASELINE configs[3])",
               n=512, dim=2, L=10.0, eq=EQ.NLSE_CUBIC, integ=INT.SS2, m=10, dt=1e-3, seg=10, batch=64),
    # weak scaling of the slab decomposition: 1024 x 1024 x (128*G) planes, G = number of GPUs
    # (G = 8 is the full 1024^3 of BASELINE configs[4]; the full cube does not fit fewer than 2 GPUs)
    "c5": dict(desc="NLSE 3D cubic 1024x1024x(128*G) complex128, z-slabs over G GPUs, in-kernel NVLink halo + "
                    "scalar exchange, m=10 (BASELINE configs[4] at G=8)",
               n=1024, dim=3, L=10.0, eq=EQ.NLSE_CUBIC, integ=INT.SS2, m=10, dt=2e-4, seg=5, slab_planes=128),
}


def slab_ic(n, z0, nz, nzg, L):
    """seed-free analytic IC for the slab workload: 8 fixed Gaussians, evaluated per slab."""
    x = np.linspace(-L, L, n)
    z = np.linspace(-L, L, nzg)[z0:z0 + nz]
    u = np.zeros((nz, n, n), dtype=np.complex128)
    for k in range(8):
        c = 0.5 * L * np.array([np.cos(1.3 * k), np.sin(2.1 * k), np.cos(0.7 * k + 1.0)])
        s = 0.8 + 0.05 * k
        gx = np.exp(-(x - c[0]) ** 2 / (2 * s * s)) * np.exp(1j * 0.3 * k * x)
        gy = np.exp(-(x - c[1]) ** 2 / (2 * s * s))
        gz = np.exp(-(z - c[2]) ** 2 / (2 * s * s))
        u += (0.5 + 0.05 * k) * gz[:, None, None] * gy[None, :, None] * gx[None, None, :]
    return u


def bytes_per_site_step(w, reorth_full=True):
    """algorithmic bytes per site-step (S = 16 complex / 8 real)."""
    m = w["m"]
    lz = lz_full(m) if reorth_full else lz_3t(m)
    if w["integ"] == INT.SS2:
        return 16 * (lz + m + 2.5)
    if w["integ"] == INT.GAUTSCHI:
        return 8 * (2 * lz + 2 * m + 11)
    return 32.0


def make_desc(w, batch=None, device=0):
    n = w["n"]
    return Desc.make(n, n, n if w["dim"] == 3 else 1, L=w["L"], equation=w["eq"], integrator=w["integ"],
                     laplacian=w.get("lap", LAP.UNIFORM), krylov_m=w["m"], batch=batch or w.get("batch", 1),
                     s1=w.get("s1", 0.0), s2=w.get("s2", 0.0), dt=w["dt"], device=device,
                     flags=int(os.environ.get("NLS_BENCH_FLAGS", w.get("flags", 0))))


def make_state(w, batch, seed0=0):
    n = w["n"]
    if w["eq"] <= EQ.NLSE_SAT:
        if w["dim"] == 3:
            u0 = np.stack([ics.gaussian_bumps_3d(n, w["L"], seed=99 + seed0 + b) for b in range(batch)])
        else:
            u0 = np.stack([ics.gaussian_bumps_2d(n, w["L"], seed=10000 + seed0 + b) for b in range(batch)])
        return u0, None
    u0 = np.stack([ics.kink_2d(n, w["L"], seed=77 + seed0 + b) for b in range(batch)])
    return u0, np.zeros_like(u0)


def load_state(s, w, u0, v0):
    d = s.desc
    s.set(FIELD.M, np.ones(d.n))
    if d.laplacian == LAP.ANISO:
        s.set(FIELD.C, np.ones(d.n))
    s.set(FIELD.U, u0)
    if v0 is not None:
        s.set(FIELD.V, v0)


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(workload, steps_per_launch):
    """ncu-measured DRAM bytes of the dominant kernel per launch (profiles/traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        ent = json.load(open(p)).get(workload)
        return None if ent is None else ent["bytes_per_step"] * steps_per_launch
    except (OSError, ValueError, KeyError):
        return None


def pinned_like(a):
    import torch
    t = torch.empty(a.shape, dtype=torch.complex128 if np.iscomplexobj(a) else torch.float64, pin_memory=True)
    out = t.numpy()
    out[...] = a
    return out, t  # keep the tensor alive


def cpu_baseline(w, budget_s=20.0):
    """The oracle (CPU restatement of the reference's Eigen path, kind 'port') timed on ONE host
    thread -- how the reference's host path actually executes (no -fopenmp in its CMake) -- on a
    bounded sample: the same grid and inputs, few steps."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    if "slab_planes" in w:  # the CPU cannot hold the slab workload: same equation on a 128^3 cube
        w = dict(w, n=128)
    d = make_desc(w, batch=1)
    u0, v0 = make_state(w, 1)
    o = orc.Oracle(d)
    load_state(o, w, u0, v0)
    t0 = time.perf_counter()
    o.step(w["dt"], 1)
    t1 = time.perf_counter() - t0
    steps = int(max(1, min(200, budget_s / max(t1, 1e-4))))
    t0 = time.perf_counter()
    o.step(w["dt"], steps)
    el = time.perf_counter() - t0
    return {"value": d.n * steps / el, "unit": "site-steps/s", "cores": 1, "kind": "port",
            "sample": f"{steps} steps of the same {w['n']}^{w['dim']} grid, 1 trajectory, 1 thread "
                      f"({el:.1f} s); reference CPU path needs Eigen3 (absent) -> oracle port"}


def run_reference(args, w, rank, world):
    """--impl reference: the reference's CPU algorithm (oracle port; its own sources need Eigen3,
    which this image lacks) on all host cores: one trajectory per OpenMP thread, the analogue of
    the reference's one-process-per-trajectory launch (device/submit_nlse.py:117-129)."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    orc.Oracle.set_threads(len(os.sched_getaffinity(0)))  # torchrun exports OMP_NUM_THREADS=1
    cores = orc.Oracle.threads()
    if "slab_planes" in w:  # the CPU cannot hold the slab workload: same equation on a 128^3 cube
        w = dict(w, n=128)
    d = make_desc(w, batch=cores)
    u0, v0 = make_state(w, 1)
    u0 = np.repeat(u0, cores, axis=0)
    v0 = None if v0 is None else np.repeat(v0, cores, axis=0)
    o = orc.Oracle(d)
    load_state(o, w, u0, v0)
    # bounded sample: every bench step = `sub` time steps of `cores` trajectories
    t0 = time.perf_counter()
    o.step(w["dt"], 1)
    t1 = time.perf_counter() - t0
    total = args.steps + args.warmup
    sub = int(max(1, min(50, 90.0 / (t1 * total))))
    for _ in range(args.warmup):
        o.step(w["dt"], sub)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        o.step(w["dt"], sub)
    el = time.perf_counter() - t0
    val = d.n * cores * sub * args.steps / el
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "site-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": args.workload, "description": w["desc"]},
            "cpu_baseline": {"value": val, "unit": "site-steps/s", "cores": cores, "kind": "port",
                             "sample": f"each step = {sub} time steps x {cores} trajectories (one per thread) "
                                       f"of the {w['n']}^{w['dim']} grid"},
            "e2e": {"value": val, "unit": "site-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_native(args, w, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    batch = w.get("batch", 1)
    if "slab_planes" in w:  # z-slab decomposition: one slab handle per GPU, wired over peer memory
        from nls_b200 import slabs
        n, nzg = w["n"], w["slab_planes"] * world
        dx = 2.0 * w["L"] / (n - 1)

        def mk(nz, nz_global, z0):
            return Desc.make(n, n, nz, dx=dx, equation=w["eq"], integrator=w["integ"], krylov_m=w["m"],
                             dt=w["dt"], device=local_rank, nz_global=nz_global, z0=z0)
        eng, z0, nz = slabs.make_slab_engine(mk, nzg, dist if world > 1 else None, rank, world,
                                             torch.device("cuda", local_rank))
        d = eng.desc
        u0, v0 = slab_ic(n, z0, nz, nzg, w["L"])[None], None
    else:
        d = make_desc(w, batch=batch, device=local_rank)
        u0, v0 = make_state(w, batch, seed0=rank * batch)
        eng = Engine(d)
    load_state(eng, w, u0, v0)
    dt, K, W = w["dt"], args.steps, args.warmup
    sites = d.n * batch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput: CUDA events on the handle's own stream
    eng.step(dt, W)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = eng.launch_count()
    ms = eng.step_timed(dt, K)
    launches = eng.launch_count() - l0
    tma_launches = eng.tma_launch_count()
    barrier()
    # ---- end to end through the C ABI with pinned HOST buffers: per segment of `seg` steps,
    # H2D of the state (u, u_past / u) and D2H of the result (u, v / u)
    seg = min(w["seg"], K)
    nseg = max(1, K // seg)
    hu, _k1 = pinned_like(eng.get(FIELD.U))
    real = not d.is_complex
    if real:
        hp, _k2 = pinned_like(eng.get(FIELD.U_PAST))
        hv, _k3 = pinned_like(np.zeros_like(hu))
    for _ in range(2):  # warm the copy path
        eng.set(FIELD.U, hu)
        eng.step(dt, 1)
        eng.get_into(FIELD.U, hu)
    barrier()
    t0 = time.perf_counter()
    for _ in range(nseg):
        eng.set(FIELD.U, hu)
        if real:
            eng.set(FIELD.U_PAST, hp)
        eng.step(dt, seg)
        eng.get_into(FIELD.U, hu)
        if real:
            eng.get_into(FIELD.V, hv)
            eng.get_into(FIELD.U_PAST, hp)
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()
    h2d = hu.nbytes * (2 if real else 1)
    d2h = hu.nbytes * (3 if real else 1)
    t_ms, t_e2e = ms, e2e_s
    if world > 1:
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_ms, t_e2e = float(t[0]), float(t[1])
    if rank != 0:
        return
    value = sites * world * K / (t_ms * 1e-3)
    e2e = sites * world * seg * nseg / t_e2e
    bpss = bytes_per_site_step(w)
    peak, peak_src = peaks()
    achieved = bpss * (sites * K / (ms * 1e-3)) / 1e9  # GB/s on this rank's GPU
    line = {
        "metric": METRIC, "value": value, "unit": "site-steps/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": t_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "description": w["desc"],
                   "grid": [w["n"], w["n"], w["slab_planes"] * world] if "slab_planes" in w else [w["n"]] * w["dim"],
                   "trajectories_per_gpu": batch, "krylov_m": w["m"], "dt": dt, "reorth": "full",
                   "l2": "no flush: state+basis working set per step exceeds/approaches L2 and every step "
                         "rewrites it (production access pattern)",
                   "parallelism": (f"z-slabs x{world} (peer-memory halo + in-kernel all-reduce)" if "slab_planes" in w
                                   else f"ensemble x{world}") if world > 1 else "single GPU"},
        "e2e": {"value": e2e, "unit": "site-steps/s", "h2d_bytes_per_step": h2d / seg,
                "d2h_bytes_per_step": d2h / seg, "segment_steps": seg,
                "note": "per segment: pinned-host state -> device, seg steps, state+velocity -> pinned host"},
        "gpu_launches": int(launches), "tma_staged_launches_total": int(tma_launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": measured_traffic(args.workload, K if launches <= 2 else 1), "peak_source": peak_src,
                     "model": f"{bpss:.0f} algorithmic B/site-step (SURVEY 8d, full re-orthogonalisation) x "
                              f"site-steps/s; whole-step graph ({launches // max(K, 1)} kernels/step)"},
    }
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(w)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="c2_gautschi", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, w, rank, world)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl")
    try:
        run_native(args, w, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
