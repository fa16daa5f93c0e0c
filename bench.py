#!/usr/bin/env python
"""bench.py -- site-steps/s of the time-stepping hot path on B200, with its HBM roofline and
the CPU baseline beside it (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference]

A "step" is one time step of the workload (Lanczos runs + tridiagonal eigensolves +
recombination/pointwise/boundary).

Default workload at EVERY N = the z-slab weak-scaling series of BASELINE configs[4] (`c5`):
NLSE 3D cubic, 1024 x 1024 x (128*N) complex128, one z-slab of 128 planes per GPU, halo planes and
Lanczos scalar sums exchanged inside the kernels over NVLink peer memory (N = 8 is the full 1024^3;
N = 1 is the plain 128-plane handle), so the driver's 1/2/4/8 series measures the path the north
star is about.  The other BASELINE configs (C1, C2 Gautschi, C2 Stoermer-Verlet, C3, C4) are timed
in short runs in the same process and reported in the "others" object of the same JSON line
(N = 1: all five; N > 1: C4, the ensemble, 64 trajectories per GPU).

The headline is the MEDIAN of 5 repeats of K steps, each repeat bracketed by a barrier
+ synchronize and timed with CUDA events on the handle's own stream, max over ranks.
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
REPEATS = 5


def lz_model(m):
    """vector passes of one Lanczos run as SURVEY 8d counts them (alpha step, block dots, block update,
    normalise: 2j+10 per iteration) -- the contract's "full" column, reported for reference only."""
    return (m - 1) * (m + 8) + (m - 2)


def lz_impl(m, gdr=False):
    """vector passes of one Lanczos run as IMPLEMENTED: two fused passes per iteration over an un-normalised
    basis, pass A (stencil + all dots) j+2, pass B (block update + norm) j+3 -> sum_j (2j+5) = (m-1)(m+3).
    gdr: the Gram-delayed form (the multi-kernel path's default for m <= 13; NLS_REORTH=0 in the environment switches it
    off): a 3-pass raw stencil sweep + a (j+3)-pass update per iteration, sum_j (j+6) = (m-1)(m+10)/2."""
    return (m - 1) * (m + 10) // 2 if gdr else (m - 1) * (m + 3)


def runs_gram_delayed(w):
    """does this workload's handle run the Gram-delayed Lanczos form?  (multi-kernel path = 3D grids, batches and slabs
    of the bench; the 2D single-trajectory workloads run the persistent kernel's one-pass block form)"""
    multi_kernel = w["dim"] == 3 or w.get("batch", 1) > 1
    return multi_kernel and w["m"] <= 13 and os.environ.get("NLS_REORTH", "3") != "0"


WORKLOADS = {
    "c1": dict(desc="NLSE 2D cubic focusing 256x256, dt=1e-3, m=10 (BASELINE configs[0])",
               n=256, dim=2, L=10.0, eq=EQ.NLSE_CUBIC, integ=INT.SS2, m=10, dt=1e-3, seg=100, short=(400, 20)),
    "c2_gautschi": dict(desc="sine-Gordon 2D 1024x1024 Gautschi, Lanczos m=10 (BASELINE configs[1])",
                        n=1024, dim=2, L=10.0, eq=EQ.SGE, integ=INT.GAUTSCHI, m=10, dt=1e-3, seg=50, short=(100, 10)),
    "c2_sv": dict(desc="sine-Gordon 2D 1024x1024 Stoermer-Verlet (BASELINE configs[1])",
                  n=1024, dim=2, L=10.0, eq=EQ.SGE, integ=INT.SV, m=10, dt=1e-3, seg=500, short=(2000, 100)),
    "c3": dict(desc="NLSE 3D cubic-quintic 256^3, Strang, m=12 (BASELINE configs[2])",
               n=256, dim=3, L=6.0, eq=EQ.NLSE_CQ, integ=INT.SS2, m=12, dt=5e-4, seg=10, lap=LAP.ANISO,
               s1=1.0, s2=-0.1, short=(10, 3)),
    "c4": dict(desc="ensemble of NLSE 2D cubic 512x512 trajectories, 64 per GPU, m=10 (BASELINE configs[3]: 512 at 8 GPUs)",
               n=512, dim=2, L=10.0, eq=EQ.NLSE_CUBIC, integ=INT.SS2, m=10, dt=1e-3, seg=10, batch=64, short=(10, 3)),
    # weak scaling of the slab decomposition: 1024 x 1024 x (128*G) planes, G = number of GPUs
    # (G = 8 is the full 1024^3 of BASELINE configs[4]; the full cube does not fit fewer than 2 GPUs)
    "c5": dict(desc="NLSE 3D cubic 1024x1024x(128*G) complex128, z-slabs over G GPUs, in-kernel NVLink halo + "
                    "scalar exchange, m=10 (BASELINE configs[4] at G=8)",
               n=1024, dim=3, L=10.0, eq=EQ.NLSE_CUBIC, integ=INT.SS2, m=10, dt=2e-4, seg=10, slab_planes=128,
               short=(5, 3), cpu_n=128),
    # STRONG scaling of the slab decomposition: the C3 grid (256^3, cubic-quintic, div(c grad), m=12) cut into
    # G z-slabs of 256/G planes; total work fixed, so the cross-GPU reductions and halo planes show undiluted
    "c3_strong": dict(desc="NLSE 3D cubic-quintic 256^3 (BASELINE configs[2]) cut into G z-slabs, strong scaling",
                      n=256, dim=3, L=6.0, eq=EQ.NLSE_CQ, integ=INT.SS2, m=12, dt=5e-4, seg=10, lap=LAP.ANISO,
                      s1=1.0, s2=-0.1, short=(10, 3), slab_total=256, cpu_n=64),
}
slab_ic = ics.slab_ic


def bytes_per_site_step(w, model=False):
    """algorithmic bytes per site-step of the algorithm actually EXECUTED (model=False)
    or as SURVEY.md 8d's table counts them (model=True).  S = 16 complex / 8 real."""
    m = w["m"]
    lz = lz_model(m) if model else lz_impl(m, runs_gram_delayed(w))
    if w["integ"] == INT.SS2:  # Lanczos + finish [read m columns, write u, write next start vector, read m(x) (1/2)]
        extra = 1.0 if (w["eq"] != EQ.NLSE_CUBIC and not model) else 0.0  # cq/sat: old-u density read (Q7)
        return 16 * (lz + m + 2.5 + extra)
    if w["integ"] == INT.GAUTSCHI:  # 2 Lanczos + two recombination passes
        return 8 * (2 * lz + 2 * m + 11)
    return 32.0


def grid_str(w, world):
    if "slab_total" in w:
        return f"{w['n']}x{w['n']}x{w['slab_total']}"
    if "slab_planes" in w:
        return f"{w['n']}x{w['n']}x{w['slab_planes'] * world}"
    return "x".join([str(w["n"])] * w["dim"])


def config_for(name, w, world):
    """identical in the native and the reference line (scalars and strings only)"""
    slab = "slab_planes" in w or "slab_total" in w
    cfg = {"workload": name, "description": w["desc"], "grid": grid_str(w, world),
           "trajectories_per_gpu": w.get("batch", 1), "krylov_m": w["m"], "dt": w["dt"], "reorth": "full",
           "l2": "no flush: the per-step working set (state + Krylov basis) is >= 20x the 126 MB L2 for the headline; "
                 "every step rewrites it (production access pattern)",
           "parallelism": ((f"z-slabs x{world} (peer-memory halo + in-kernel all-reduce)" if slab else f"ensemble x{world}")
                           if world > 1 else "single GPU"),
           "repeats": REPEATS}
    if "cpu_n" in w:
        cfg["reference_grid"] = "x".join([str(w["cpu_n"])] * w["dim"]) + \
            " (the CPU arm runs the same equation, dt and m on this grid: it cannot hold the slab)"
    return cfg


def make_desc(w, batch=None, device=0, n=None):
    n = n or w["n"]
    return Desc.make(n, n, n if w["dim"] == 3 else 1, L=w["L"], equation=w["eq"], integrator=w["integ"],
                     laplacian=w.get("lap", LAP.UNIFORM), krylov_m=w["m"], batch=batch or w.get("batch", 1),
                     s1=w.get("s1", 0.0), s2=w.get("s2", 0.0), dt=w["dt"], device=device,
                     flags=int(os.environ.get("NLS_BENCH_FLAGS", w.get("flags", 0))))


def make_state(w, batch, seed0=0, n=None):
    n = n or w["n"]
    if "slab_planes" in w or "slab_total" in w:
        return slab_ic(n, 0, n, n, w["L"])[None].repeat(batch, axis=0), None
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


def measured_traffic(workload):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of one time step from the committed ncu
    capture of this workload (profiles/traffic.json names the capture), or None.  Static: an ncu figure
    cannot be taken inside a timed run."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        ent = json.load(open(p)).get(workload)
        return (None, None) if ent is None else (ent["bytes_per_step"], ent.get("source"))
    except (OSError, ValueError, KeyError):
        return None, None


def cpu_baseline(w, budget_s=20.0):
    """The oracle (CPU restatement of the reference's Eigen path, kind 'port') timed on ONE host
    thread -- how the reference's host path actually executes (no -fopenmp in its CMake) -- on a
    bounded sample: the same grid and inputs (or `cpu_n`^3 for the slab workload), few steps."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    n = w.get("cpu_n", w["n"])
    d = make_desc(w, batch=1, n=n)
    u0, v0 = make_state(w, 1, n=n)
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
            "sample": f"{steps} steps of the {n}^{w['dim']} grid, 1 trajectory, 1 thread "
                      f"({el:.1f} s); reference CPU path needs Eigen3 (absent) -> oracle port"}


def run_reference(args, name, w, rank, world):
    """--impl reference: the reference's CPU algorithm (oracle port; its own sources need Eigen3,
    which this image lacks) on all host cores: one trajectory per OpenMP thread, the analogue of
    the reference's one-process-per-trajectory launch (device/submit_nlse.py:117-129)."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    orc.Oracle.set_threads(len(os.sched_getaffinity(0)))  # torchrun exports OMP_NUM_THREADS=1
    cores = orc.Oracle.threads()
    n = w.get("cpu_n", w["n"])
    d = make_desc(w, batch=cores, n=n)
    u0, v0 = make_state(w, 1, n=n)
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
            "data": "synthetic", "config": config_for(name, w, world),
            "cpu_baseline": {"value": val, "unit": "site-steps/s", "cores": cores, "kind": "port",
                             "sample": f"each step = {sub} time steps x {cores} trajectories (one per thread) "
                                       f"of the {n}^{w['dim']} grid"},
            "e2e": {"value": val, "unit": "site-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


class Ctx:
    def __init__(self, rank, world, local_rank, dist, torch):
        self.rank, self.world, self.local_rank, self.dist, self.torch = rank, world, local_rank, dist, torch

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        if self.world == 1:
            return list(vals)
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]


def measure(cx, name, w, K, W, repeats, sample_clocks):
    """one workload on this rank's GPU (z-slab workloads: collectively on all ranks).  Returns the result dict
    on rank 0 (None elsewhere)."""
    torch = cx.torch
    rank, world, lr = cx.rank, cx.world, cx.local_rank
    batch = w.get("batch", 1)
    slab = "slab_planes" in w or "slab_total" in w
    if slab:  # z-slab decomposition: one slab handle per GPU, wired over peer memory
        from nls_b200 import slabs
        n, nzg = w["n"], (w["slab_total"] if "slab_total" in w else w["slab_planes"] * world)
        dx = 2.0 * w["L"] / (n - 1)

        def mk(nz, nz_global, z0):
            return Desc.make(n, n, nz, dx=dx, equation=w["eq"], integrator=w["integ"], krylov_m=w["m"],
                             laplacian=w.get("lap", LAP.UNIFORM), s1=w.get("s1", 0.0), s2=w.get("s2", 0.0),
                             dt=w["dt"], device=lr, nz_global=nz_global, z0=z0,
                             flags=int(os.environ.get("NLS_BENCH_FLAGS", w.get("flags", 0))))
        eng, z0, nz = slabs.make_slab_engine(mk, nzg, cx.dist if world > 1 else None, rank, world,
                                             torch.device("cuda", lr))
        d = eng.desc
        # m(x) == 1 and the analytic initial condition are generated ON THE DEVICE from the global grid coordinates
        # (nls_generate_field): no 2 GiB-per-rank host arrays
        from nls_b200 import fieldgen, GEN
        eng.generate(FIELD.M, [(GEN.CONST, 0, [1.0])], L=w["L"])
        if d.laplacian == LAP.ANISO:
            eng.generate(FIELD.C, [(GEN.CONST, 0, [1.0])], L=w["L"])
        eng.generate(FIELD.U, *fieldgen.slab_ic(w["L"]), L=w["L"])
    else:
        d = make_desc(w, batch=batch, device=lr)
        u0, v0 = make_state(w, batch, seed0=rank * batch)
        eng = Engine(d)
        load_state(eng, w, u0, v0)
    dt = w["dt"]
    sites = d.n * batch

    # ---- device-resident throughput: CUDA events on the handle's own stream, median of `repeats`
    eng.step(dt, W)
    cx.barrier()
    sampler = ClockSampler(lr) if sample_clocks else None
    if sampler:
        sampler.start()
    l0 = eng.launch_count()
    ms_runs = []
    for _ in range(repeats):
        cx.barrier()
        ms = eng.step_timed(dt, K)
        cx.barrier()
        ms_runs.append(cx.max_over_ranks([ms])[0])
    launches = (eng.launch_count() - l0) // repeats
    tma_launches = eng.tma_launch_count()
    # ---- end to end through the C ABI with pinned HOST buffers: per segment of `seg` steps the state (u [, u_past])
    # is uploaded from page-locked host memory and the result (u [, v, u_past]) is read back to page-locked host
    # memory.  "e2e": the asynchronous API (nls_set_field_async / nls_get_field_async: staging + snapshot rings, the
    # copies overlap the neighbouring segments' steps); "e2e_blocking": nls_set_field / nls_get_field per segment.
    seg = min(w["seg"], K)
    nseg = max(1, K // seg)
    real = not d.is_complex
    st = eng.get(FIELD.U)
    h_in = nls_b200.pinned_empty(st.shape, st.dtype)
    h_in[...] = st
    h_out = [nls_b200.pinned_empty(st.shape, st.dtype) for _ in range(2)]
    pins = [h_in] + h_out
    if real:
        hp_in = nls_b200.pinned_empty(st.shape, st.dtype)
        hp_in[...] = eng.get(FIELD.U_PAST)
        hv_out = [nls_b200.pinned_empty(st.shape, st.dtype) for _ in range(2)]
        hp_out = [nls_b200.pinned_empty(st.shape, st.dtype) for _ in range(2)]
        pins += [hp_in] + hv_out + hp_out
    del st

    def e2e_async():
        for k in range(nseg):
            eng.set_async(FIELD.U, h_in)
            if real:
                eng.set_async(FIELD.U_PAST, hp_in)
            eng.step(dt, seg, sync=False)
            eng.get_async(FIELD.U, h_out[k & 1])
            if real:
                eng.get_async(FIELD.V, hv_out[k & 1])
                eng.get_async(FIELD.U_PAST, hp_out[k & 1])
        eng.wait()

    def e2e_blocking():
        for k in range(nseg):
            eng.set(FIELD.U, h_in)
            if real:
                eng.set(FIELD.U_PAST, hp_in)
            eng.step(dt, seg)
            eng.get_into(FIELD.U, h_out[k & 1])
            if real:
                eng.get_into(FIELD.V, hv_out[k & 1])
                eng.get_into(FIELD.U_PAST, hp_out[k & 1])

    e2e_runs, e2e_blk_runs = [], []
    for fn, runs in ((e2e_async, e2e_runs), (e2e_blocking, e2e_blk_runs)):
        fn()  # warm the copy path (allocates the rings)
        for _ in range(min(repeats, 3)):
            cx.barrier()
            t0 = time.perf_counter()
            fn()
            cx.barrier()
            runs.append(cx.max_over_ranks([time.perf_counter() - t0])[0])
    hbytes = h_in.nbytes
    clocks = sampler.stop() if sampler else None
    eng.close()
    for a in pins:
        nls_b200.pinned_free(a)
    if rank != 0:
        return None
    h2d = hbytes * (2 if real else 1)
    d2h = hbytes * (3 if real else 1)
    t_ms = float(np.median(ms_runs))
    t_e2e = float(np.median(e2e_runs))
    t_e2e_blk = float(np.median(e2e_blk_runs))
    value = sites * world * K / (t_ms * 1e-3)
    e2e = sites * world * seg * nseg / t_e2e
    bpss, bpss_model = bytes_per_site_step(w), bytes_per_site_step(w, model=True)
    peak, peak_src = peaks()
    achieved = bpss * (sites * K / (t_ms * 1e-3)) / 1e9  # GB/s on one GPU (max-over-ranks time)
    traffic, traffic_src = measured_traffic(name)
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": traffic, "traffic_unit": "DRAM bytes per time step (ncu dram__bytes_read.sum + dram__bytes_write.sum)",
            "traffic_source": traffic_src, "peak_source": peak_src,
            "bytes_per_site_step": bpss, "bytes_per_site_step_survey_model": bpss_model,
            "frac_survey_model": bpss_model / bpss * achieved / peak,
            "model": f"{bpss:.0f} algorithmic B/site-step of the algorithm executed ({lz_impl(w['m'], runs_gram_delayed(w))} vector passes per Lanczos run "
                     f"+ recombination/pointwise tail; SURVEY 8d's table counts {bpss_model:.0f}) x site-steps/s of one GPU; "
                     f"{launches // max(K, 1)} kernels/step"}
    if traffic is not None:
        dram = traffic * K / (t_ms * 1e-3) / 1e9
        roof["dram_gbs"] = dram
        roof["dram_frac"] = dram / peak
        if traffic < 0.5 * bpss * sites:  # the working set is L2-resident: HBM is not what bounds this workload
            roof["bound_note"] = "l2/latency: measured DRAM traffic is far below the algorithmic bytes (Krylov basis resident in L2)"
    return {"value": value, "ms_per_step": t_ms / K, "ms_per_step_runs": [x / K for x in ms_runs],
            "e2e": {"value": e2e, "unit": "site-steps/s", "h2d_bytes_per_step": h2d / seg,
                    "d2h_bytes_per_step": d2h / seg, "segment_steps": seg,
                    "blocking_value": sites * world * seg * nseg / t_e2e_blk,
                    "note": "per segment: pinned-host state -> device, seg steps, state(+velocity) -> pinned host through "
                            "nls_set_field_async / nls_get_field_async (copies overlap the steps); blocking_value: the same "
                            "with nls_set_field / nls_get_field; wall clock, max over ranks, median of 3"},
            "gpu_launches": int(launches), "tma_staged_launches_total": int(tma_launches),
            "clocks": clocks, "roofline": roof, "steps": K, "warmup": W, "sites_per_gpu": sites}


def run_native(args, name, w, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    cx = Ctx(rank, world, local_rank, dist, torch)
    K, W = args.steps, args.warmup
    head = measure(cx, name, w, K, W, REPEATS, sample_clocks=True)
    others = {}
    if not args.no_others:
        names = [k for k in ("c1", "c2_gautschi", "c2_sv", "c3", "c4") if k != name] if world == 1 else \
                [k for k in ("c4",) if k != name]
        for k in names:
            wk = WORKLOADS[k]
            Ks, Ws = wk["short"]
            r = measure(cx, k, wk, Ks, Ws, 3, sample_clocks=False)
            if r is not None:
                others[k] = {"description": wk["desc"], "grid": grid_str(wk, world), "value": r["value"],
                             "unit": "site-steps/s", "ms_per_step": r["ms_per_step"], "steps": Ks, "warmup": Ws,
                             "repeats": 3, "e2e": r["e2e"]["value"], "gpu_launches": r["gpu_launches"],
                             "roofline": {kk: r["roofline"][kk] for kk in
                                          ("achieved", "frac", "frac_survey_model", "bytes_per_site_step", "dram_frac",
                                           "bound_note") if kk in r["roofline"]}}
    if rank != 0:
        return
    line = {
        "metric": METRIC, "value": head["value"], "unit": "site-steps/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": head["ms_per_step"], "ms_per_step_runs": head["ms_per_step_runs"],
        "higher_is_better": True, "scaling": "strong" if "slab_total" in w else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": config_for(name, w, world),
        "e2e": head["e2e"], "gpu_launches": head["gpu_launches"] * REPEATS,
        "gpu_launches_per_repeat": head["gpu_launches"],
        "tma_staged_launches_total": head["tma_staged_launches_total"],
        "clocks": head["clocks"], "roofline": head["roofline"], "others": others,
    }
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(w)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="c5", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-others", action="store_true", help="skip the short runs of the other BASELINE configs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    w = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, args.workload, w, rank, world)
        return
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl")
    try:
        run_native(args, args.workload, w, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
