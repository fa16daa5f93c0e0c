#!/usr/bin/env python
"""launcher_shim.py -- hand the reference launchers' run loop to the engine as ENSEMBLE batches.

The reference spawns one driver process per run: finalized_scripts/complex_launcher_2d.py:252-268 (`run()` calls
`run_simulation` -> subprocess per run, :130-171) and device/submit_nlse.py:83-138 (one rank per initial condition).
Every process pays CUDA context creation, matrix assembly and file staging.  The B200 drivers accept ensemble files
`(B, ny, nx)` and advance all B trajectories in one batched handle per GPU (host/drivers.cpp), so this shim

  * `run_batch(...)`: groups the runs that share the same m(x) and c(x) fields (a handle's fields are shared by its
    batch), writes ONE ensemble file per group, calls the driver ONCE per group (optionally `--devices G`,
    `--downsample N`) and returns the trajectories in run order -- bit-identical to what the per-run calls write;
  * `patch_launcher(cls)`: replaces `run()` of a reference launcher class (anything with the attribute/method
    names of complex_launcher_2d.NLSELauncher) by a version that first draws every run's (u0, c, m) in the original
    RNG order, runs them through `run_batch`, then post-processes each run with the launcher's own
    `downsample_trajectory` / `save_to_hdf5` / `cleanup`.

The launcher's `run_simulation` saves m under the c file name (`np.save(c_file, m)`, complex_launcher_2d.py:140-141),
so the driver it spawns receives c == m; `launcher_c_is_m=True` (default in `patch_launcher`) keeps that behaviour.
"""
import hashlib
import os
import subprocess
import tempfile
import time

import numpy as np


def _key(a):
    return None if a is None else hashlib.sha1(np.ascontiguousarray(a).view(np.uint8)).hexdigest()


def group_runs(ms, cs):
    """indices of the runs grouped by identical (m, c) content, groups in order of first appearance"""
    groups = {}
    for i, (m, c) in enumerate(zip(ms, cs)):
        groups.setdefault((_key(m), _key(c)), []).append(i)
    return list(groups.values())


def run_batch(exe, nx, ny, Lx, Ly, u0s, ms, cs, T, nt, snapshots, scalars=(), v0s=None, workdir=None, devices=1,
              downsample=0, extra_args=()):
    """Run len(u0s) trajectories through `exe` (a reference driver name built by host/Makefile) with as few
    processes as the shared-field rule allows.  Returns (list of trajectories [(u[, v])...] in run order, seconds)."""
    n = len(u0s)
    ms = list(ms) if ms is not None else [None] * n
    cs = list(cs) if cs is not None else [None] * n
    out = [None] * n
    t0 = time.time()
    with tempfile.TemporaryDirectory(dir=workdir) as tmp:
        for g, idx in enumerate(group_runs(ms, cs)):
            f = lambda s: os.path.join(tmp, f"g{g}_{s}.npy")  # noqa: E731
            ens = len(idx) > 1
            np.save(f("u0"), np.stack([u0s[i] for i in idx]) if ens else u0s[idx[0]])
            cmd = [str(exe), str(nx), str(ny), str(Lx), str(Ly)] + [str(s) for s in scalars] + [f("u0")]
            real = v0s is not None
            if real:
                np.save(f("v0"), np.stack([v0s[i] for i in idx]) if ens else v0s[idx[0]])
                cmd += [f("v0"), f("traj"), f("vel")]
            else:
                cmd += [f("traj")]
            cmd += [str(T), str(nt), str(snapshots)]
            if ms[idx[0]] is not None:
                np.save(f("m"), ms[idx[0]])
                cmd.append(f("m"))
            if cs[idx[0]] is not None:
                np.save(f("c"), cs[idx[0]])
                cmd.append(f("c"))
            if devices > 1 and ens:
                cmd += ["--devices", str(min(devices, len(idx)))]
            if downsample:
                cmd += ["--downsample", str(downsample)]
            cmd += list(extra_args)
            r = subprocess.run(cmd, check=True, capture_output=True, text=True)
            if r.stderr:
                print(f"Warnings/Errors: {r.stderr}")
            tu = np.load(f("traj"))
            tv = np.load(f("vel")) if real else None
            for k, i in enumerate(idx):
                u = tu[k] if ens else tu
                out[i] = (u, tv[k] if ens else tv) if real else u
    return out, time.time() - t0


def patch_launcher(cls, launcher_c_is_m=True, devices=1):
    """Replace cls.run by the batched loop (see the module docstring).  Returns cls."""

    def run(self):
        a = self.args
        runs = []
        for i in range(a.num_runs):  # same RNG call order as the original loop (simulation draws nothing)
            u0, params = self.generate_initial_condition(i)
            c = self.generate_anisotropy(i)
            m = self.generate_spatial_amplification(i, c)
            runs.append((u0, params, m, c))
        trajs, wall = run_batch(a.exe, a.nx, a.ny, a.Lx, a.Ly, [r[0] for r in runs], [r[2] for r in runs],
                                [(r[2] if launcher_c_is_m else r[3]) for r in runs], a.T, a.nt, a.snapshots,
                                devices=devices)
        for i, ((u0, params, m, c), traj) in enumerate(zip(runs, trajs)):
            try:
                traj = self.downsample_trajectory(traj)
                self.save_to_hdf5(i, u0, m, c, traj, params, wall / max(1, a.num_runs))
                self.cleanup(i)
            except Exception as e:  # the reference loop continues with the next run
                print(f"Error in run {i+1}: {e}")
        return trajs

    cls.run = run
    return cls


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser(description="run B initial conditions (one .npy of shape (B, ny, nx)) through a driver "
                                             "as one ensemble; the batched replacement of device/submit_nlse.py's rank loop")
    ap.add_argument("exe"); ap.add_argument("u0"); ap.add_argument("out")
    ap.add_argument("--L", type=float, default=10.0); ap.add_argument("--T", type=float, default=1.0)
    ap.add_argument("--nt", type=int, default=1000); ap.add_argument("--snapshots", type=int, default=100)
    ap.add_argument("--m"); ap.add_argument("--c"); ap.add_argument("--devices", type=int, default=1)
    args = ap.parse_args()
    u0 = np.load(args.u0)
    m = np.load(args.m) if args.m else None
    c = np.load(args.c) if args.c else None
    B, ny, nx = u0.shape
    trajs, wall = run_batch(args.exe, nx, ny, args.L, args.L, list(u0), [m] * B, [c] * B, args.T, args.nt, args.snapshots,
                            devices=args.devices)
    np.save(args.out, np.stack(trajs))
    print(f"{B} trajectories in {wall:.2f} s")
