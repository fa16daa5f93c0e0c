"""z-slab parity worker: run under torchrun with >= 2 ranks, one GPU per rank (backend nccl).
Every rank owns a slab of the same global problem; rank 0 gathers the slabs and checks them
against the CPU oracle on the full grid and against a single-GPU handle.  Prints SLAB_OK."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nls_b200  # noqa: E402
from nls_b200 import Desc, Engine, EQ, INT, LAP, FN, FIELD, RED, slabs  # noqa: E402
import ics  # noqa: E402


def rel_l2(a, b):
    return float(np.linalg.norm(a.ravel() - b.ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def gather(local, dev):
    """rank 0 gets the list of every rank's (possibly differently sized) slab array"""
    local = np.ascontiguousarray(local)
    cplx = np.iscomplexobj(local)
    t = torch.from_numpy(local.view(np.float64) if cplx else local).to(dev)
    n = torch.tensor([t.numel()], device=dev)
    ns = [torch.zeros_like(n) for _ in range(dist.get_world_size())]
    dist.all_gather(ns, n)
    mx = int(max(x.item() for x in ns))
    pad = torch.zeros(mx, dtype=t.dtype, device=dev)
    pad[:t.numel()] = t.ravel()
    parts = [torch.empty_like(pad) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, pad)
    out = [p[:int(k.item())].cpu().numpy() for p, k in zip(parts, ns)]
    return [o.view(np.complex128) for o in out] if cplx else out


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl")
    import orc
    failures = []

    def case(name, n, nzg, kw, steps, dt, real, tol, with_oracle=True, tol_1gpu=1e-12):
        L = 4.0
        rng = np.random.default_rng(11)
        cfield = 1.0 + 0.1 * rng.standard_normal((nzg, n, n))
        mfield = 1.0 + 0.2 * rng.standard_normal((nzg, n, n))
        if real:
            u0 = 0.3 * np.abs(ics.gaussian_bumps_3d(max(n, nzg), L))[:nzg, :n, :n]
            v0 = 0.05 * np.cos(u0)
        else:
            u0 = ics.gaussian_bumps_3d(max(n, nzg), L)[:nzg, :n, :n]
            v0 = None
        x = rng.standard_normal((nzg, n, n)) + (0 if real else 1j * rng.standard_normal((nzg, n, n)))
        dx = 2 * L / (n - 1)

        def mk(nz, nz_global, z0, device=lr):
            return Desc.make(n, n, nz, dx=dx, nz_global=nz_global, z0=z0, device=device, dt=dt, **kw)

        eng, z0, nz = slabs.make_slab_engine(mk, nzg, dist, rank, world, dev)
        aniso = kw.get("laplacian", 0) == LAP.ANISO
        if aniso:
            eng.set(FIELD.C, slabs.local_planes(cfield, z0, nz))
        eng.set(FIELD.M, slabs.local_planes(mfield, z0, nz))
        res = {}
        res["stencil"] = gather(eng.stencil_apply(slabs.local_planes(x, z0, nz)), dev)
        fn = FN.COS_SQRT if real else FN.EXP
        res["matfunc"] = gather(eng.matfunc(fn, 1e-2 if real else 1e-2j, slabs.local_planes(x, z0, nz)), dev)
        eng.set(FIELD.U, slabs.local_planes(u0, z0, nz))
        if real:
            eng.set(FIELD.V, slabs.local_planes(v0, z0, nz))
        eng.step(dt, steps)
        res["u"] = gather(eng.get(FIELD.U), dev)
        mass = eng.reduce(RED.MASS)[0]
        dist.barrier()
        eng.close()
        if rank == 0:
            full = mk(nzg, 0, 0)
            e1 = Engine(full)
            o = orc.Oracle(full) if with_oracle else e1  # big grids: the single-GPU handle stands in
            for s in (o, e1) if with_oracle else (e1,):
                if aniso:
                    s.set(FIELD.C, cfield)
                s.set(FIELD.M, mfield)
            checks = []
            for key, fun in (("stencil", lambda s: s.stencil_apply(x)),
                             ("matfunc", lambda s: s.matfunc(fn, 1e-2 if real else 1e-2j, x))):
                got = np.concatenate(res[key])
                checks.append((key + " vs oracle", rel_l2(got, fun(o)), 1e-15 if key == "stencil" else 1e-11))
            for s in (o, e1) if with_oracle else (e1,):
                s.set(FIELD.U, u0)
                if real:
                    s.set(FIELD.V, v0)
                s.step(dt, steps)
            got = np.concatenate(res["u"])
            checks.append(("steps vs oracle", rel_l2(got, o.get()), tol))
            checks.append(("steps vs 1 GPU", rel_l2(got, e1.get()), tol_1gpu))
            checks.append(("mass", abs(mass - o.reduce(RED.MASS)[0]) / o.reduce(RED.MASS)[0], 1e-12))
            for what, err, bound in checks:
                okay = err <= bound
                print(f"[slab x{world}] {name}: {what}: {err:.3e} (<= {bound:g}) {'ok' if okay else 'FAIL'}", flush=True)
                if not okay:
                    failures.append((name, what, err))
        dist.barrier()

    def Z(base):  # at least 3 planes on every rank, and not a multiple of the rank count
        return max(base, 3 * world + 1 + base % 2)

    case("nlse cubic 3D uniform", 20, 20 + world, dict(equation=EQ.NLSE_CUBIC, krylov_m=10), 8, 1e-3, False, 1e-11)
    case("nlse cq 3D aniso m=12", 24, Z(26), dict(equation=EQ.NLSE_CQ, laplacian=LAP.ANISO, krylov_m=12, s1=1.0, s2=-0.1),
         6, 5e-4, False, 1e-11)
    case("nlse 3-term", 16, Z(18), dict(equation=EQ.NLSE_CUBIC, krylov_m=8, reorth=1), 5, 1e-3, False, 1e-11)
    case("nlse cq 3D aniso, TMA passes", 32, Z(28), dict(equation=EQ.NLSE_CQ, laplacian=LAP.ANISO, krylov_m=12, s1=1.0, s2=-0.1,
                                                     flags=nls_b200.FLAG_TMA), 5, 5e-4, False, 1e-11)
    case("nlse sewi 3D", 18, Z(20), dict(equation=EQ.NLSE_CUBIC, integrator=INT.SEWI, krylov_m=8), 5, 1e-3, False, 1e-11)
    case("kge 3D gautschi aniso", 20, Z(22), dict(equation=EQ.KGE, integrator=INT.GAUTSCHI, laplacian=LAP.ANISO,
                                               krylov_m=10, op_sign=-1.0), 6, 1e-3, True, 1e-11)
    case("sge 3D stoermer-verlet", 20, Z(23), dict(equation=EQ.SGE, integrator=INT.SV), 40, 1e-3, True, 1e-11)
    if world > 1:  # thinnest legal decomposition: 2 planes on all ranks but one
        case("thin slabs, 2 planes per rank", 16, 2 * world + 1, dict(equation=EQ.NLSE_CUBIC, krylov_m=8), 5, 1e-3, False, 1e-11)
        case("thin slabs, kge gautschi", 16, 2 * world + 1, dict(equation=EQ.KGE, integrator=INT.GAUTSCHI, krylov_m=8,
                                                               op_sign=-1.0), 5, 1e-3, True, 1e-11)
    # BASELINE configs[4] semantics at 256^3: slab-count invariance (G slabs vs one GPU) <= 1e-13, no oracle
    case("C5 slab-count invariance 256^3 cubic m=10", 256, 256, dict(equation=EQ.NLSE_CUBIC, krylov_m=10), 3, 2e-4, False,
         1e-13, with_oracle=False, tol_1gpu=1e-13)
    case("C3 slab-count invariance 256^3 cq aniso m=12", 256, 256,
         dict(equation=EQ.NLSE_CQ, laplacian=LAP.ANISO, krylov_m=12, s1=1.0, s2=-0.1), 2, 5e-4, False, 1e-13,
         with_oracle=False, tol_1gpu=1e-13)
    if rank == 0:
        print("SLAB_FAIL" if failures else "SLAB_OK", failures, flush=True)
    dist.destroy_process_group()
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
