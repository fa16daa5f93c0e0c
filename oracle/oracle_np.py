"""Independent numpy/scipy twin of the CPU oracle (TEST INFRASTRUCTURE ONLY).

Restates the same reference functions as oracle/nls_oracle.cpp with different machinery
(scipy.sparse COO assembly, numpy.linalg.eigh on the full m x m T) so that a slip in one
restatement shows up as a disagreement between the two.  Parity status: see oracle/nls_oracle.h
(pinned against the reference sources via oracle/_ref; Eigen's own arithmetic substituted).  Small grids only (pure-Python assembly loops).
"""
import numpy as np
import scipy.sparse as sp


def lap2d_uniform(nx, ny, dx, dy):
    """build_laplacian_noflux(nx-2, ny-2, dx, dy): nlsolvers/common/include/laplacians.hpp:9-52."""
    N = nx * ny
    r, c, v = [], [], []
    for i in range(N):  # :23-30
        ring = i < nx or i >= N - nx or i % nx == 0 or i % nx == nx - 1
        r.append(i); c.append(i); v.append(-3.0 if ring else -4.0)
    for i in range(N - 1):  # :32-37
        if (i + 1) % nx != 0:
            r += [i, i + 1]; c += [i + 1, i]; v += [1.0, 1.0]
    for i in range(N - nx):  # :39-42
        r += [i, i + nx]; c += [i + nx, i]; v += [1.0, 1.0]
    return sp.coo_matrix((v, (r, c)), shape=(N, N)).tocsr() * (1.0 / (dx * dy))


def lap2d_aniso(nx, ny, dx, dy, cf):
    """build_anisotropic_laplacian_noflux: laplacians.hpp:54-103."""
    N = nx * ny
    cf = np.asarray(cf).ravel()
    r, c, v = [], [], []
    diag = np.zeros(N)
    for i in range(N - 1):
        if (i + 1) % nx != 0:
            a = (cf[i] + cf[i + 1]) / 2.0
            r += [i, i + 1]; c += [i + 1, i]; v += [a, a]
            diag[i] += a; diag[i + 1] += a
    for i in range(N - nx):
        a = (cf[i] + cf[i + nx]) / 2.0
        r += [i, i + nx]; c += [i + nx, i]; v += [a, a]
        diag[i] += a; diag[i + nx] += a
    r += list(range(N)); c += list(range(N)); v += list(-diag)
    return sp.coo_matrix((v, (r, c)), shape=(N, N)).tocsr() * (1.0 / (dx * dy))


def lap3d(nx, ny, nz, dx, cf=None):
    """build_laplacian_noflux_3d (:105-156) / build_anisotropic_laplacian_noflux_3d (:158-218),
    including the y-neighbour loop without a plane check (:140-143, :192-198)."""
    P = nx * ny
    N = P * nz
    r, c, v = [], [], []
    diag = np.zeros(N)
    w = (lambda a, b: 1.0) if cf is None else (lambda a, b: (cf[a] + cf[b]) / 2.0)
    if cf is not None:
        cf = np.asarray(cf).ravel()

    def link(a, b):
        t = w(a, b)
        r.extend([a, b]); c.extend([b, a]); v.extend([t, t])
        diag[a] += t; diag[b] += t
    for i in range(N - 1):
        if (i + 1) % nx != 0:
            link(i, i + 1)
    for i in range(N - nx):
        link(i, i + nx)
    for i in range(N - P):
        link(i, i + P)
    if cf is None:  # :119-131 uniform diagonal: -5 on any face, else -6
        k, j, i = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
        face = (k == 0) | (k == nz - 1) | (j == 0) | (j == ny - 1) | (i == 0) | (i == nx - 1)
        dg = np.where(face, -5.0, -6.0).ravel()
    else:
        dg = -diag
    r += list(range(N)); c += list(range(N)); v += list(dg)
    return sp.coo_matrix((v, (r, c)), shape=(N, N)).tocsr() * (1.0 / (dx * dx))


def lanczos(A, u, m, reorth=True):
    """lanczos_L_complex eigen_krylov_complex.hpp:13-52 / lanczos_L eigen_krylov_real.hpp:8-55.
    Returns V (n x m), T (m x m, T[m-1,m-1] stays 0), beta."""
    n = u.shape[0]
    V = np.zeros((n, m), dtype=u.dtype)
    T = np.zeros((m, m), dtype=u.dtype)
    beta = np.linalg.norm(u)
    V[:, 0] = u / beta
    for j in range(m - 1):
        w = A @ V[:, j]
        if j > 0:
            w = w - T[j - 1, j] * V[:, j - 1]
        T[j, j] = np.vdot(V[:, j], w)
        w = w - T[j, j] * V[:, j]
        if reorth:
            for i in range(j + 1):
                w = w - np.vdot(V[:, i], w) * V[:, i]
        T[j + 1, j] = T[j, j + 1] = np.linalg.norm(w)
        V[:, j + 1] = w / T[j + 1, j]
    return V, T, beta


def _sinc(x):
    x = np.asarray(x)
    out = np.ones_like(x)
    nz = np.abs(x) >= 1e-8
    out[nz] = np.sin(x[nz]) / x[nz]
    return out


def f_table(fn, t, lam):
    """Function tables of eigen_krylov_complex.hpp:70-76,:89-96 and eigen_krylov_real.hpp."""
    if fn == "exp":
        return np.exp(t * lam)
    if fn == "sinc":
        return _sinc(t * lam)
    th = np.real(t) * np.sqrt(np.abs(lam))
    if fn == "cos_sqrt":
        return np.cos(th)
    if fn == "sinc2_sqrt":
        return _sinc(th) ** 2
    if fn == "id_sqrt":
        return th
    if fn == "sinc2_half":
        return _sinc(th / 2.0) ** 2
    if fn == "mod_cosine":
        out = np.ones_like(th)
        nz = np.abs(th) >= 1e-12
        out[nz] = np.cos(th[nz] / 2) ** 2 * np.sin(th[nz]) / th[nz]
        return out ** 2
    raise ValueError(fn)


def matfunc(A, u, t, fn, m=10, reorth=True):
    """beta * V * Q f(Lambda) Q^H * e1  (expm_multiply eigen_krylov_complex.hpp:54-83 etc.)."""
    V, T, beta = lanczos(A, u, m, reorth)
    lam, Q = np.linalg.eigh(T)  # lower triangle, like SelfAdjointEigenSolver
    fT = (Q * f_table(fn, t, lam)) @ Q.conj().T
    r = beta * (V @ fT[:, 0])
    return r if np.iscomplexobj(u) else r.real


def neumann_bc(u):
    """neumann_bc_no_velocity boundaries.hpp:40-57 / _3d boundaries_3d.hpp:8-32; u shaped
    (ny,nx) or (nz,ny,nx), modified in place."""
    if u.ndim == 2:
        u[1:-1, 0] = u[1:-1, 1]
        u[1:-1, -1] = u[1:-1, -2]
        u[0, :] = u[1, :]
        u[-1, :] = u[-2, :]
    else:
        u[1:-1, 1:-1, 0] = u[1:-1, 1:-1, 1]
        u[1:-1, 1:-1, -1] = u[1:-1, 1:-1, -2]
        u[1:-1, 0, :] = u[1:-1, 1, :]
        u[1:-1, -1, :] = u[1:-1, -2, :]
        u[0] = u[1]
        u[-1] = u[-2]
    return u


def nlse_step(A, u, mfield, dt, m=10, kind="cubic", s1=0.0, s2=0.0, kappa=0.0):
    """NLSESolver::step nlse_cubic_solver.hpp:55-74, NLSEQuinticSolver::step
    nlse_cubic_quintic_solver.hpp:10-29, NLSESaturatingSolver::step nlse_saturating_solver.hpp:11-33.
    u, mfield flat; returns new flat u (before BC)."""
    tau = 1j * dt

    def g(dens_src, den_src):
        r = np.abs(dens_src) ** 2
        if kind == "cubic":
            return mfield * r
        if kind == "cq":
            return mfield * (s1 * r + s2 * r * r)
        return mfield * (r / (1.0 + kappa * den_src))
    rho = np.exp(0.5 * tau * g(u, u)) * u
    buf = matfunc(A, rho, tau, "exp", m)
    if kind == "cubic":
        return np.exp(0.5 * tau * g(buf, buf)) * buf
    return np.exp(0.5 * tau * g(u, buf)) * buf


def sewi_step(A, u, uprev, mfield, dt, m=10, kind="cubic", s1=0.0, s2=0.0, kappa=0.0, host2d=False):
    """sEWI step (nlse_cubic_gautschi_solver_3d.hpp:8-35 and twins; host2d: the 2D cubic host
    convention of nlse_cubic_gautschi_solver.hpp:21-39).  Returns (u_new before BC, u_prev_new)."""
    tau = 1j * dt
    r = np.abs(u) ** 2
    if kind == "cubic":
        g = mfield * r
    elif kind == "cq":
        g = mfield * (s1 * r + s2 * r * r)
    else:
        g = mfield * (r / (1.0 + kappa * r))
    B = -g * u
    sg = -1.0 if host2d else 1.0
    phi = matfunc(A, B, dt, "sinc", m)
    a = matfunc(A, phi, sg * tau, "exp", m)
    b = matfunc(A, uprev, sg * 2 * tau, "exp", m)
    return b - sg * 2 * tau * a, u.copy()


_F = {"sge": np.sin, "kge": lambda x: x ** 3, "phi4": lambda x: x - x ** 3,
      "sge_double": lambda x: np.sin(x) + .6 * np.sin(2. * x), "sge_sinh": np.sinh}


def gautschi_step(A, u, up, mfield, tau, m=10, kind="sge"):
    """SGESolver::step sg_single_solver.hpp:43-59 (mod-cosine filter); id_sqrt family kg_solver.hpp:13-22."""
    filt = "mod_cosine" if kind == "sge" else "id_sqrt"
    b = matfunc(A, u, tau, filt, m)
    b = -mfield * _F[kind](b)
    b = matfunc(A, b, tau, "sinc2_half", m)
    un = 2 * matfunc(A, u, tau, "cos_sqrt", m) - up + tau * tau * b
    return un, u.copy()


def sv_step(A, u, up, mfield, tau, kind="sge"):
    """SGESolverSV::step sg_single_sv_solver.hpp:9-19 / KGESVSolver::step kg_sv_solver.hpp:13-21."""
    b = A @ u - mfield * _F[kind](u)
    return 2 * u - up + tau * tau * b, u.copy()


def downsample_fft(u, target_shape):
    """finalized_scripts/downsampling.py:10-34 restated with numpy.fft: ortho fft2 over the last two axes, fftshift,
    centred crop to target_shape, ifftshift, ortho ifft2; real input keeps the real part."""
    u = np.asarray(u)
    n0, n1 = u.shape[-2:]
    t0, t1 = target_shape
    ft = np.fft.fftshift(np.fft.fft2(u.astype(np.complex128), norm="ortho", axes=(-2, -1)), axes=(-2, -1))
    s0, s1 = (n0 - t0) // 2, (n1 - t1) // 2
    crop = np.fft.ifftshift(ft[..., s0:s0 + t0, s1:s1 + t1], axes=(-2, -1))
    out = np.fft.ifft2(crop, norm="ortho", axes=(-2, -1))
    return out if np.iscomplexobj(u) else out.real
