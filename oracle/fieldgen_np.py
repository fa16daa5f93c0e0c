"""numpy twin of the on-device field generators (k_generate in nls_kernels.cuh; include/nls.h nls_generate_field).
TEST INFRASTRUCTURE ONLY: evaluates a term table exactly as the kernel does, on the engine's axes (x = fastest)."""
import numpy as np


def coords(n, L):
    return np.linspace(-L, L, n)


def evaluate(terms, post, pp, nx, ny, nz=1, L=(1.0, 1.0, 1.0), cplx=False, c_field=None, z0=0, nz_glob=None):
    Lx, Ly, Lz = L
    x, y = coords(nx, Lx), coords(ny, Ly)
    dim3 = nz > 1
    if dim3:
        z = coords(nz_glob or nz, Lz)[z0:z0 + nz]
        Z, Y, X = np.meshgrid(z, y, x, indexing="ij")
    else:
        Y, X = np.meshgrid(y, x, indexing="ij")
        Z = 0.0
    if post == 4:  # GRADIENT_MASS (m_fields_2d.py:38-49), script axis 0 = engine y
        c = np.asarray(c_field, dtype=np.float64).reshape(ny, nx)
        gy, gx = np.zeros_like(c), np.zeros_like(c)
        gy[1:-1, :] = (c[2:, :] - c[:-2, :]) / 2
        gx[:, 1:-1] = (c[:, 2:] - c[:, :-2]) / 2
        gs = gy ** 2 + gx ** 2
        return pp[0] * (1 + pp[1] * gs / (gs + pp[2] ** 2))
    v = np.zeros(X.shape, dtype=np.complex128)
    for kind, ip, p in terms:
        p = list(p) + [0.0] * (10 - len(p))
        if kind == 0:
            v += p[0]
        elif kind == 1:
            r2 = (X - p[1]) ** 2 + (Y - p[2]) ** 2 + ((Z - p[3]) ** 2 if dim3 else 0.0)
            v += p[0] * np.exp(-r2 / (2 * (p[4] * p[4])))
        elif kind == 2:
            kd = p[1] * X + p[2] * Y + (p[3] * Z if dim3 else 0.0)
            a = p[5] * kd + p[4]
            v += p[0] * (np.sin(a) if ip else np.cos(a))
        elif kind == 3:
            b = [np.sqrt(X * X + Y * Y) - p[2], np.maximum(np.abs(X), np.abs(Y)) - p[2], X, Y, Y + X][min(ip, 4)]
            v += p[0] + (p[1] - p[0]) * (0.5 * (1 + np.tanh(b / p[3])))
        elif kind == 4:
            v += p[0] * (1 + p[1] * (np.sin(p[2] * Y) * np.sin(p[2] * X)))
        elif kind == 5:
            cc = Y if p[3] != 0 else X
            lw = 2 * p[2] / ip
            t = np.full(X.shape, p[0])
            for i in range(1, ip, 2):
                t[(cc >= -p[2] + i * lw) & (cc < -p[2] + (i + 1) * lw)] = p[0] * p[1]
            v += t
        elif kind == 6:
            pat = np.sin(np.pi * Y / p[1]) * np.sin(np.pi * X / p[1])
            v += p[0] * (np.tanh(p[2] * pat) if p[2] > 0 else np.sign(pat))
        elif kind == 7:
            v += p[0] * np.tanh(p[1] * (Y if p[3] != 0 else X) / p[2])
        elif kind == 8:
            r2 = (X - p[1]) ** 2 + (Y - p[2]) ** 2 + ((Z - p[3]) ** 2 if dim3 else 0.0)
            ph = p[5] * X + p[6] * Y + (p[7] * Z if dim3 else 0.0)
            v += p[0] * np.exp(-r2 / (2 * (p[4] * p[4]))) * np.exp(1j * ph)
        else:
            r = np.sqrt((X - Y) ** 2 / 3.0 + (X + Y) ** 2 / 2.0)
            v += p[0] * np.arctan(np.exp(p[1] - p[2] * r))
    if post == 1:
        v = np.maximum(v.real, pp[0]) + 0j
    elif post == 2:
        v = pp[0] * ((v.real - v.real.min()) / (v.real.max() - v.real.min())) + 0j
    elif post == 3:
        v = v / np.abs(v).max()
    return v if cplx else v.real
