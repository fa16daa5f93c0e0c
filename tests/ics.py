"""Deterministic synthetic inputs: random Gaussian bumps, radial kink."""
import numpy as np


def grid2d(n, L):
    x = np.linspace(-L, L, n)
    return np.meshgrid(x, x)  # X[j,i] = x_i, Y[j,i] = y_j  -> shape (ny, nx)


def gaussian_bumps_2d(n, L, seed=1234, K=5):
    """C1 initial condition: K Gaussian bumps with plane-wave phases, max|u0| = 1."""
    rng = np.random.default_rng(seed)
    X, Y = grid2d(n, L)
    u = np.zeros((n, n), dtype=np.complex128)
    for _ in range(K):
        x0, y0 = rng.uniform(-0.6 * L, 0.6 * L, 2)
        s = rng.uniform(0.6, 1.2)
        a = rng.uniform(0.5, 1.0)
        kx, ky = rng.uniform(-1, 1, 2)
        u += a * np.exp(-((X - x0) ** 2 + (Y - y0) ** 2) / (2 * s * s)) * np.exp(1j * (kx * X + ky * Y))
    return u / np.abs(u).max()


def gaussian_bumps_3d(n, L, seed=99, K=8):
    rng = np.random.default_rng(seed)
    x = np.linspace(-L, L, n)
    Z, Y, X = np.meshgrid(x, x, x, indexing="ij")
    u = np.zeros((n, n, n), dtype=np.complex128)
    for _ in range(K):
        c = rng.uniform(-0.5 * L, 0.5 * L, 3)
        s = rng.uniform(0.5, 1.0)
        a = rng.uniform(0.5, 1.0)
        k = rng.uniform(-1, 1, 3)
        u += a * np.exp(-((X - c[0]) ** 2 + (Y - c[1]) ** 2 + (Z - c[2]) ** 2) / (2 * s * s)) \
            * np.exp(1j * (k[0] * X + k[1] * Y + k[2] * Z))
    return u / np.abs(u).max()


def kink_2d(n, L, seed=77, bumps=3):
    """C2 initial condition: elliptic radial kink (root sg_driver.cpp:62-64) + small bumps."""
    rng = np.random.default_rng(seed)
    X, Y = grid2d(n, L)
    u = 4.0 * np.arctan(np.exp(3.0 - 5.0 * np.sqrt((X - Y) ** 2 / 3.0 + (X + Y) ** 2 / 2.0)))
    for _ in range(bumps):
        x0, y0 = rng.uniform(-0.5 * L, 0.5 * L, 2)
        s = rng.uniform(0.5, 1.0)
        u += 0.1 * np.exp(-((X - x0) ** 2 + (Y - y0) ** 2) / (2 * s * s))
    return u


def slab_ic(n, z0, nz, nzg, L):
    """seed-free analytic IC of the C5 workload (same formula as bench.py): 8 fixed Gaussians."""
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
