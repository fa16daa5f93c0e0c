"""Term tables for the on-device field generators (nls_generate_field).

Each builder mirrors one closed-form generator of the reference's finalized_scripts/m_fields_2d.py:14-147 or
c_fields_2d.py:12-169: it draws the random PARAMETERS from numpy's global RNG in exactly the order the script
does (so `np.random.seed(s)` + builder == `np.random.seed(s)` + script function) and returns
`(terms, post, post_params)` for `Engine.generate`.  The field itself is then evaluated on the device.

Axes: the scripts build `X, Y = meshgrid(x, y, indexing='ij')`, i.e. their X runs along the FIRST (slow) array
axis, which is the engine's y; their Y is the engine's x (fastest index).  The builders do that swap.

Not closed-form (n x n random arrays, image filters, point-set distances) and therefore host-side only:
multiscale_mass, topological_mass, phase_shifted_mass, wave_guiding_structures, turbulent_like_fields.
"""
import numpy as np

from .binding import GEN, GENPOST

NONE = (GENPOST.NONE, (0.0, 0.0, 0.0))


# ------------------------------------------------------------------ m fields (m_fields_2d.py)
def constant_mass(n, L, base_value=1.0):
    return [(GEN.CONST, 0, [base_value])], *NONE


def piecewise_constant_mass(n, L, m1=1.0, m2=2.0, boundary_type="circle", boundary_param=0.5, smooth_width=0.05):
    # script X -> engine y, script Y -> engine x:  'horizontal' (boundary = Y) is the engine's x axis
    kind = {"circle": 0, "square": 1, "horizontal": 2, "vertical": 3, "diagonal": 4}[boundary_type]
    return [(GEN.TANH_BOUNDARY, kind, [m1, m2, boundary_param * L, smooth_width * L])], *NONE


def localized_mass_defects(n, L, m0=1.0, num_defects=10, min_strength=-0.5, max_strength=1.0, min_width=0.05,
                           max_width=0.2):
    terms = [(GEN.CONST, 0, [m0])]
    for _ in range(num_defects):
        x0 = np.random.uniform(-L, L)   # script X centre -> engine y
        y0 = np.random.uniform(-L, L)
        strength = np.random.uniform(min_strength, max_strength) * m0
        width = np.random.uniform(min_width, max_width) * L
        terms.append((GEN.GAUSS, 0, [strength, y0, x0, 0.0, width]))
    return terms, GENPOST.CLAMP_MIN, (0.1 * m0, 0.0, 0.0)


def _golden_waves(num_waves):
    golden_ratio = (1 + np.sqrt(5)) / 2
    ks = []
    for i in range(num_waves):
        angle = i * np.pi / num_waves
        kx, ky = np.cos(angle), np.sin(angle)
        scale = golden_ratio ** i
        ks.append((kx * scale, ky * scale))
    return ks


def quasi_periodic_mass(n, L, m0=1.0, num_waves=5, min_amp=0.1, max_amp=0.5):
    terms = [(GEN.CONST, 0, [m0])]
    for kx, ky in _golden_waves(num_waves):
        amplitude = np.random.uniform(min_amp, max_amp) * m0
        phase = np.random.uniform(0, 2 * np.pi)
        terms.append((GEN.COSWAVE, 0, [amplitude, ky, kx, 0.0, phase, 1.0]))
    return terms, GENPOST.CLAMP_MIN, (0.1 * m0, 0.0, 0.0)


def gradient_based_mass(n, L, m0=1.0, gamma=1.0, epsilon=1e-6):
    """m from the handle's c field: call Engine.generate(FIELD.M, [], GENPOST.GRADIENT_MASS, (m0, gamma, epsilon))"""
    return [], GENPOST.GRADIENT_MASS, (m0, gamma, epsilon)


# ------------------------------------------------------------------ c fields (c_fields_2d.py)
def constant(n, L):
    return [(GEN.CONST, 0, [1.0])], *NONE


def periodic_structure(n, L, base_value=1.0, amplitude=0.5, frequency=3):
    k = np.pi * frequency / L
    return [(GEN.SINSIN, 0, [base_value, amplitude, k])], *NONE


def piecewise_constant(n, L, num_layers=3, base_value=1.0, contrast_factor=2.0):
    return [(GEN.LAYERS, num_layers, [base_value, contrast_factor, L, 1.0])], *NONE  # layers along script X = engine y


def sign_changing_mass(n, L, base_value=1.0, regions="checkerboard", scale=2, sharpness=5.0):
    if regions == "checkerboard":
        return [(GEN.CHECKER, 0, [base_value, L / scale, sharpness])], *NONE
    if regions == "half_space":
        return [(GEN.HALFSPACE, 0, [base_value, sharpness, L, 1.0])], *NONE
    raise NotImplementedError(regions)


def layered_materials(n, L, base_value=1.0, num_layers=3, min_amplitude=0.2, max_amplitude=0.8, min_freq=2, max_freq=10):
    terms = [(GEN.CONST, 0, [base_value])]
    for _ in range(num_layers):
        theta = np.random.uniform(0, np.pi)
        nx, ny = np.cos(theta), np.sin(theta)
        amplitude = np.random.uniform(min_amplitude, max_amplitude)
        freq = np.random.uniform(min_freq, max_freq)
        phase = np.random.uniform(0, 2 * np.pi)
        terms.append((GEN.COSWAVE, 1, [amplitude, ny, nx, 0.0, phase, freq]))  # amplitude sin(freq (nx X + ny Y) + phase)
    return terms, GENPOST.MINMAX, (base_value, 0.0, 0.0)


def quasi_periodic_structures(n, L, base_value=1.0, num_waves=5, min_amp=0.1, max_amp=0.5):
    terms = [(GEN.CONST, 0, [base_value])]
    for kx, ky in _golden_waves(num_waves):
        amplitude = np.random.uniform(min_amp, max_amp)
        phase = np.random.uniform(0, 2 * np.pi)
        terms.append((GEN.COSWAVE, 0, [amplitude, ky, kx, 0.0, phase, 1.0]))
    return terms, GENPOST.MINMAX, (base_value, 0.0, 0.0)


# ------------------------------------------------------------------ initial conditions of the bench / tests
def gaussian_bumps_2d(n, L, seed=1234, K=5):
    """tests/ics.py gaussian_bumps_2d (C1 / C4 initial condition): K bumps with plane-wave phases, max|u0| = 1"""
    rng = np.random.default_rng(seed)
    terms = []
    for _ in range(K):
        x0, y0 = rng.uniform(-0.6 * L, 0.6 * L, 2)
        s = rng.uniform(0.6, 1.2)
        a = rng.uniform(0.5, 1.0)
        kx, ky = rng.uniform(-1, 1, 2)
        terms.append((GEN.GAUSS_WAVE, 0, [a, x0, y0, 0.0, s, kx, ky, 0.0]))
    return terms, GENPOST.MAXABS, (0.0, 0.0, 0.0)


def slab_ic(L):
    """tests/ics.py slab_ic (C5 initial condition): 8 fixed Gaussians with a phase along x, seed-free"""
    terms = []
    for k in range(8):
        c = 0.5 * L * np.array([np.cos(1.3 * k), np.sin(2.1 * k), np.cos(0.7 * k + 1.0)])
        s = 0.8 + 0.05 * k
        terms.append((GEN.GAUSS_WAVE, 0, [0.5 + 0.05 * k, c[0], c[1], c[2], s, 0.3 * k, 0.0, 0.0]))
    return terms, *NONE


def kink_2d(n, L):
    """the radial kink of the C2 initial condition (root sg_driver.cpp:62-64), without the random bumps"""
    return [(GEN.KINK, 0, [4.0, 3.0, 5.0])], *NONE
