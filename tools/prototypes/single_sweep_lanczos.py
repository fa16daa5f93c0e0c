"""Numerics prototype of the SINGLE-SWEEP Lanczos iteration ("Still open"): numpy only.

two_pass()     -- the scheme the CUDA kernels implement today: un-normalised basis C (C[0] = start, C[j+1] = w_j as it left
                  the orthogonalisation), pass A = stencil + recurrence term + dots, pass B = Gram-corrected block update.
single_sweep() -- the fused form: sweep j reads C[0..j-1] and the RAW vector r_{j-1} = L C[j-1] once, forms
                  C[j] = r/beta_{j-1} - (beta_{j-1}/beta_{j-2}) C[j-2] - sum_i h_i C[i], r_j = L C[j], and the raw dots
                  C[i]^H r_j, the Gram entries C[i]^H C[j] and ||C[j]||^2; the scalars of the next sweep follow from
                  those without touching the vectors again.
Both return (V, alpha, beta) in the normalised convention of the reference (oracle_np.lanczos) so that they can be compared
with it.  tests/test_single_sweep_prototype.py holds the comparison: if the fused form did not track the reference's MGS
sweep at rounding level on white noise, building the kernel would be pointless.
"""
import numpy as np


def _finish(C, nrm2, alpha, m):
    beta = np.sqrt(np.array(nrm2[:m]))
    V = np.stack([C[i] / beta[i] for i in range(m)], axis=1)
    return V, np.array(alpha), beta


def two_pass(A, u, m):
    C = [u.copy()]
    nrm2 = [np.vdot(u, u).real]
    alpha, g = [], []
    for j in range(m - 1):
        bj = np.sqrt(nrm2[j])
        w = A @ (C[j] * (1.0 / bj))
        if j > 0:
            w = w + (-bj / np.sqrt(nrm2[j - 1])) * C[j - 1]
        s = [np.vdot(C[i], w) for i in range(j + 1)]
        alpha.append(s[j].real / bj)
        q = s[j].real / nrm2[j]
        h = [(s[i] - q * g[i]) / nrm2[i] for i in range(j)] + [s[j] / nrm2[j]]
        for i in range(j + 1):
            w = w - h[i] * C[i]
        g = [np.vdot(C[i], w) for i in range(j + 1)]
        C.append(w)
        nrm2.append(np.vdot(w, w).real)
    return _finish(C, nrm2, alpha, m)


def single_sweep(A, u, m):
    C = [u.copy()]
    nrm2 = [np.vdot(u, u).real]
    r = A @ C[0]                      # sweep 0: raw stencil + raw dot
    q = [np.vdot(C[0], r)]
    gram = {0: []}                    # gram[j][i] = C[i]^H C[j], i < j
    alpha = []
    for j in range(1, m):
        # ---- scalar prologue of sweep j (what a kernel's first warp does from the reduced sums of sweep j-1)
        b1 = np.sqrt(nrm2[j - 1])
        inv = 1.0 / b1
        cmul = -b1 / np.sqrt(nrm2[j - 2]) if j >= 2 else 0.0
        s = []
        for i in range(j):
            if j < 2:
                G = 0.0
            elif i < j - 2:
                G = gram[j - 2][i]
            elif i == j - 2:
                G = nrm2[j - 2]
            else:
                G = np.conj(gram[j - 1][j - 2])
            s.append(inv * q[i] + cmul * G)
        alpha.append(s[j - 1].real / b1)
        qq = s[j - 1].real / nrm2[j - 1]
        h = [(s[i] - qq * gram[j - 1][i]) / nrm2[i] for i in range(j - 1)] + [s[j - 1] / nrm2[j - 1]]
        # ---- the sweep: one pass over C[0..j-1] and r
        w = r * inv
        if j >= 2:
            w = w + cmul * C[j - 2]
        for i in range(j):
            w = w - h[i] * C[i]
        C.append(w)
        nrm2.append(np.vdot(w, w).real)
        gram[j] = [np.vdot(C[i], w) for i in range(j)]
        if j < m - 1:                 # the last sweep needs no stencil (T[m-1, m-1] stays 0)
            r = A @ w
            q = [np.vdot(C[i], r) for i in range(j + 1)]
    return _finish(C, nrm2, alpha, m)


def gram_delayed(A, u, m, exact_h=False):
    """Gram-delayed re-orthogonalisation: per iteration ONE narrow stencil pass S(j) (reads C[j], C[j-1]; writes the raw
    r_j = L C[j]; true dots only with C[j] and C[j-1]) and ONE update pass U(j+1) (reads r_j and C[0..j]; writes C[j+1];
    ||C[j+1]||^2 and the Gram entries C[i]^H C[j+1]).  The projections on the older columns are not measured on r_j but
    derived from the Gram entries of C[j] through the recurrence L C[i] = beta_i (C[i+1] - cmul_i C[i-1] + (alpha_i /
    beta_i) C[i]):  j + 5 vector passes per iteration instead of 2j + 5.
    exact_h: keep every stored update coefficient in that recurrence (not only the alpha term)."""
    C = [u.copy()]
    nrm2 = [np.vdot(u, u).real]
    gram = {0: []}
    alpha = []
    hs = {}                                   # hs[i] = coefficients used to form C[i+1] (exact_h)
    cmuls = {}
    for j in range(m - 1):
        # ---- S(j)
        r = A @ C[j]
        a = np.vdot(C[j], r)
        b = np.vdot(C[j - 1], r) if j > 0 else 0.0
        # ---- scalar prologue of U(j+1)
        bj = np.sqrt(nrm2[j])
        inv = 1.0 / bj
        cmul = -bj / np.sqrt(nrm2[j - 1]) if j > 0 else 0.0
        cmuls[j] = cmul
        g = gram[j]                           # g[i] = C[i]^H C[j], i < j
        G = lambda k: g[k] if k < j else nrm2[j]   # C[k]^H C[j] for k <= j
        s = [0.0] * (j + 1)
        s[j] = inv * a + (cmul * np.conj(g[j - 1]) if j > 0 else 0.0)
        if j > 0:
            s[j - 1] = inv * b + cmul * nrm2[j - 1]
        for i in range(j - 1):                # i <= j-2: estimated from Gram entries
            bi = np.sqrt(nrm2[i])
            est = G(i + 1) + (alpha[i] / bi) * G(i)
            if i > 0:
                est = est - cmuls[i] * G(i - 1)
            if exact_h:
                for k in range(i):
                    est = est + np.conj(hs[i][k]) * G(k)
            raw = bi * est                    # ~ C[i]^H r_j
            gprev = gram[j - 1][i] if i < j - 1 else nrm2[j - 1]
            s[i] = inv * raw + cmul * gprev
        alpha.append(s[j].real / bj)
        q = s[j].real / nrm2[j]
        h = [(s[i] - q * g[i]) / nrm2[i] for i in range(j)] + [s[j] / nrm2[j]]
        hs[j] = h
        # ---- U(j+1)
        w = r * inv
        if j > 0:
            w = w + cmul * C[j - 1]
        for i in range(j + 1):
            w = w - h[i] * C[i]
        C.append(w)
        nrm2.append(np.vdot(w, w).real)
        gram[j + 1] = [np.vdot(C[i], w) for i in range(j + 1)]
    return _finish(C, nrm2, alpha, m)
