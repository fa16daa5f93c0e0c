"""Golden vectors for the on-device field generators: the reference's own generator functions
(finalized_scripts/m_fields_2d.py, c_fields_2d.py) are imported from /root/reference, run with a fixed numpy seed,
and their outputs stored next to the term table that nonlinear-solvers_b200/fieldgen.py builds from the SAME seed
(it mirrors the scripts' RNG call order).  The GPU box has no /root/reference: tests/test_fieldgen.py only reads
the .npz written here.     python tests/golden/make_field_golden.py"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF = "/root/reference/nlsolvers/finalized_scripts"
sys.path.insert(0, REF)
# the scripts import matplotlib at module level only for their plotting helpers
mpl = types.ModuleType("matplotlib")
mpl.pyplot = types.ModuleType("matplotlib.pyplot")
sys.modules.setdefault("matplotlib", mpl)
sys.modules.setdefault("matplotlib.pyplot", mpl.pyplot)
import m_fields_2d as refm  # noqa: E402
import c_fields_2d as refc  # noqa: E402
import nls_b200  # noqa: E402
from nls_b200 import fieldgen  # noqa: E402

N, L = 64, 10.0
CASES = [  # (name, reference function, builder, kwargs, seed)
    ("m_constant", refm.constant_mass, fieldgen.constant_mass, dict(base_value=1.3), 1),
    ("m_piecewise_circle", refm.piecewise_constant_mass, fieldgen.piecewise_constant_mass,
     dict(m1=1.0, m2=2.2, boundary_type="circle", boundary_param=0.45, smooth_width=0.04), 2),
    ("m_piecewise_square", refm.piecewise_constant_mass, fieldgen.piecewise_constant_mass,
     dict(m1=1.0, m2=1.7, boundary_type="square", boundary_param=0.6, smooth_width=0.08), 3),
    ("m_piecewise_horizontal", refm.piecewise_constant_mass, fieldgen.piecewise_constant_mass,
     dict(m1=0.8, m2=2.5, boundary_type="horizontal", smooth_width=0.02), 4),
    ("m_piecewise_vertical", refm.piecewise_constant_mass, fieldgen.piecewise_constant_mass,
     dict(m1=0.8, m2=2.5, boundary_type="vertical", smooth_width=0.05), 5),
    ("m_piecewise_diagonal", refm.piecewise_constant_mass, fieldgen.piecewise_constant_mass,
     dict(m1=1.0, m2=3.0, boundary_type="diagonal", smooth_width=0.1), 6),
    ("m_defects", refm.localized_mass_defects, fieldgen.localized_mass_defects, dict(m0=1.0, num_defects=12), 7),
    ("m_quasiperiodic", refm.quasi_periodic_mass, fieldgen.quasi_periodic_mass, dict(m0=1.2, num_waves=5), 8),
    ("c_constant", refc.constant, fieldgen.constant, dict(), 9),
    ("c_periodic", refc.periodic_structure, fieldgen.periodic_structure, dict(base_value=1.0, amplitude=0.5, frequency=3), 10),
    ("c_piecewise_layers", refc.piecewise_constant, fieldgen.piecewise_constant, dict(num_layers=5, contrast_factor=2.5), 11),
    ("c_checkerboard", refc.sign_changing_mass, fieldgen.sign_changing_mass, dict(regions="checkerboard", scale=2, sharpness=5.0), 12),
    ("c_checkerboard_sign", refc.sign_changing_mass, fieldgen.sign_changing_mass, dict(regions="checkerboard", scale=3, sharpness=0.0), 13),
    ("c_half_space", refc.sign_changing_mass, fieldgen.sign_changing_mass, dict(regions="half_space", sharpness=4.0), 14),
    ("c_layered", refc.layered_materials, fieldgen.layered_materials, dict(num_layers=4), 15),
    ("c_quasiperiodic", refc.quasi_periodic_structures, fieldgen.quasi_periodic_structures, dict(num_waves=6), 16),
]


def pack(terms):
    kind = np.array([t[0] for t in terms], dtype=np.int32)
    ip = np.array([t[1] for t in terms], dtype=np.int32)
    p = np.zeros((len(terms), 10))
    for r, t in enumerate(terms):
        p[r, :len(t[2])] = t[2]
    return kind, ip, p


def main():
    out = {"names": np.array([c[0] for c in CASES] + ["m_gradient"]), "n": N, "L": L}
    for name, ref, build, kw, seed in CASES:
        np.random.seed(seed)
        expected = np.asarray(ref(N, L, **kw), dtype=np.float64)
        np.random.seed(seed)
        terms, post, pp = build(N, L, **kw)
        kind, ip, p = pack(terms)
        out[f"{name}/kind"], out[f"{name}/iparam"], out[f"{name}/p"] = kind, ip, p
        out[f"{name}/post"], out[f"{name}/post_params"] = post, np.array(pp, dtype=np.float64)
        out[f"{name}/expected"] = expected
    # gradient_based_mass needs a c field: the periodic structure above
    c_field = refc.periodic_structure(N, L, 1.0, 0.5, 3)
    out["m_gradient/c_field"] = c_field
    out["m_gradient/expected"] = refm.gradient_based_mass(N, L, c_field, m0=1.1, gamma=1.5, epsilon=2e-2)
    out["m_gradient/post_params"] = np.array([1.1, 1.5, 2e-2])
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "fields_2d.npz"), **out)
    print("wrote fields_2d.npz:", len(CASES) + 1, "cases")


if __name__ == "__main__":
    main()
