#!/usr/bin/env python
"""Generates tests/golden/riemann_approx.npz: golden vectors for the approximate Riemann solvers.

Sources, all the REFERENCE's own code run in the dev container (never shipped):
  roe_*   : /root/reference/python/riemann.py:397-443 (class Roe, 1-D, NumPy) imported as it lies
  ausm_*  : Ausm<Gas,3>::GetFluxUpwind (riemann/euler/ausm.hpp:153-217) through oracle/_ref/libref_fr_p2.so
  hllc_*  : Hllc<Gas,2>::GetFluxUpwind (riemann/euler/hllc.hpp:100-180) through the same library
States are normal-frame primitives [n][5] = rho, u, v, w, p.  Seeded; re-running reproduces the file.

  python tests/golden/make_golden_riemann_approx.py
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import refharness  # noqa: E402

REF_PY = "/root/reference/python"


def known_cases():
    """the left/right primitives of test/riemann/euler/{ausm,hllc}.cpp:35-110 (Sod, shock collision, blast waves)"""
    rows = [((1.0, 0.0, 1.0), (0.125, 0.0, 0.1)), ((0.125, 0.0, 0.1), (1.0, 0.0, 1.0)),
            ((5.99924, 19.5975, 460.894), (5.99242, 6.19633, 46.0950)),
            ((1.0, 0.0, 1000.0), (1.0, 0.0, 0.01)), ((1.0, 0.0, 0.01), (1.0, 0.0, 100.0)),
            ((1.0, -2.0, 0.4), (1.0, 2.0, 0.4))]
    left = np.array([[r[0][0], r[0][1], 1.5, -0.7, r[0][2]] for r in rows])
    right = np.array([[r[1][0], r[1][1], 2.5, 0.3, r[1][2]] for r in rows])
    return left, right


def random_cases(rng, n):
    def side():
        rho = rng.uniform(0.1, 3.0, n); p = rng.uniform(0.1, 3.0, n)
        vel = rng.uniform(-2.5, 2.5, (n, 3))  # up to Mach ~ 5 either way: every branch of the three solvers
        return np.column_stack([rho, vel, p])
    return side(), side()


def main():
    rng = np.random.default_rng(20241105)
    kl, kr = known_cases()
    rl, rr = random_cases(rng, 400)
    left = np.ascontiguousarray(np.vstack([kl, rl])); right = np.ascontiguousarray(np.vstack([kr, rr]))
    out = {"left": left, "right": right}
    out["ausm_flux"] = refharness.ausm3_flux(left, right)
    l2, r2 = left.copy(), right.copy()
    l2[:, 3] = 0.0; r2[:, 3] = 0.0
    out["hllc_flux_w0"] = refharness.hllc2_flux(l2, r2)  # valid for w = 0 on both sides
    # Roe, 1-D: v = w = 0
    sys.path.insert(0, REF_PY)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import riemann as ref_riemann  # noqa: E402
    roe = ref_riemann.Roe(1.4)
    eq = roe.equation()
    flux = np.zeros((left.shape[0], 3))
    for i in range(left.shape[0]):
        ul = eq.primitive_to_conservative(left[i, 0], left[i, 1], left[i, 4])
        ur = eq.primitive_to_conservative(right[i, 0], right[i, 1], right[i, 4])
        flux[i] = roe.get_upwind_flux(np.array(ul, dtype=float), np.array(ur, dtype=float))
    out["roe_flux_1d"] = flux  # mass, momentum, energy
    np.savez(os.path.join(HERE, "riemann_approx.npz"), **out)
    print("wrote riemann_approx.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
