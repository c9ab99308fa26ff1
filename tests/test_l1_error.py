"""Part::MeasureL1Error (mesh/part.hpp:1082-1096) on the device, and the one check of a run that
does not lean on any oracle: the measured convergence order of the isentropic vortex (a smooth exact solution).

CPU: the oracle's restatement against the reference's own loop (compiled, oracle/_ref).  GPU: mcfd_measure_l1_error
against the oracle, then mesh refinement at p = 1, 2, 3 for FR and modal DG.
"""
import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle, refharness
from tests import common

SCHEMES = [(abi.SCHEME_FR, 2), (abi.SCHEME_DGSEM, 3), (abi.SCHEME_DG_MODAL, 2), (abi.SCHEME_FR, 1)]


def _vortex(sim, t=0.0):
    common.set_vortex_case(sim)
    sim.approximate(abi.FUNC_VORTEX, common.VORTEX, t)


@pytest.mark.parametrize("scheme,degree", SCHEMES)
def test_oracle_l1_error_vs_compiled_reference(lib, scheme, degree):
    if not refharness.available(scheme, degree):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    mesh = abi.BoxMesh((4, 3, 3), (0, 0, 0), (10, 9, 8), distortion=0.12)
    ref = refharness.RefSim([mesh.view], scheme, degree)
    cpu = oracle.OracleSim([mesh.view], scheme, degree)
    for sim in (ref, cpu):
        _vortex(sim)
    a, b = cpu.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.3), ref.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.3)
    assert b[0] > 1e-3                                   # the vortex has moved: a real error, not rounding
    assert np.abs(a - b).max() <= 1e-13 * np.abs(b).max()
    # against the function it was projected from the nodal schemes are exact at their own nodes
    zero = ref.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    if scheme != abi.SCHEME_DG_MODAL:
        assert np.abs(zero).max() < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("scheme,degree", SCHEMES + [(abi.SCHEME_DG_MODAL, 3), (abi.SCHEME_DG_MODAL, 1)])
def test_gpu_l1_error_matches_oracle(lib, oracle_lib, scheme, degree):
    mesh = abi.BoxMesh((5, 4, 3), (0, 0, 0), (10, 9, 8), distortion=0.12)
    gpu = abi.Solver(mesh.view, scheme=scheme, degree=degree)
    cpu = oracle.OracleSim([mesh.view], scheme, degree)
    for sim in (gpu, cpu):
        _vortex(sim)
    for t in (0.0, 0.3):
        a, b = gpu.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, t), cpu.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, t)
        # (at t = 0 the nodal schemes are exact at their own nodes: both sums are rounding, ~1e-14)
        assert np.abs(a - b).max() <= 1e-12 * np.abs(b).max() + 1e-12, (t, a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("scheme,degree,cells", [(abi.SCHEME_FR, 1, (32, 64)), (abi.SCHEME_FR, 2, (12, 24)),
                                                 (abi.SCHEME_FR, 3, (16, 32)), (abi.SCHEME_DGSEM, 2, (12, 24)),
                                                 (abi.SCHEME_DG_MODAL, 1, (16, 32)), (abi.SCHEME_DG_MODAL, 2, (12, 24)),
                                                 (abi.SCHEME_DG_MODAL, 3, (8, 16))])
def test_vortex_convergence_order(lib, scheme, degree, cells):
    """Isentropic vortex advected to t = 0.2 on n x n x 2 cubes of [0,10]^2 (exact state on the x/y boundaries, slip
    walls on z), mesh halved: the L1 errors must fall at the scheme's order -- a property that no oracle enters.
    Observed on B200 (scripts/convergence_probe.py): modal DG p + 1 (1.9, 3.0, 3.8 at p = 1, 2, 3); the schemes
    collocated on Lobatto points (FR with Huynh's lumped-Lobatto correction, DGSEM) p + 1/2 .. p + 1 (1.5-1.7, 2.5-2.6,
    3.1-3.4), the known price of their inexact mass matrix.  Bounds: p + 0.6 and p + 0.35."""
    errs = []
    for n in cells:
        h = 10.0 / n
        mesh = abi.BoxMesh((n, n, 2), (0, 0, 0), (10, 10, 2 * h))
        gpu = abi.Solver(mesh.view, scheme=scheme, degree=degree, rk_order=3)
        _vortex(gpu)
        steps = 8 * n // cells[0] * 5
        dt = 0.2 / steps
        gpu.rk_update(0.0, dt, steps)
        e = gpu.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.2)
        assert np.isfinite(e).all()
        errs.append(e / (2 * h))                       # per unit depth: the two meshes differ in z extent
    order = np.log2(errs[0] / errs[1])
    print(f"scheme {scheme} p={degree}: L1(rho) {errs[0][0]:.3e} -> {errs[1][0]:.3e}, orders {np.round(order[[0, 1, 2, 4]], 2)}")
    bound = degree + (0.6 if scheme == abi.SCHEME_DG_MODAL else 0.35)
    for c in (0, 1, 2, 4):
        assert order[c] > bound, (c, order)
