"""mcfd_measure_l1_error_static: the L1 error of a run for a caller that is not allowed to allocate.

Who that is: an error monitor called in the middle of a run by a host code that has already given the rest of the
GPU's memory to its own arena (so a cudaMalloc inside the library would fail), or that shares the GPU with peers
spinning on this GPU's halo counters (cudaMalloc / cudaFree stall the whole device), or whose own integrator owns all
three stage columns -- the ordinary entry point parks its per-block partial sums in an idle stage column and adds
them in a host vector.  The variant keeps them in a fixed 2.6 MB array inside the library and adds them on the device,
in the same order, so the five sums are the ordinary entry point's.

This file is named to be collected last: `pytest -x` stops at the first failure, and a failure here must not hide the
parity tests.  It holds no state of its own, so it passes alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

from minicfd_b200 import abi
from tests import common

TOL = 1e-12   # the bound tests/test_l1_error.py uses for this quantity (by construction the sums agree bit for bit)


def _close(a, b):
    return np.abs(a - b).max() <= TOL * np.abs(b).max() + TOL


def _solver(n, scheme, degree, distortion=0.12):
    mesh = abi.BoxMesh(n, (0, 0, 0), (10, 9, 8), distortion=distortion)
    sim = abi.Solver(mesh.view, scheme=scheme, degree=degree)
    common.set_vortex_case(sim)
    sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    return mesh, sim


@pytest.mark.gpu
@pytest.mark.parametrize("scheme,degree", [(abi.SCHEME_FR, 2), (abi.SCHEME_DGSEM, 3), (abi.SCHEME_DG_MODAL, 2),
                                           (abi.SCHEME_DG_MODAL, 1)])
def test_static_scratch_l1_error_equals_the_allocating_entry_point(lib, scheme, degree):
    """Two calls in a row on a large part, then two on a small one (which leaves most of the shared array holding the
    large part's partial sums): every result equals mcfd_measure_l1_error's at the same time."""
    big_mesh, big = _solver((9, 8, 7), scheme, degree)
    small_mesh, small = _solver((3, 2, 2), scheme, degree)
    for sim in (big, small):
        for t in (0.3, 0.1):                        # two calls in a row, different answers
            want = sim.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, t)
            got = sim.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, t)
            assert want[0] > 1e-4, want             # a real error (the vortex has moved), not rounding
            assert _close(got, want), (t, got, want)
        again = sim.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 0.3)
        assert np.array_equal(again, sim.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 0.3))
        assert _close(again, sim.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.3))
    # the solution column is an input: the monitor must not have touched it
    before = big.get_solution()
    big.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 0.2)
    assert np.array_equal(before, big.get_solution())


@pytest.mark.gpu
def test_static_scratch_l1_error_after_time_steps(lib):
    """The use it is meant for: between Runge-Kutta steps, where the ordinary entry point's scratch column is the
    integrator's.  The two agree after the steps, and the steps that follow are unchanged by the monitor."""
    _, a = _solver((6, 6, 2), abi.SCHEME_FR, 2, distortion=0.0)
    _, b = _solver((6, 6, 2), abi.SCHEME_FR, 2, distortion=0.0)
    dt = 0.01
    a.rk_update(0.0, dt, 3)
    b.rk_update(0.0, dt, 3)
    got = a.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 3 * dt)
    want = b.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 3 * dt)
    assert _close(got, want), (got, want)
    a.rk_update(3 * dt, dt, 2)
    b.rk_update(3 * dt, dt, 2)
    assert np.array_equal(a.get_solution(), b.get_solution())


@pytest.mark.gpu
def test_static_scratch_l1_error_refuses_what_it_cannot_hold(lib):
    """Above 16 777 216 quadrature points the array is too small: a clear error, no partial answer.  A nodal solver
    whose corner table was never uploaded is refused as well (uploading it would be an allocation)."""
    mesh = abi.BoxMesh((66, 64, 64), (0, 0, 0), (10, 9, 8))           # 270 336 cells x 64 nodes = 17.3 M points
    sim = abi.Solver(mesh.view, scheme=abi.SCHEME_FR, degree=3)
    common.set_vortex_case(sim)
    sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    with pytest.raises(RuntimeError, match="16777216"):
        sim.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 0.1)
    assert sim.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.1)[0] > 0    # the ordinary entry point still serves it
    sim.close()

    mesh = abi.BoxMesh((3, 3, 3), (0, 0, 0), (10, 9, 8))
    cold = abi.Solver(mesh.view, scheme=abi.SCHEME_FR, degree=2)
    common.set_vortex_case(cold)
    cold.set_solution(np.ones(cold.n))
    with pytest.raises(RuntimeError, match="corner table"):
        cold.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 0.1)
    want = cold.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.1)         # uploads the table
    assert _close(cold.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 0.1), want)
