"""Parity at the sizes the headline numbers are quoted on ("What's missing" 6).

* 64 x 32 x 32 cells (the sub-box bench.py's reference arm runs), FR p = 2, 3 RK3 steps: the CUDA path against the
  COMPILED REFERENCE (oracle/_ref, the unmodified templates) on one part -- 65 536 cells, 8.8 M DOF, ~1 min of CPU.
* 16^3 cells, FR p = 3 (configs[4]'s scheme), 3 RK3 steps against the compiled reference.
* the full 64^3 of configs[1]: size-independent properties the scheme guarantees -- a closed box (slip walls all
  round) conserves mass and energy to rounding under the Lobatto rule, and the solution stays finite.
Without oracle/_ref (it travels with the snapshot; it cannot be rebuilt on the GPU box) the plain-C oracle stands in.
"""
import math

import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle, refharness
from tests import common

pytestmark = pytest.mark.gpu


def _checker(views, scheme, degree, rk):
    if refharness.available(scheme, degree):
        return refharness.RefSim(views, scheme, degree, 0, rk), "compiled reference"
    return oracle.OracleSim(views, scheme, degree, rk), "oracle"


@pytest.mark.parametrize("n,degree,steps", [((64, 32, 32), 2, 3), ((16, 16, 16), 3, 3)])
def test_fr_steps_at_bench_size_vs_reference(lib, n, degree, steps):
    h = 10.0 / 64 if degree == 2 else 10.0 / 16
    mesh = abi.BoxMesh(n, (0, 0, 0), (n[0] * h, n[1] * h, n[2] * h))
    gpu = abi.Solver(mesh.view, scheme=abi.SCHEME_FR, degree=degree, rk_order=3)
    ref, who = _checker([mesh.view], abi.SCHEME_FR, degree, 3)
    for sim in (gpu, ref):
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    a0, b0 = gpu.get_solution(), ref.get_solution(0)
    assert common.rel_err(a0, b0) < 1e-14, who                   # Cell::Approximate
    dt = 1e-3 if degree == 2 else 2e-3
    gpu.rk_update(0.0, dt, steps)
    ref.rk_update(0.0, dt, steps)
    a, b = gpu.get_solution(), ref.get_solution(0)
    assert np.isfinite(a).all()
    per_comp, per_dof = common.parity(a, b)
    print(f"{n} p={degree}: vs {who}: per component {per_comp:.2e}, per DOF {per_dof:.2e}, column {common.col_err(a, b):.2e}")
    # the planar vortex leaves z-momentum at the noise level of the exact solver's 1e-8 stop (common.NOISE_TOL)
    assert common.strong_err(a, b) < 1e-10, who
    assert per_comp < common.NOISE_TOL, who


def _lobatto_weights(q):
    return {2: [1.0, 1.0], 3: [1 / 3, 4 / 3, 1 / 3], 4: [1 / 6, 5 / 6, 5 / 6, 1 / 6]}[q]


def test_closed_box_conserves_mass_and_energy_at_full_size(lib):
    """64^3 FR p = 2 (35 389 440 DOF, the bench workload), slip walls on all six sides, the vortex as initial data:
    sum_q w_q U^_q (= the integral of u, U^ = det J u) of mass and energy may only move by rounding."""
    n, q = 64, 3
    mesh = abi.BoxMesh((n, n, n), (0, 0, 0), (10, 10, 10))
    gpu = abi.Solver(mesh.view, scheme=abi.SCHEME_FR, degree=2, rk_order=3)
    for g in range(6):
        gpu.set_boundary(g, abi.BC_SOLID_WALL)
    gpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    w1 = np.array(_lobatto_weights(q))
    w = np.einsum("i,j,k->ijk", w1, w1, w1).ravel()

    def integrals():   # per cell, then an exactly rounded sum over the cells (a plain sum of 7 M terms drifts by 1e-11)
        per_cell = np.einsum("cnk,n->ck", gpu.get_solution().reshape(-1, q ** 3, 5), w)
        return np.array([math.fsum(per_cell[:, k]) for k in range(5)])

    before = integrals()
    gpu.rk_update(0.0, 1e-3, 5)
    u = gpu.get_solution()
    assert np.isfinite(u).all()
    after = integrals()
    for comp in (0, 4):
        assert abs(after[comp] - before[comp]) <= 1e-14 * abs(before[comp]), (comp, before, after)
    assert np.abs(gpu.get_solution() - u).max() == 0.0
