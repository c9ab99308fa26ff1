"""Pins the plain-C oracle (oracle/euler_oracle.c) against the reference's OWN templates compiled
into oracle/_ref/libref_*.so (oracle/ref/harness.cpp; built from /root/reference by oracle/Makefile).

CPU only.  The _ref libraries are built in the dev container and travel to the GPU box as binaries;
if they are absent (a checkout that never had /root/reference) these tests skip and
tests/test_golden.py still pins the oracle through the committed fixtures.
"""
import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle, refharness
from tests import common

# rounding-level agreement: both sides are FP64 evaluations of the same formulas in (almost) the
# same order; 1e-12 relative to the column's largest entry leaves room for Eigen's reassociations.
TOL = 1e-12


def _need(scheme, degree):
    if not refharness.available(scheme, degree):
        pytest.skip("oracle/_ref not built (needs /root/reference)")


@pytest.mark.parametrize("scheme", [abi.SCHEME_FR, abi.SCHEME_DGSEM])
@pytest.mark.parametrize("degree", [1, 2, 3])
@pytest.mark.parametrize("distortion", [0.0, 0.15])
def test_residual(lib, scheme, degree, distortion):
    _need(scheme, degree)
    meshes = common.make_parts((4, 3, 3), (0, 0, 0), (10, 9, 8), (1, 1, 1), distortion)
    views = [m.view for m in meshes]
    ref = refharness.RefSim(views, scheme, degree)
    cpu = oracle.OracleSim(views, scheme, degree)
    for sim in (ref, cpu):
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    assert common.rel_err(cpu.get_solution(0), ref.get_solution(0)) < 1e-15
    u = common.perturbed(ref.get_solution(0))
    ref.set_solution([u]); cpu.set_solution([u])
    common.assert_parity(cpu.get_residual(0.25)[0], ref.get_residual(0.25)[0], TOL)


@pytest.mark.parametrize("scheme,degree,rk_order", [(abi.SCHEME_FR, 2, 3), (abi.SCHEME_FR, 1, 2), (abi.SCHEME_FR, 3, 3),
                                                    (abi.SCHEME_DGSEM, 2, 3), (abi.SCHEME_DGSEM, 1, 1),
                                                    (abi.SCHEME_DGSEM, 3, 2)])
def test_rk_steps(lib, scheme, degree, rk_order):
    _need(scheme, degree)
    meshes = common.make_parts((4, 4, 3), (0, 0, 0), (10, 10, 10), (1, 1, 1), 0.1)
    views = [m.view for m in meshes]
    ref = refharness.RefSim(views, scheme, degree, 0, rk_order)
    cpu = oracle.OracleSim(views, scheme, degree, rk_order)
    for sim in (ref, cpu):
        common.set_vortex_case(sim, wall_groups=())
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    ref.rk_update(0.0, 5e-3, 6)
    cpu.rk_update(0.0, 5e-3, 6)
    assert common.rel_err(cpu.get_solution(0), ref.get_solution(0)) < common.NOISE_TOL  # z-momentum is 1e-3 of the rest
    assert common.col_err(cpu.get_solution(0), ref.get_solution(0)) < TOL


@pytest.mark.parametrize("kind", range(6))
def test_boundary_kinds_in_residual(lib, kind):
    _need(abi.SCHEME_FR, 2)
    meshes = common.make_parts((3, 3, 2), (0, 0, 0), (10, 10, 10), (1, 1, 1), 0.1)
    views = [m.view for m in meshes]
    ref = refharness.RefSim(views, abi.SCHEME_FR, 2)
    cpu = oracle.OracleSim(views, abi.SCHEME_FR, 2)
    for sim in (ref, cpu):
        for g in range(6):
            sim.set_boundary(g, kind, abi.FUNC_VORTEX, common.VORTEX)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    u = common.perturbed(ref.get_solution(0), amp=1e-2)
    ref.set_solution([u]); cpu.set_solution([u])
    common.assert_parity(cpu.get_residual(0.5)[0], ref.get_residual(0.5)[0], TOL)


@pytest.mark.parametrize("parts", [(2, 1, 1), (2, 2, 1), (2, 2, 2)])
def test_partitioned_mesh(lib, parts):
    """Ghost exchange + ghost faces: reference parts run as threads (its own overlap order)."""
    _need(abi.SCHEME_FR, 2)
    meshes = common.make_parts((4, 4, 4), (0, 0, 0), (10, 10, 10), parts, 0.1)
    views = [m.view for m in meshes]
    ref = refharness.RefSim(views, abi.SCHEME_FR, 2)
    cpu = oracle.OracleSim(views, abi.SCHEME_FR, 2)
    for sim in (ref, cpu):
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    ref.rk_update(0.0, 5e-3, 3)
    cpu.rk_update(0.0, 5e-3, 3)
    for r in range(len(meshes)):
        common.assert_parity(cpu.get_solution(r), ref.get_solution(r), TOL)


def test_partition_does_not_change_the_answer(lib):
    """One part vs 2x2x2 parts of the same mesh, gathered by global cell id (oracle only)."""
    n = (4, 4, 4)
    whole = common.make_parts(n, (0, 0, 0), (10, 10, 10), (1, 1, 1), 0.1)
    split = common.make_parts(n, (0, 0, 0), (10, 10, 10), (2, 2, 2), 0.1)
    sims = [oracle.OracleSim([m.view for m in ms], abi.SCHEME_FR, 2) for ms in (whole, split)]
    for sim in sims:
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
        sim.rk_update(0.0, 5e-3, 3)
    fields = 135
    full = np.empty((64, fields))
    gid = whole[0].cell_global_ids()[:64]
    full[gid] = sims[0].get_solution(0).reshape(-1, fields)
    for r, m in enumerate(split):
        ids = m.cell_global_ids()[:m.n_local_cells]
        part = sims[1].get_solution(r).reshape(-1, fields)
        # holder/sharer roles change with the partition, i.e. left/right of the Riemann problem swap and
        # Newton starts from the other side's pressure (exact.hpp:139); it stops at |f| < 1e-8
        # (exact.hpp:163-174), so fluxes move by up to ~1e-8 and the solution by ~1e-8 * dt per step
        assert common.rel_err(part, full[ids]) < 1e-6  # worst component: z-momentum, 1e-3 of the others (measured 2.7e-7)
        assert common.col_err(part, full[ids]) < 1e-9


def _random_states(rng, n, lo=0.1, hi=2.0, vmax=1.5):
    rho = rng.uniform(lo, hi, n); p = rng.uniform(lo, hi, n)
    vel = rng.uniform(-vmax, vmax, (n, 3))
    e = p / 0.4 + 0.5 * rho * (vel ** 2).sum(1)
    return np.ascontiguousarray(np.column_stack([rho, rho[:, None] * vel, e]))


def _random_frames(rng, n):
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    q[:, :, 2] *= np.sign(np.linalg.det(q))[:, None]
    return np.ascontiguousarray(np.transpose(q, (0, 2, 1)).reshape(n, 9))


def test_riemann_flux_random(lib):
    _need(abi.SCHEME_FR, 2)
    rng = np.random.default_rng(3)
    n = 20000
    ul, ur, fr = _random_states(rng, n), _random_states(rng, n), _random_frames(rng, n)
    a = oracle.riemann_flux(common.GAMMA, fr, ul, ur)
    b = refharness.riemann_flux(fr, ul, ur)
    scale = np.abs(b).max(axis=1, keepdims=True)
    assert (np.abs(a - b) / scale).max() < 1e-12


def test_riemann_flux_strong_and_vacuum(lib):
    """Strong rarefactions (near-vacuum / vacuum branches, exact.hpp:87-126,315-342) and strong shocks."""
    _need(abi.SCHEME_FR, 2)
    rng = np.random.default_rng(5)
    n = 4000
    fr = _random_frames(rng, n)
    ul, ur = _random_states(rng, n, 0.05, 1.0, 0.5), _random_states(rng, n, 0.05, 1.0, 0.5)
    # push the two sides apart / together along the face normal
    push = rng.uniform(-6.0, 6.0, n)
    for u, sgn in ((ul, -1.0), (ur, +1.0)):
        rho = u[:, 0].copy()
        v = u[:, 1:4] / rho[:, None]
        p = 0.4 * (u[:, 4] - 0.5 * rho * (v ** 2).sum(1))
        v = v + sgn * push[:, None] * fr[:, 0:3]
        u[:, 1:4] = rho[:, None] * v
        u[:, 4] = p / 0.4 + 0.5 * rho * (v ** 2).sum(1)
    a = oracle.riemann_flux(common.GAMMA, fr, ul, ur)
    b = refharness.riemann_flux(fr, ul, ur)
    scale = np.maximum(np.abs(b).max(axis=1, keepdims=True), 1e-30)
    assert (np.abs(a - b) / scale).max() < 1e-12


def test_boundary_flux_random(lib):
    _need(abi.SCHEME_FR, 2)
    rng = np.random.default_rng(11)
    n = 5000
    ui, uo, fr = _random_states(rng, n, 0.5, 1.5, 2.0), _random_states(rng, n, 0.5, 1.5, 2.0), _random_frames(rng, n)
    for kind in range(6):
        a = oracle.boundary_flux(kind, common.GAMMA, fr, ui, uo)
        b = refharness.boundary_flux(kind, fr, ui, uo)
        assert common.rel_err(a, b) < 1e-14, kind


@pytest.mark.parametrize("scheme,degree", [(abi.SCHEME_FR, 2), (abi.SCHEME_DGSEM, 2), (abi.SCHEME_DG_MODAL, 2)])
def test_oracle_vacuum_branch_follows_the_reference_member(lib, scheme, degree):
    """Exact::star_u is a MEMBER of the Riemann object (exact.hpp:64): a vacuum solve upwinds (v, w) by the contact speed
    of the last ordinary solve of the same face point"""
    if not refharness.available(scheme, degree, 0):
        pytest.skip("oracle/_ref not built here (needs /root/reference)")
    mesh = abi.BoxMesh((4, 2, 2), (0, 0, 0), (4.0, 2.0, 2.0))
    ref = refharness.RefSim([mesh.view], scheme, degree, 0, 3)
    cpu = oracle.OracleSim([mesh.view], scheme, degree, 3)
    a, col = common.vacuum_sequence(cpu, mesh, scheme, degree, lambda c: cpu.set_solution([c]), lambda: cpu.get_residual(0.0)[0])
    b, _ = common.vacuum_sequence(ref, mesh, scheme, degree, lambda c: ref.set_solution([c]), lambda: ref.get_residual(0.0)[0])
    assert common.col_err(a, b) < 1e-10 and common.transverse_momentum_err(a, b) < 1e-10
    fresh = oracle.OracleSim([mesh.view], scheme, degree, 3)   # the second solve alone: star_u of a fresh object is 0
    for g in range(6):
        fresh.set_boundary(g, abi.BC_SOLID_WALL)
    fresh.set_solution([col])
    assert common.transverse_momentum_err(fresh.get_residual(0.0)[0], b) > 1e-2   # the case does exercise the member
