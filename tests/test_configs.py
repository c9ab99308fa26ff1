"""BASELINE.json's named configurations as parity cases.

configs[0]  demo/euler sod: 5 x 1 x 0.5 shock tube, 50 x 10 x 5 hexahedra, Left/Right smart boundaries with the
            initial states, walls elsewhere (demo/euler/shock_tube_sod.cpp:5-38), 2 parts as `mpirun -n 2`.
            The shipped demo is FR p=2 + RK3; "DG p=1" exists as the modal dg::General with RK2.
configs[2]  double Mach reflection: 4 x 1 x 0.2 domain, Mach-10 shock at 60 degrees from x = 1/6, moving-shock
            supersonic inlets on the top and left, outlet on the right and on the bottom "Gap" x < 1/6, walls on
            the rest (demo/euler/double_mach.cpp:5-52), modal DG p=2 + weno::Eigen.

CPU tests hold the plain-C oracle to the COMPILED REFERENCE on these cases; the gpu tests hold the CUDA path
(through the C-ABI) to the oracle.  Tolerance: 1e-10 relative per column (north_star).
"""
import ctypes as C
import math

import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle, refharness
from tests import common

TUBE_N, TUBE_HI = (50, 10, 5), (5.0, 1.0, 0.5)
X_GAP = 1.0 / 6.0


def sod_tube_case(sim):
    sim.set_boundary(0, abi.BC_SMART, abi.FUNC_CONSTANT, common.SOD_LEFT)    # Left  (4_S_31)
    sim.set_boundary(1, abi.BC_SMART, abi.FUNC_CONSTANT, common.SOD_RIGHT)   # Right (4_S_23)
    for g in (2, 3, 4, 5):
        sim.set_boundary(g, abi.BC_SOLID_WALL)
    sim.approximate(abi.FUNC_PLANE_SPLIT, common.plane_split((1, 0, 0), 2.5, 0.0, common.SOD_LEFT, common.SOD_RIGHT), 0.0)


def double_mach_states():
    """double_mach.cpp:5-20"""
    gp, g, gm = 2.4, 1.4, 0.4
    rho_b, p_b, mach, a_b = 1.4, 1.0, 10.0, 1.0
    u_nu = mach * a_b
    rho_a = rho_b * (mach * mach * gp / 2) / (1 + mach * mach * gm / 2)
    p_a = p_b * (mach * mach * g - gm / 2) / (gp / 2)
    u_n_a = u_nu * (rho_a - rho_b) / rho_a
    tan60 = math.sqrt(3.0)
    cos30, sin30 = tan60 * 0.5, 0.5
    after = common.prim_to_cons(rho_a, u_n_a * cos30, u_n_a * (-sin30), 0.0, p_a)
    before = common.prim_to_cons(rho_b, 0.0, 0.0, 0.0, p_b)
    return after, before, tan60, u_nu / cos30


def with_gap(mesh):
    """Boundary group 6 = the bottom faces left of x = 1/6 (section "Gap", double_mach.cpp:49)."""
    a = mesh.arrays()
    v = mesh.view
    nb = v.n_boundary_faces
    if nb:
        first = v.n_local_faces + v.n_ghost_faces
        centre = a["node_xyz"][a["face_nodes"][first:first + nb]].mean(axis=1)
        group = a["bface_group"].astype(np.int32).copy()
        group[(group == 2) & (centre[:, 0] < X_GAP)] = 6
        mesh._gap_groups = np.ascontiguousarray(group)   # keeps the array alive as long as the mesh
        v.bface_group = mesh._gap_groups.ctypes.data_as(C.POINTER(C.c_int32))
    v.n_bc_groups = 7
    return mesh


def double_mach_case(sim):
    after, before, tan60, u_x = double_mach_states()
    # (x - (x_gap + u_x t)) tan60 < y  <=>  tan60 x - y < tan60 x_gap + tan60 u_x t
    shock = common.plane_split((tan60, -1.0, 0.0), tan60 * X_GAP, tan60 * u_x, after, before)
    sim.set_boundary(3, abi.BC_SUPERSONIC_INLET, abi.FUNC_PLANE_SPLIT, shock)   # Top
    sim.set_boundary(0, abi.BC_SUPERSONIC_INLET, abi.FUNC_PLANE_SPLIT, shock)   # Left
    sim.set_boundary(4, abi.BC_SOLID_WALL)                                      # Back
    sim.set_boundary(5, abi.BC_SOLID_WALL)                                      # Front
    sim.set_boundary(2, abi.BC_SOLID_WALL)                                      # Bottom
    sim.set_boundary(1, abi.BC_SUPERSONIC_OUTLET)                               # Right
    sim.set_boundary(6, abi.BC_SUPERSONIC_OUTLET)                               # Gap
    sim.approximate(abi.FUNC_PLANE_SPLIT, shock, 0.0)


DM_N, DM_HI = (48, 12, 2), (4.0, 1.0, 0.2)   # h = 1/12: x = 1/6 is a grid line


def dm_meshes(parts=(1, 1, 1)):
    world = parts[0] * parts[1] * parts[2]
    return [with_gap(abi.BoxMesh(DM_N, (0, 0, 0), DM_HI, parts=parts, rank=r)) for r in range(world)]


# ---- CPU: the oracle against the compiled reference --------------------------------------------------------------
@pytest.mark.parametrize("scheme,degree,rk", [(abi.SCHEME_FR, 2, 3), (abi.SCHEME_DG_MODAL, 1, 2)])
def test_config0_oracle_vs_reference_two_parts(lib, scheme, degree, rk):
    if not refharness.available(scheme, degree, 0):
        pytest.skip("oracle/_ref not built")
    meshes = common.make_parts(TUBE_N, (0, 0, 0), TUBE_HI, (2, 1, 1))
    views = [m.view for m in meshes]
    ref = refharness.RefSim(views, scheme, degree, 0, rk)
    cpu = oracle.OracleSim(views, scheme, degree, rk)
    for sim in (ref, cpu):
        sod_tube_case(sim)
        sim.rk_update(0.0, 1e-3, 3)
    for r in range(2):
        assert common.rel_err(cpu.get_solution(r), ref.get_solution(r)) < common.NOISE_TOL  # x-momentum is still weak
        assert common.col_err(cpu.get_solution(r), ref.get_solution(r)) < 1e-11


def test_config2_oracle_vs_reference(lib):
    if not refharness.available(abi.SCHEME_DG_MODAL, 2, abi.LIMITER_EIGEN):
        pytest.skip("oracle/_ref not built")
    meshes = dm_meshes((2, 1, 1))
    views = [m.view for m in meshes]
    ref = refharness.RefSim(views, abi.SCHEME_DG_MODAL, 2, abi.LIMITER_EIGEN, 3)
    cpu = oracle.OracleSim(views, abi.SCHEME_DG_MODAL, 2, 3, limiter=abi.LIMITER_EIGEN)
    for sim in (ref, cpu):
        double_mach_case(sim)
        sim.set_solution([sim.get_solution(r) for r in range(2)])   # SetSolutionColumn: first limiter pass
        sim.rk_update(0.0, 1e-4, 2)
    assert cpu.troubled_count() == ref.troubled_count() > 0
    for r in range(2):
        a, b = cpu.get_solution(r), ref.get_solution(r)
        assert np.isfinite(b).all()
        common.assert_parity(a, b, 1e-10)


# ---- configs[3]: forward-facing step on tetrahedra (demo/euler/forward_step.cpp:5-40), the reference-backed form: modal DG
# p = 2 + weno::Eigen + exact Riemann (FR is hexahedral and Hllc<Gas,3> does not exist in the reference).
# The impulsive start (Mach 3 against the step's front wall) makes troubled cells from the first stage on.
def step_meshes(parts=(1, 1, 1), n=5):
    import bench
    world = parts[0] * parts[1] * parts[2]
    return [bench.forward_step_mesh(abi, n, parts, r) for r in range(world)]


def forward_step_case(sim):
    import bench
    bench.forward_step_case(sim, abi)


def test_config3_oracle_vs_reference(lib):
    if not refharness.available(abi.SCHEME_DG_MODAL, 2, abi.LIMITER_EIGEN, cell=1):
        pytest.skip("oracle/_ref not built")
    meshes = step_meshes((3, 1, 1))
    views = [m.view for m in meshes]
    assert sum(v.n_local_cells for v in views) == 6 * (15 * 5 * 2 - 12 * 1 * 2)
    ref = refharness.RefSim(views, abi.SCHEME_DG_MODAL, 2, abi.LIMITER_EIGEN, 3)
    cpu = oracle.OracleSim(views, abi.SCHEME_DG_MODAL, 2, 3, limiter=abi.LIMITER_EIGEN)
    for sim in (ref, cpu):
        forward_step_case(sim)
        sim.set_solution([sim.get_solution(r) for r in range(3)])   # SetSolutionColumn: first limiter pass
        sim.rk_update(0.0, 2e-4, 5)
    assert cpu.troubled_count() == ref.troubled_count() > 0
    for r in range(3):
        a, b = cpu.get_solution(r), ref.get_solution(r)
        assert np.isfinite(b).all()
        common.assert_parity(a, b, 1e-10)


# ---- GPU: the CUDA path against the oracle -----------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("scheme,degree,rk", [(abi.SCHEME_FR, 2, 3), (abi.SCHEME_DG_MODAL, 1, 2), (abi.SCHEME_DGSEM, 1, 2)])
def test_config0_gpu_sod_tube(lib, scheme, degree, rk):
    mesh = abi.BoxMesh(TUBE_N, (0, 0, 0), TUBE_HI)
    gpu = abi.Solver(mesh.view, scheme=scheme, degree=degree, rk_order=rk)
    cpu = oracle.OracleSim([mesh.view], scheme, degree, rk)
    for sim in (gpu, cpu):
        sod_tube_case(sim)
        sim.rk_update(0.0, 1e-3, 20)   # README: t in [0, 0.2], 200 steps; the first 20 here
    a, b = gpu.get_solution(), cpu.get_solution(0)
    assert np.isfinite(a).all()
    common.assert_parity(a, b, 1e-10)


@pytest.mark.gpu
def test_config2_gpu_double_mach(lib):
    mesh = dm_meshes()[0]
    gpu = abi.Solver(mesh.view, scheme=abi.SCHEME_DG_MODAL, degree=2, rk_order=3, limiter=abi.LIMITER_EIGEN)
    cpu = oracle.OracleSim([mesh.view], abi.SCHEME_DG_MODAL, 2, 3, limiter=abi.LIMITER_EIGEN)
    double_mach_case(gpu)
    double_mach_case(cpu)
    gpu.set_solution(gpu.get_solution())
    cpu.set_solution([cpu.get_solution(0)])
    for sim in (gpu, cpu):
        sim.rk_update(0.0, 1e-4, 5)
    assert gpu.troubled_cells() == cpu.troubled_count() > 0
    a, b = gpu.get_solution(), cpu.get_solution(0)
    assert np.isfinite(a).all()
    common.assert_parity(a, b, 1e-10)


@pytest.mark.gpu
def test_config3_gpu_forward_step(lib):
    mesh = step_meshes()[0]
    gpu = abi.Solver(mesh.view, scheme=abi.SCHEME_DG_MODAL, degree=2, rk_order=3, limiter=abi.LIMITER_EIGEN)
    cpu = oracle.OracleSim([mesh.view], abi.SCHEME_DG_MODAL, 2, 3, limiter=abi.LIMITER_EIGEN)
    forward_step_case(gpu)
    forward_step_case(cpu)
    gpu.set_solution(gpu.get_solution())
    cpu.set_solution([cpu.get_solution(0)])
    for sim in (gpu, cpu):
        sim.rk_update(0.0, 2e-4, 10)
    assert gpu.troubled_cells() == cpu.troubled_count() > 0
    a, b = gpu.get_solution(), cpu.get_solution(0)
    assert np.isfinite(a).all()
    common.assert_parity(a, b, 1e-10)
