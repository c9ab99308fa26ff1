"""GPU parity tests: the CUDA path, called through the C-ABI, against the oracle on the same
seeded inputs.  Tolerances: north_star asks for a relative 1e-10 per DOF on solution columns;
single residual evaluations are held to 1e-10 relative to the column's largest entry as well.
"""
import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle
from tests import common

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _pair(mesh_views, scheme, degree, rk_order=3, **kw):
    gpu = [abi.Solver(v, scheme=scheme, degree=degree, rk_order=rk_order, rank=r, world=len(mesh_views), **kw)
           for r, v in enumerate(mesh_views)]
    cpu = oracle.OracleSim(mesh_views, scheme, degree, rk_order)
    return gpu, cpu


def _exchange_ghosts(gpu, meshes, columns):
    """What NCCL does in production, done by hand: ghost coefficients from the peers' columns."""
    for r, m in enumerate(meshes):
        a = m.arrays()
        fields = gpu[r].fields_per_cell
        chunks = []
        for i, peer in enumerate(a["nbr_rank"]):
            pa = meshes[peer].arrays()
            j = list(pa["nbr_rank"]).index(r)
            cells = pa["send_cells"][pa["send_ptr"][j]:pa["send_ptr"][j + 1]]
            chunks.append(columns[peer].reshape(-1, fields)[cells].ravel())
        gpu[r].set_ghost_solution(np.concatenate(chunks) if chunks else np.zeros(0))


@pytest.mark.parametrize("scheme", [abi.SCHEME_FR, abi.SCHEME_DGSEM])
@pytest.mark.parametrize("degree", [1, 2, 3])
@pytest.mark.parametrize("distortion", [0.0, 0.15])
def test_residual_matches_oracle(lib, scheme, degree, distortion):
    meshes = common.make_parts((5, 4, 3), (0, 0, 0), (10, 9, 8), (1, 1, 1), distortion)
    gpu, cpu = _pair([m.view for m in meshes], scheme, degree)
    for sim in (gpu[0], cpu):
        common.set_vortex_case(sim)
    cpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    u = common.perturbed(cpu.get_solution(0))
    cpu.set_solution([u])
    gpu[0].set_solution(u)
    gpu[0].set_time(0.25)
    r_gpu = gpu[0].get_residual()
    r_cpu = cpu.get_residual(0.25)[0]
    assert np.isfinite(r_gpu).all()
    common.assert_parity(r_gpu, r_cpu, TOL)


@pytest.mark.parametrize("scheme,degree,rk_order", [(abi.SCHEME_FR, 2, 3), (abi.SCHEME_FR, 1, 2), (abi.SCHEME_FR, 3, 3),
                                                    (abi.SCHEME_DGSEM, 2, 3), (abi.SCHEME_DGSEM, 1, 1),
                                                    (abi.SCHEME_DGSEM, 3, 2)])
def test_rk_steps_match_oracle(lib, scheme, degree, rk_order):
    meshes = common.make_parts((6, 5, 4), (0, 0, 0), (10, 10, 10), (1, 1, 1), 0.1)
    gpu, cpu = _pair([m.view for m in meshes], scheme, degree, rk_order)
    for sim in (gpu[0], cpu):
        common.set_vortex_case(sim, wall_groups=())
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    assert common.rel_err(gpu[0].get_solution(), cpu.get_solution(0)) < 1e-14  # Approximate itself
    gpu[0].rk_update(0.0, 5e-3, 10)
    cpu.rk_update(0.0, 5e-3, 10)
    common.assert_parity(gpu[0].get_solution(), cpu.get_solution(0), TOL)


def test_sod_walls_and_smart_boundaries(lib):
    """demo/euler/shock_tube_sod.cpp on a coarse 5 x 1 x 0.5 box: discontinuous IC, walls + smart ends."""
    meshes = common.make_parts((20, 4, 2), (0, 0, 0), (5, 1, 0.5), (1, 1, 1))
    gpu, cpu = _pair([m.view for m in meshes], abi.SCHEME_FR, 2, 3)
    ic = common.plane_split((1, 0, 0), 2.5, 0.0, common.SOD_LEFT, common.SOD_RIGHT)
    for sim in (gpu[0], cpu):
        sim.set_boundary(0, abi.BC_SMART, abi.FUNC_CONSTANT, common.SOD_LEFT)
        sim.set_boundary(1, abi.BC_SMART, abi.FUNC_CONSTANT, common.SOD_RIGHT)
        for g in (2, 3, 4, 5):
            sim.set_boundary(g, abi.BC_SOLID_WALL)
        sim.approximate(abi.FUNC_PLANE_SPLIT, ic, 0.0)
    assert np.array_equal(gpu[0].get_solution(), cpu.get_solution(0))
    gpu[0].rk_update(0.0, 1e-3, 5)
    cpu.rk_update(0.0, 1e-3, 5)
    common.assert_parity(gpu[0].get_solution(), cpu.get_solution(0), TOL)


@pytest.mark.parametrize("kind", [abi.BC_SOLID_WALL, abi.BC_SUPERSONIC_INLET, abi.BC_SUPERSONIC_OUTLET,
                                  abi.BC_SUBSONIC_INLET, abi.BC_SUBSONIC_OUTLET, abi.BC_SMART])
def test_every_boundary_kind(lib, kind):
    meshes = common.make_parts((3, 3, 3), (0, 0, 0), (10, 10, 10), (1, 1, 1), 0.1)
    gpu, cpu = _pair([m.view for m in meshes], abi.SCHEME_FR, 2)
    for sim in (gpu[0], cpu):
        for g in range(6):
            sim.set_boundary(g, kind, abi.FUNC_VORTEX, common.VORTEX)
    cpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    u = common.perturbed(cpu.get_solution(0), amp=1e-2)
    cpu.set_solution([u]); gpu[0].set_solution(u)
    gpu[0].set_time(0.5)
    common.assert_parity(gpu[0].get_residual(), cpu.get_residual(0.5)[0], TOL)


def test_unregistered_boundary_contributes_nothing(lib):
    meshes = common.make_parts((3, 3, 3), (0, 0, 0), (10, 10, 10), (1, 1, 1))
    gpu, cpu = _pair([m.view for m in meshes], abi.SCHEME_DGSEM, 2)
    for sim in (gpu[0], cpu):
        sim.set_boundary(0, abi.BC_SOLID_WALL)  # groups 1..5 never registered (fem.hpp:185-226)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    common.assert_parity(gpu[0].get_residual(), cpu.get_residual(0.0)[0], TOL)


def test_host_evaluated_boundary_values(lib):
    """The std::function route: the caller evaluates u(x, t) at mcfd_bc_points and uploads the values."""
    meshes = common.make_parts((4, 3, 3), (0, 0, 0), (10, 10, 10), (1, 1, 1), 0.1)
    gpu, cpu = _pair([m.view for m in meshes], abi.SCHEME_FR, 2)
    for g in range(6):
        cpu.set_boundary(g, abi.BC_SMART, abi.FUNC_PLANE_SPLIT,
                         common.plane_split((1, 1, 0), 9.0, 0.5, common.SOD_LEFT, common.SOD_RIGHT))
        gpu[0].set_boundary(g, abi.BC_SMART, abi.FUNC_HOST_VALUES)
        xyz = gpu[0].bc_points(g)
        s = xyz[:, 0] + xyz[:, 1]
        vals = np.where((s < 9.0 + 0.5 * 0.3)[:, None], np.array(common.SOD_LEFT), np.array(common.SOD_RIGHT))
        gpu[0].set_bc_values(g, vals)
    cpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    u = cpu.get_solution(0)
    gpu[0].set_solution(u)
    gpu[0].set_time(0.3)
    common.assert_parity(gpu[0].get_residual(), cpu.get_residual(0.3)[0], TOL)


@pytest.mark.parametrize("parts", [(2, 1, 1), (2, 2, 1), (1, 2, 2)])
def test_ghost_faces_on_partitioned_mesh(lib, parts):
    """Each part's residual with hand-fed ghost coefficients = the oracle's multi-part residual."""
    meshes = common.make_parts((4, 4, 4), (0, 0, 0), (10, 10, 10), parts, 0.1)
    gpu, cpu = _pair([m.view for m in meshes], abi.SCHEME_FR, 2)
    for sim in gpu + [cpu]:
        common.set_vortex_case(sim)
    cpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    cols = [common.perturbed(cpu.get_solution(r), seed=r + 1) for r in range(len(meshes))]
    cpu.set_solution(cols)
    for r, g in enumerate(gpu):
        g.set_solution(cols[r])
    _exchange_ghosts(gpu, meshes, cols)
    ref = cpu.get_residual(0.0)
    for r, g in enumerate(gpu):
        common.assert_parity(g.get_residual(), ref[r], TOL)


@pytest.mark.parametrize("riemann", [abi.RIEMANN_EXACT])
def test_riemann_flux_random_states(lib, riemann):
    rng = np.random.default_rng(7)
    n = 20000
    def states():
        rho = rng.uniform(0.1, 2.0, n); p = rng.uniform(0.1, 2.0, n)
        vel = rng.uniform(-1.5, 1.5, (n, 3))
        e = p / 0.4 + 0.5 * rho * (vel ** 2).sum(1)
        return np.ascontiguousarray(np.column_stack([rho, rho[:, None] * vel, e]))
    ul, ur = states(), states()
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    q[:, :, 2] *= np.sign(np.linalg.det(q))[:, None]
    frames = np.ascontiguousarray(np.transpose(q, (0, 2, 1)).reshape(n, 9))
    f_gpu = abi.riemann_flux(riemann, common.GAMMA, frames, ul, ur)
    f_cpu = oracle.riemann_flux(common.GAMMA, frames, ul, ur)
    scale = np.abs(f_cpu).max(axis=1, keepdims=True)
    assert (np.abs(f_gpu - f_cpu) / scale).max() < 1e-9  # Newton stops at |f| < 1e-8 (exact.hpp:163)
    assert np.median(np.abs(f_gpu - f_cpu) / scale) < 1e-14


def test_riemann_flux_degenerate_data(lib):
    """Left and right states equal up to a few ulp (smooth IC sampled at shared nodes): the sign of
    f(p_L) is rounding noise and selects FindRoot's halving branch (exact.hpp:163-166); the device must
    take the reference's branch, so agreement stays at rounding level, not at Newton's 1e-8."""
    rng = np.random.default_rng(29)
    n = 200000
    rho = rng.uniform(0.5, 1.5, n); p = rng.uniform(0.5, 1.5, n); vel = rng.uniform(-1.5, 1.5, (n, 3))
    ul = np.ascontiguousarray(np.column_stack([rho, rho[:, None] * vel, p / 0.4 + 0.5 * rho * (vel ** 2).sum(1)]))
    ur = ul.copy()
    for k in range(5):
        ur[:, k] = ul[:, k] * (1.0 + rng.integers(-3, 4, n) * 2.0 ** -53)
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    frames = np.ascontiguousarray(np.transpose(q, (0, 2, 1)).reshape(n, 9))
    f_gpu = abi.riemann_flux(abi.RIEMANN_EXACT, common.GAMMA, frames, ul, ur)
    f_cpu = oracle.riemann_flux(common.GAMMA, frames, ul, ur)
    err = np.abs(f_gpu - f_cpu).max(axis=1) / np.abs(f_cpu).max(axis=1)
    assert err.max() < 1e-12


def test_boundary_flux_random_states(lib):
    rng = np.random.default_rng(11)
    n = 5000
    rho = rng.uniform(0.5, 1.5, (2, n)); p = rng.uniform(0.5, 1.5, (2, n)); vel = rng.uniform(-2, 2, (2, n, 3))
    u = [np.ascontiguousarray(np.column_stack([rho[i], rho[i][:, None] * vel[i], p[i] / 0.4 + 0.5 * rho[i] * (vel[i] ** 2).sum(1)]))
         for i in range(2)]
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    frames = np.ascontiguousarray(np.transpose(q, (0, 2, 1)).reshape(n, 9))
    for kind in range(6):
        f_gpu = abi.boundary_flux(kind, common.GAMMA, frames, u[0], u[1])
        f_cpu = oracle.boundary_flux(kind, common.GAMMA, frames, u[0], u[1])
        assert common.rel_err(f_gpu, f_cpu) < 1e-13, kind


def test_error_paths(lib):
    mesh = abi.BoxMesh((2, 2, 2))
    with pytest.raises(RuntimeError):
        abi.Solver(mesh.view, degree=7)
    with pytest.raises(RuntimeError):
        abi.Solver(mesh.view, scheme=7)
    with pytest.raises(RuntimeError):
        abi.Solver(mesh.view, scheme=abi.SCHEME_DGSEM, limiter=abi.LIMITER_LAZY)  # limiters: modal scheme only
    s = abi.Solver(mesh.view)
    with pytest.raises(RuntimeError):
        s.set_solution(np.zeros(3))
    with pytest.raises(RuntimeError):
        s.set_boundary(9, abi.BC_SOLID_WALL)


@pytest.mark.parametrize("scheme,degree", [(abi.SCHEME_FR, 2), (abi.SCHEME_DGSEM, 2), (abi.SCHEME_FR, 3), (abi.SCHEME_FR, 1)])
@pytest.mark.parametrize("cells", [(3, 3, 3), (7, 3, 1)])
def test_cell_counts_that_do_not_fill_a_tile(lib, scheme, degree, cells):
    """27 and 21 cells: the last block of either cell kernel holds an odd number of cells, whose bytes are not a multiple
    of 16 at p = 2 -- the bulk-copy (TMA) path must hand over to plain copies there."""
    meshes = common.make_parts(cells, (0, 0, 0), (10, 9, 8), (1, 1, 1), 0.1)
    gpu, cpu = _pair([m.view for m in meshes], scheme, degree, 3)
    for sim in (gpu[0], cpu):
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    gpu[0].rk_update(0.0, 2e-3, 3)
    cpu.rk_update(0.0, 2e-3, 3)
    common.assert_parity(gpu[0].get_solution(), cpu.get_solution(0), TOL)
    gpu[0].set_time(6e-3)
    common.assert_parity(gpu[0].get_residual(), cpu.get_residual(6e-3)[0], TOL)


def test_trilinear_only_distortion_is_not_taken_for_affine(lib):
    """nodes displaced in a checkerboard pattern give every hexahedron a pure xi*eta*zeta term -- the
    Jacobian agrees at two opposite corners and differs at the other six.  Such a mesh needs per-node metrics; with a
    single cell-centre metric the residual would be visibly wrong."""
    mesh = abi.BoxMesh((3, 3, 3), (0, 0, 0), (3.0, 3.0, 3.0))
    v = mesh.view
    xyz = np.ctypeslib.as_array(v.node_xyz, shape=(v.n_nodes, 3))   # a view: the solver and the oracle read these
    ijk = np.rint(xyz).astype(int)
    xyz[:, 0] += 0.08 * (-1.0) ** ijk.sum(1)
    gpu, cpu = _pair([v], abi.SCHEME_FR, 2)
    for sim in (gpu[0], cpu):
        common.set_vortex_case(sim)
    cpu.approximate(abi.FUNC_VORTEX, [5.0, 1.5, 1.5, 1.0, 1.0, 1.4], 0.0)
    u = cpu.get_solution(0)
    gpu[0].set_solution(u)
    gpu[0].set_time(0.0)
    common.assert_parity(gpu[0].get_residual(), cpu.get_residual(0.0)[0], TOL)


# ---- Exact::star_u is a MEMBER of the Riemann object (exact.hpp:64): a vacuum solve upwinds (v, w) by the contact speed
# of the last ordinary solve of the same face point ----------------------------------------------------------------
@pytest.mark.parametrize("scheme,degree", [(abi.SCHEME_FR, 2), (abi.SCHEME_FR, 3), (abi.SCHEME_DGSEM, 2), (abi.SCHEME_DG_MODAL, 2)])
def test_vacuum_branch_uses_the_contact_speed_of_the_previous_solve(lib, scheme, degree):
    mesh = abi.BoxMesh((4, 2, 2), (0, 0, 0), (4.0, 2.0, 2.0))
    gpu = abi.Solver(mesh.view, scheme=scheme, degree=degree, rk_order=3)
    cpu = oracle.OracleSim([mesh.view], scheme, degree, 3)

    def gpu_residual():
        gpu.set_time(0.0)
        return gpu.get_residual()
    a, col = common.vacuum_sequence(gpu, mesh, scheme, degree, gpu.set_solution, gpu_residual)
    b, _ = common.vacuum_sequence(cpu, mesh, scheme, degree, lambda c: cpu.set_solution([c]), lambda: cpu.get_residual(0.0)[0])
    assert common.col_err(a, b) < 1e-10 and common.transverse_momentum_err(a, b) < 1e-10
    fresh = abi.Solver(mesh.view, scheme=scheme, degree=degree, rk_order=3)
    for g in range(6):
        fresh.set_boundary(g, abi.BC_SOLID_WALL)
    fresh.set_solution(col)
    fresh.set_time(0.0)
    assert common.transverse_momentum_err(fresh.get_residual(), b) > 1e-2   # a fresh object upwinds from the other side
