"""Tetrahedral parts (BASELINE configs[3] in its reference-backed form: TETRA_4 + modal DG +
weno::Eigen + exact Riemann).

What the reference builds for a tetrahedron (mesh/part.hpp:437-452, 663-670, 726-736): geometry::Tetrahedron4 with
gauss::Tetrahedron<4|14|24> for degree 1|2|3 and geometry::Triangle3 with gauss::Triangle<3|6|12> on its faces, the
same polynomial::Projection (orthonormalised Taylor basis) as on hexahedra.  The meshes are boxes of hexahedra cut
into six tetrahedra each (mcfd_box_mesh_split_tetra; gmsh, which meshes input/forward_step.geo, is absent here).

CPU (`-m "not gpu"`): the generator's invariants; the C oracle against the COMPILED REFERENCE (oracle/_ref/libref_dgtet_*)
on fresh inputs and against the fixtures tests/golden/dgtet_*.npz that tests/golden/make_golden.py wrote from it.
GPU (`-m gpu`): the CUDA path through the C-ABI against the fixtures and the oracle, same tolerances as tests/test_modal.py.
"""
import os

import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle, refharness
from tests import common
from tests.test_modal import sod_split

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["dgtet_p2", "dgtet_p1", "dgtet_p3", "dgtet_lazy_p2", "dgtet_eigen_p2", "dgtet_eigen_p3", "dgtet_eigen_p2_2parts"]
HI = (1.2, 1.0, 0.8)


def tetra_parts(n, lo, hi, parts=(1, 1, 1), distortion=0.0, hole=None):
    world = parts[0] * parts[1] * parts[2]
    return [abi.BoxMesh(n, lo, hi, parts=parts, rank=r, distortion=distortion, hole=hole, tetra=True) for r in range(world)]


def set_walls(sim, n_groups=6):
    for g in range(n_groups):
        sim.set_boundary(g, abi.BC_SOLID_WALL)


# ---- the generator ---------------------------------------------------------------------------------------------
def test_tetra_split_is_conforming_and_oriented(lib):
    m = abi.BoxMesh((4, 3, 2), (0, 0, 0), (2.0, 1.5, 1.0), distortion=0.2, tetra=True)
    v, a = m.view, m.arrays()
    assert v.cell_type == abi.CELL_TETRA_4 and v.n_local_cells == 6 * 24 and v.n_ghost_cells == 0
    # every triangle belongs to two tetrahedra or to the boundary
    assert 4 * v.n_local_cells == 2 * v.n_local_faces + v.n_boundary_faces
    assert v.n_boundary_faces == 2 * 2 * (4 * 3 + 4 * 2 + 3 * 2)
    x = a["node_xyz"][a["cell_nodes"]]                    # [cell][4][3]
    vol6 = np.einsum("ci,ci->c", np.cross(x[:, 1] - x[:, 0], x[:, 2] - x[:, 0]), x[:, 3] - x[:, 0])
    assert (vol6 > 0).all()                               # CGNS TETRA_4 orientation
    assert abs(vol6.sum() / 6 - 2.0 * 1.5 * 1.0) < 1e-12   # the tetrahedra fill the box
    # boundary triangles lie on the plane of their group and the groups cover the six sides equally per area
    tri = a["node_xyz"][a["face_nodes"][v.n_local_faces:]]
    for g, (axis, value) in enumerate([(0, 0.0), (0, 2.0), (1, 0.0), (1, 1.5), (2, 0.0), (2, 1.0)]):
        sel = tri[a["bface_group"] == g]
        assert len(sel) and np.abs(sel[:, :, axis] - value).max() < 1e-14
    # interior faces: holder < sharer, both contain the three face nodes
    for f in range(v.n_local_faces):
        h, s = a["face_holder"][f], a["face_sharer"][f]
        assert h < s and set(a["face_nodes"][f]) <= set(a["cell_nodes"][h]) and set(a["face_nodes"][f]) <= set(a["cell_nodes"][s])


def test_tetra_split_of_a_partitioned_box_keeps_the_part_layout(lib):
    hexa = common.make_parts((4, 2, 2), (0, 0, 0), (4, 2, 2), (2, 1, 1))
    tets = tetra_parts((4, 2, 2), (0, 0, 0), (4, 2, 2), (2, 1, 1))
    for hm, tm in zip(hexa, tets):
        hv, tv, ta = hm.view, tm.view, tm.arrays()
        assert (tv.n_local_cells, tv.n_inner_cells, tv.n_ghost_cells) == (6 * hv.n_local_cells, 6 * hv.n_inner_cells, 6 * hv.n_ghost_cells)
        assert list(ta["recv_ptr"]) == [6 * x for x in hm.arrays()["recv_ptr"]]
        assert ((ta["send_cells"] >= tv.n_inner_cells) & (ta["send_cells"] < tv.n_local_cells)).all()
        ghost_f = slice(tv.n_local_faces, tv.n_local_faces + tv.n_ghost_faces)
        assert (ta["face_holder"][ghost_f] >= tv.n_inner_cells).all() and (ta["face_sharer"][ghost_f] >= tv.n_local_cells).all()
        # inner tetrahedra touch no ghost
        local_f = slice(0, tv.n_local_faces)
        assert (ta["face_sharer"][local_f] < tv.n_local_cells).all()
    # what rank 0 sends is what rank 1 receives, tetrahedron by tetrahedron (same global ids, same order)
    g0, g1 = tets[0].cell_global_ids(), tets[1].cell_global_ids()
    a0, a1 = tets[0].arrays(), tets[1].arrays()
    sent = g0[a0["send_cells"]]
    received = g1[tets[1].view.n_local_cells:]
    assert list(sent) == list(received)


def test_box_with_a_hole_has_a_seventh_boundary_group(lib):
    # the forward-facing step of demo/euler/forward_step.cpp in miniature: 15 x 5 x 1 cells, the step takes x >= 3, y < 1
    m = abi.BoxMesh((15, 5, 1), (0, 0, 0), (3.0, 1.0, 0.2), hole=((3, 0, 0), (15, 1, 1)))
    v, a = m.view, m.arrays()
    assert v.n_local_cells == 15 * 5 - 12 and v.n_bc_groups == 7
    step = a["node_xyz"][a["face_nodes"][v.n_local_faces:][a["bface_group"] == 6]]
    assert len(step) == 12 + 1   # the top of the step and its front
    assert ((step[:, :, 0] >= 0.6 - 1e-12) & (step[:, :, 1] <= 0.2 + 1e-12)).all()
    t = abi.BoxMesh((15, 5, 1), (0, 0, 0), (3.0, 1.0, 0.2), hole=((3, 0, 0), (15, 1, 1)), tetra=True)
    assert t.view.n_local_cells == 6 * v.n_local_cells and (t.arrays()["bface_group"] == 6).sum() == 2 * 13


# ---- oracle against the compiled reference ---------------------------------------------------------------------
@pytest.mark.parametrize("degree,limiter,parts", [(1, 0, (1, 1, 1)), (2, 0, (1, 1, 1)), (3, 0, (1, 1, 1)), (2, 1, (1, 1, 1)),
                                                  (2, 2, (2, 1, 1)), (3, 2, (1, 1, 1))])
def test_oracle_tetra_vs_compiled_reference(lib, degree, limiter, parts):
    if not refharness.available(abi.SCHEME_DG_MODAL, degree, limiter, cell=1):
        pytest.skip("oracle/_ref not built here (needs /root/reference)")
    meshes = tetra_parts((4, 3, 2), (0, 0, 0), HI, parts, 0.12)
    views = [m.view for m in meshes]
    ref = refharness.RefSim(views, abi.SCHEME_DG_MODAL, degree, limiter, 3)
    cpu = oracle.OracleSim(views, abi.SCHEME_DG_MODAL, degree, 3, limiter=limiter)
    for sim in (ref, cpu):
        set_walls(sim)
        sim.approximate(abi.FUNC_PLANE_SPLIT, sod_split(HI, weak=(limiter == 0)), 0.0)
    for r in range(len(views)):
        assert common.rel_err(cpu.get_solution(r), ref.get_solution(r)) < 1e-13  # Projection::Approximate
    cols = [ref.get_solution(r) for r in range(len(views))]
    ref.set_solution(cols)
    cpu.set_solution(cols)
    res_ref, res_cpu = ref.get_residual(0.0), cpu.get_residual(0.0)
    for r in range(len(views)):
        assert common.col_err(res_cpu[r], res_ref[r]) < 1e-8   # discontinuous data: the exact solver's own tolerance
    ref.rk_update(0.0, 1e-3, 3)
    cpu.rk_update(0.0, 1e-3, 3)
    for r in range(len(views)):
        assert common.rel_err(cpu.get_solution(r), ref.get_solution(r)) < common.NOISE_TOL  # weak transverse momenta
        assert common.col_err(cpu.get_solution(r), ref.get_solution(r)) < 1e-10
    if limiter:
        assert cpu.troubled_count() == ref.troubled_count() > 0


def test_oracle_tetra_vortex_vs_compiled_reference(lib):
    if not refharness.available(abi.SCHEME_DG_MODAL, 2, 0, cell=1):
        pytest.skip("oracle/_ref not built here (needs /root/reference)")
    meshes = tetra_parts((4, 4, 2), (0, 0, 0), (10, 10, 5), (1, 1, 1), 0.1)
    ref = refharness.RefSim([meshes[0].view], abi.SCHEME_DG_MODAL, 2, 0, 3)
    cpu = oracle.OracleSim([meshes[0].view], abi.SCHEME_DG_MODAL, 2, 3)
    for sim in (ref, cpu):
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    # smooth data: where the two face values agree to rounding the SIGN of f(p_L) decides whether FindRoot halves and
    # returns the 1e-8 iterate (tests/common.py NOISE_TOL); the residual carries that, the time-stepped solution dt x it
    assert common.col_err(cpu.get_residual(0.2)[0], ref.get_residual(0.2)[0]) < 1e-9
    ref.rk_update(0.0, 5e-3, 5)
    cpu.rk_update(0.0, 5e-3, 5)
    common.assert_parity(cpu.get_solution(0), ref.get_solution(0), 1e-10)


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def _golden_meshes(d):
    return tetra_parts(tuple(int(x) for x in d["cells"]), tuple(d["lo"]), tuple(d["hi"]), tuple(int(x) for x in d["parts"]),
                       float(d["distortion"]))


@pytest.mark.parametrize("name", CASES)
def test_oracle_tetra_golden(lib, name):
    d = _load(name)
    meshes = _golden_meshes(d)
    n = len(meshes)
    cpu = oracle.OracleSim([m.view for m in meshes], abi.SCHEME_DG_MODAL, int(d["degree"]), int(d["rk_order"]),
                           limiter=int(d["limiter"]))
    set_walls(cpu)
    cpu.set_solution([d[f"u0_{r}"] for r in range(n)])
    if int(d["limiter"]):
        assert cpu.troubled_count() == int(d["troubled_first"])
    for r in range(n):
        assert common.rel_err(cpu.get_solution(r), d[f"limited_{r}"]) < 1e-12
    res = cpu.get_residual(float(d["t0"]))
    for r in range(n):
        common.assert_parity(res[r], d[f"residual_{r}"], 1e-8)
    cpu.rk_update(float(d["t0"]), float(d["dt"]), int(d["steps"]))
    for r in range(n):
        common.assert_parity(cpu.get_solution(r), d[f"u1_{r}"], 1e-10)
    if int(d["limiter"]):
        assert cpu.troubled_count() == int(d["troubled_total"])


# ---- the CUDA path ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", [c for c in CASES if "parts" not in c])
def test_gpu_tetra_golden(lib, name):
    d = _load(name)
    meshes = _golden_meshes(d)
    gpu = abi.Solver(meshes[0].view, scheme=abi.SCHEME_DG_MODAL, degree=int(d["degree"]), limiter=int(d["limiter"]),
                     rk_order=int(d["rk_order"]))
    set_walls(gpu)
    gpu.set_solution(d["u0_0"])            # SetSolutionColumn: limits
    common.assert_parity(gpu.get_solution(), d["limited_0"], 1e-10)
    gpu.set_time(float(d["t0"]))
    common.assert_parity(gpu.get_residual(), d["residual_0"], 1e-8)
    gpu.rk_update(float(d["t0"]), float(d["dt"]), int(d["steps"]))
    common.assert_parity(gpu.get_solution(), d["u1_0"], 2e-9 if int(d["limiter"]) == 0 else 1e-10)


@pytest.mark.gpu
@pytest.mark.parametrize("degree,limiter,rk_order", [(1, 0, 2), (2, 0, 3), (3, 0, 3), (2, 1, 3), (1, 2, 2), (2, 2, 3), (3, 2, 3)])
def test_gpu_tetra_matches_oracle(lib, degree, limiter, rk_order):
    """Fresh inputs: device projection of the IC over the tetrahedral rule, one limiter pass, limited RK steps"""
    meshes = tetra_parts((5, 4, 3), (0, 0, 0), HI, (1, 1, 1), 0.12)
    gpu = abi.Solver(meshes[0].view, scheme=abi.SCHEME_DG_MODAL, degree=degree, limiter=limiter, rk_order=rk_order)
    cpu = oracle.OracleSim([meshes[0].view], abi.SCHEME_DG_MODAL, degree, rk_order, limiter=limiter)
    for sim in (gpu, cpu):
        set_walls(sim)
        sim.approximate(abi.FUNC_PLANE_SPLIT, sod_split(HI, weak=(limiter == 0)), 0.0)
    assert common.rel_err(gpu.get_solution(), cpu.get_solution(0)) < 1e-13   # Projection::Approximate
    u0 = cpu.get_solution(0)
    cpu.set_solution([u0])
    gpu.set_solution(u0)
    common.assert_parity(gpu.get_solution(), cpu.get_solution(0), 1e-10)  # one limiter pass
    gpu.rk_update(0.0, 1e-3, 4)
    cpu.rk_update(0.0, 1e-3, 4)
    assert np.isfinite(gpu.get_solution()).all()
    common.assert_parity(gpu.get_solution(), cpu.get_solution(0), 2e-9)   # see tests/test_modal.py: noise-sign mechanism
    if limiter:
        assert cpu.troubled_count() > 0
        assert gpu.troubled_cells() == cpu.troubled_count()


@pytest.mark.gpu
def test_gpu_tetra_vortex_residual_steps_and_l1(lib):
    meshes = tetra_parts((5, 5, 2), (0, 0, 0), (10, 10, 4), (1, 1, 1), 0.1)
    gpu = abi.Solver(meshes[0].view, scheme=abi.SCHEME_DG_MODAL, degree=2, rk_order=3)
    cpu = oracle.OracleSim([meshes[0].view], abi.SCHEME_DG_MODAL, 2, 3)
    for sim in (gpu, cpu):
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    gpu.set_time(0.3)
    # the residual of smooth data carries the noise-sign mechanism of the exact solver (tests/common.py NOISE_TOL: a
    # face point whose f(p_L) changes sign with the evaluation order returns the 1e-8 iterate); dt x that reaches the
    # solution, which is held to 1e-10
    common.assert_parity(gpu.get_residual(), cpu.get_residual(0.3)[0], 2e-9)
    gpu.rk_update(0.0, 5e-3, 10)
    cpu.rk_update(0.0, 5e-3, 10)
    common.assert_parity(gpu.get_solution(), cpu.get_solution(0), 1e-10)
    l1_gpu = gpu.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.05)
    l1_cpu = cpu.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.05)
    assert np.abs(l1_gpu - l1_cpu).max() <= 1e-10 * np.abs(l1_cpu).max()


@pytest.mark.gpu
def test_gpu_tetra_every_boundary_kind(lib):
    """the six boundary kinds on triangles, one per side of the box (fem.hpp:185-226 order inside the library)"""
    meshes = tetra_parts((4, 3, 3), (0, 0, 0), (10, 10, 10), (1, 1, 1), 0.1)
    gpu = abi.Solver(meshes[0].view, scheme=abi.SCHEME_DG_MODAL, degree=2, rk_order=3)
    cpu = oracle.OracleSim([meshes[0].view], abi.SCHEME_DG_MODAL, 2, 3)
    kinds = [abi.BC_SUBSONIC_INLET, abi.BC_SUBSONIC_OUTLET, abi.BC_SMART, abi.BC_SUPERSONIC_INLET,
             abi.BC_SOLID_WALL, abi.BC_SUPERSONIC_OUTLET]
    for sim in (gpu, cpu):
        for g, kind in enumerate(kinds):
            if kind in (abi.BC_SOLID_WALL, abi.BC_SUPERSONIC_OUTLET):
                sim.set_boundary(g, kind)
            else:
                sim.set_boundary(g, kind, abi.FUNC_VORTEX, common.VORTEX)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    gpu.set_time(0.1)
    common.assert_parity(gpu.get_residual(), cpu.get_residual(0.1)[0], 1e-10)


@pytest.mark.gpu
def test_gpu_tetra_two_parts_with_hand_fed_ghosts(lib):
    """a 2-part tetrahedral mesh on one GPU: each part's ghost tail is filled by hand from the other part's column
    (what UpdateGhostCellCoeffs unpacks), residuals against the oracle's two-part residual"""
    meshes = tetra_parts((4, 3, 2), (0, 0, 0), HI, (2, 1, 1), 0.1)
    views = [m.view for m in meshes]
    cpu = oracle.OracleSim(views, abi.SCHEME_DG_MODAL, 2, 3)
    gpus = [abi.Solver(v, scheme=abi.SCHEME_DG_MODAL, degree=2, rk_order=3) for v in views]
    for sim in [cpu] + gpus:
        common.set_vortex_case(sim)
    cpu.approximate(abi.FUNC_VORTEX, [5.0, 0.6, 0.5, 1.0, 1.0, 1.4], 0.0)
    cols = [cpu.get_solution(r) for r in range(2)]
    fields = 50
    for r, g in enumerate(gpus):
        g.set_solution(cols[r])
        a = meshes[1 - r].arrays()
        ghosts = cols[1 - r].reshape(-1, fields)[a["send_cells"]]   # what the neighbour sends, in its send order
        g.set_ghost_solution(np.ascontiguousarray(ghosts.reshape(-1)))
    res = cpu.get_residual(0.0)
    for r, g in enumerate(gpus):
        g.set_time(0.0)
        common.assert_parity(g.get_residual(), res[r], 1e-10)


@pytest.mark.gpu
def test_tetrahedra_need_the_modal_scheme(lib):
    m = abi.BoxMesh((2, 2, 2), tetra=True)
    with pytest.raises(RuntimeError, match="modal"):
        abi.Solver(m.view, scheme=abi.SCHEME_FR, degree=2)
    gpu = abi.Solver(m.view, scheme=abi.SCHEME_DG_MODAL, degree=2)
    with pytest.raises(RuntimeError):
        gpu.vtk_sample()
