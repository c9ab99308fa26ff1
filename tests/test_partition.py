"""Partition + ghost-index pipeline (minicfd_b200/csrc/partition.cpp).

Pinned pieces: mesh::GetNewOrder / ShuffleConnectivity and metis::GetNodeParts against (a) the golden vectors of the
reference's own tests (test/mesh/shuffler.cpp:29-65, test/mesh/metis.cpp:169-181) and (b) the reference templates
compiled as oracle/_ref/libref_partition.so, on random inputs -- bit-exact (integer work).
Unpinned (METIS is an un-vendored submodule): the neighbour ORDER of the dual graph; its neighbour SETS are checked.
Part's build order is checked through its consequences: identical cell / ghost / send / recv indexing to the box
generator for block partitions, and identical solutions for an arbitrary partition against the unpartitioned mesh.
"""
import ctypes as C
import os

import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle
from tests import common

REF_LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libref_partition.so")
i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))


def new_order(lib, parts):
    parts = i32(parts)
    a, b = np.zeros(len(parts), np.int32), np.zeros(len(parts), np.int32)
    assert lib.mcfd_get_new_order(ip(parts), len(parts), ip(a), ip(b)) == abi.OK
    return a, b


def shuffle_connectivity(lib, old_to_new_nodes, new_to_old_cells, npe, conn):
    o, c, conn = i32(old_to_new_nodes), i32(new_to_old_cells), i32(conn).copy()
    assert lib.mcfd_shuffle_connectivity(ip(o), len(o), ip(c), len(c), npe, ip(conn)) == abi.OK
    return conn


def node_parts(lib, rng, idx, cell_parts, n_nodes, n_parts):
    rng, idx, cp = i32(rng), i32(idx), i32(cell_parts)
    out = np.zeros(n_nodes, np.int32)
    assert lib.mcfd_get_node_parts(len(cp), n_nodes, ip(rng), ip(idx), ip(cp), n_parts, ip(out)) == abi.OK
    return out


# ---- the reference's own known-answer tests -------------------------------------------------------------------------
def test_get_new_order_golden(lib):   # test/mesh/shuffler.cpp:29-47
    new_to_old, old_to_new = new_order(lib, [1, 0, 0, 2, 0, 1, 2, 1, 0, 2])
    assert new_to_old.tolist() == [1, 2, 4, 8, 0, 5, 7, 3, 6, 9]
    assert (old_to_new[new_to_old] == np.arange(10)).all()
    data = np.array([1.0, 0.0, 0.1, 2.0, 0.2, 1.1, 2.1, 1.2, 0.3, 2.2])
    assert data[new_to_old].tolist() == [0.0, 0.1, 0.2, 0.3, 1.0, 1.1, 1.2, 2.0, 2.1, 2.2]   # ShuffleData


def test_shuffle_connectivity_golden(lib):   # test/mesh/shuffler.cpp:48-65 (the argument is used as old_to_new)
    conn = shuffle_connectivity(lib, [5, 4, 3, 2, 1, 0], [1, 2, 3, 0], 3, [1, 2, 6, 2, 3, 6, 6, 3, 5, 3, 4, 5])
    assert conn.tolist() == [5, 4, 1, 1, 4, 2, 4, 3, 2, 6, 5, 1]


def test_get_node_parts_golden(lib):   # test/mesh/metis.cpp:169-181
    out = node_parts(lib, [0, 4, 8, 11, 14], [0, 2, 3, 1, 2, 4, 5, 3, 6, 8, 7, 8, 9, 7], [2, 0, 0, 1], 10, 3)
    assert out.tolist() == [2, 2, 0, 0, 0, 0, 0, 0, 0, 1]


# ---- against the compiled reference templates -----------------------------------------------------------------------
@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(REF_LIB):
        pytest.skip("oracle/_ref/libref_partition.so not built")
    return C.CDLL(REF_LIB)


@pytest.mark.parametrize("seed", range(5))
def test_helpers_match_compiled_reference(lib, ref, seed):
    rng = np.random.default_rng(20241104 + seed)
    n_cells, n_nodes, n_parts, npe = int(rng.integers(5, 400)), int(rng.integers(8, 300)), int(rng.integers(1, 9)), 8
    parts = i32(rng.integers(0, n_parts, n_cells))
    a, b = new_order(lib, parts)
    ra, rb = np.zeros(n_cells, np.int32), np.zeros(n_cells, np.int32)
    ref.ref_get_new_order(ip(parts), n_cells, ip(ra), ip(rb))
    assert (a == ra).all() and (b == rb).all()
    conn = i32(rng.integers(0, n_nodes, n_cells * npe))
    rng_csr = i32(np.arange(n_cells + 1) * npe)
    mine = node_parts(lib, rng_csr, conn, parts, n_nodes, n_parts)
    theirs = np.zeros(n_nodes, np.int32)
    ref.ref_get_node_parts(n_cells, n_nodes, ip(rng_csr), ip(conn), ip(parts), n_parts, ip(theirs))
    assert (mine == theirs).all()
    _, node_old_to_new = new_order(lib, mine)
    conn1 = i32(conn + 1)
    shuffled = shuffle_connectivity(lib, node_old_to_new, a, npe, conn1)
    expected = conn1.copy()
    ref.ref_shuffle_connectivity(ip(node_old_to_new), n_nodes, ip(a), n_cells, npe, ip(expected))
    assert (shuffled == expected).all()


# ---- Part's build order: block partitions against the box generator -------------------------------------------------
def global_box(n, lo, hi):
    """The unpartitioned box as a general mesh: cell id = i + nx (j + ny k), HEXA_8 nodes in CGNS order, boundary
    faces grouped x-min, x-max, y-min, y-max, z-min, z-max like the box generator."""
    nx, ny, nz = n
    mx, my = nx + 1, ny + 1
    gx, gy, gz = [np.linspace(lo[d], hi[d], n[d] + 1) for d in range(3)]
    kk, jj, ii = np.meshgrid(np.arange(nz + 1), np.arange(ny + 1), np.arange(nx + 1), indexing="ij")
    xyz = np.stack([gx[ii].ravel(), gy[jj].ravel(), gz[kk].ravel()], axis=1)
    nid = lambda i, j, k: i + mx * (j + my * k)
    corner = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)]
    side_corners = [(0, 4, 7, 3), (1, 2, 6, 5), (0, 1, 5, 4), (2, 3, 7, 6), (0, 3, 2, 1), (4, 5, 6, 7)]
    cells, bnodes, bgroup = [], [], []
    for k in range(nz):
        for j in range(ny):
            for i in range(nx):
                c = [nid(i + a, j + b, k + d) for a, b, d in corner]
                cells.append(c)
                on = [i == 0, i == nx - 1, j == 0, j == ny - 1, k == 0, k == nz - 1]
                for s in range(6):
                    if on[s]:
                        bnodes.append([c[v] for v in side_corners[s]])
                        bgroup.append(s)
    return xyz, np.array(cells), np.array(bnodes), np.array(bgroup)


def block_parts(n, parts):
    def owner_1d(count, blocks):
        q, r = divmod(count, blocks)
        out = np.zeros(count, np.int32)
        for b in range(blocks):
            lo = b * q + min(b, r)
            out[lo:lo + q + (1 if b < r else 0)] = b
        return out
    ox, oy, oz = [owner_1d(n[d], parts[d]) for d in range(3)]
    k, j, i = np.meshgrid(np.arange(n[2]), np.arange(n[1]), np.arange(n[0]), indexing="ij")
    return (ox[i] + parts[0] * (oy[j] + parts[1] * oz[k])).ravel().astype(np.int32)


def face_keys(mesh, first, count):
    a, gid = mesh.arrays(), mesh.cell_global_ids()
    h, s = a["face_holder"][first:first + count], a["face_sharer"][first:first + count]
    return sorted((int(gid[x]), int(gid[y]) if y >= 0 else -1) for x, y in zip(h, s))


@pytest.mark.parametrize("n,parts", [((6, 5, 4), (2, 1, 1)), ((6, 5, 4), (2, 2, 1)), ((5, 4, 4), (2, 2, 2)), ((7, 3, 3), (3, 1, 1))])
def test_block_partition_matches_box_generator(lib, n, parts):
    lo, hi = (0, 0, 0), (3.0, 2.0, 1.0)
    xyz, cells, bnodes, bgroup = global_box(n, lo, hi)
    cell_parts = block_parts(n, parts)
    world = parts[0] * parts[1] * parts[2]
    pt = abi.Partitioner(xyz, cells, cell_parts, world, bnodes, bgroup, 6)
    # dual graph: the face neighbours, symmetric
    rng, idx = pt.dual_graph()
    nx, ny, nz = n
    for c in (0, len(cells) // 2, len(cells) - 1):
        i, j, k = c % nx, (c // nx) % ny, c // (nx * ny)
        want = {a + nx * (b + ny * d) for a, b, d in [(i - 1, j, k), (i + 1, j, k), (i, j - 1, k), (i, j + 1, k), (i, j, k - 1), (i, j, k + 1)]
                if 0 <= a < nx and 0 <= b < ny and 0 <= d < nz}
        assert set(idx[rng[c]:rng[c + 1]].tolist()) == want
    assert (pt.face_parts() == cell_parts[[np.flatnonzero((cells == f[0]).any(1) & (cells == f[1]).any(1) & (cells == f[2]).any(1) & (cells == f[3]).any(1))[0] for f in bnodes]]).all()
    for r in range(world):
        mine, box = pt.part_mesh(r), abi.BoxMesh(n, lo, hi, parts=parts, rank=r)
        a, b = mine.view, box.view
        for field in ("n_local_cells", "n_inner_cells", "n_ghost_cells", "n_local_faces", "n_ghost_faces", "n_boundary_faces", "n_neighbours", "n_nodes"):
            assert getattr(a, field) == getattr(b, field), field
        assert (mine.cell_global_ids() == box.cell_global_ids()).all()          # local inner | inter, then ghosts
        ma, ba = mine.arrays(), box.arrays()
        for key in ("nbr_rank", "send_ptr", "recv_ptr", "send_cells"):          # halo lists, bit-exact
            assert (ma[key] == ba[key]).all(), key
        nl, ng, nb = a.n_local_faces, a.n_ghost_faces, a.n_boundary_faces
        assert face_keys(mine, 0, nl) == face_keys(box, 0, nl)
        assert face_keys(mine, nl, ng) == face_keys(box, nl, ng)
        assert face_keys(mine, nl + ng, nb) == face_keys(box, nl + ng, nb)
        assert sorted(ma["bface_group"].tolist()) == sorted(ba["bface_group"].tolist())
        # the partition text: one line per local face in the inner list, one per ghost face in the interpart list
        sections = pt.info_text(r).split("#")[1:]
        body = [s.strip().split("\n")[1:] for s in sections]
        assert len(body[4]) == nl and len(body[5]) == ng
        head, tail = [int(x) for x in body[3][0].split()[2:]]
        assert tail - head == a.n_local_cells
        nh, nt = [int(x) for x in body[0][0].split()[1:]]
        assert nt - nh == int((pt.node_parts() == r).sum())


def test_arbitrary_partition_reproduces_the_unpartitioned_solution(lib):
    """3 ragged parts (not blocks): the halo indexing is right iff the partitioned run equals the single-part run."""
    n, lo, hi = (5, 4, 3), (0, 0, 0), (10.0, 8.0, 6.0)
    xyz, cells, bnodes, bgroup = global_box(n, lo, hi)
    many, parts, _n = ragged_three_parts(n, lo, hi)
    one = abi.Partitioner(xyz, cells, np.zeros(len(cells), np.int32), 1, bnodes, bgroup, 6)
    whole = one.part_mesh(0)
    sims = [oracle.OracleSim([m.view for m in parts], abi.SCHEME_FR, 2, 3), oracle.OracleSim([whole.view], abi.SCHEME_FR, 2, 3)]
    for sim in sims:
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
        sim.rk_update(0.0, 5e-3, 3)
    full = sims[1].get_solution(0).reshape(len(cells), -1)     # single part: local cell index == metis id
    assert (whole.cell_global_ids() == np.arange(len(cells))).all()
    for r, m in enumerate(parts):
        gid = m.cell_global_ids()[:m.view.n_local_cells]
        got = sims[0].get_solution(r).reshape(len(gid), -1)
        # not 1e-10: a ghost face is solved once per side with left and right swapped, and the reference's Newton loop
        # stops at |f| < 1e-8 (exact.hpp:163-174), so the two sides' fluxes -- and a partitioned run against an
        # unpartitioned one -- agree to ~1e-10, in the reference as well.  A wrong ghost index shows up at 1e-3.
        assert common.rel_err(got, full[gid]) < 1e-8


def ragged_three_parts(n=(5, 4, 3), lo=(0, 0, 0), hi=(10.0, 8.0, 6.0)):
    """3 ragged (non-block) parts of a box through the partition pipeline: (Partitioner, [PartMesh x 3], n_cells)"""
    xyz, cells, bnodes, bgroup = global_box(n, lo, hi)
    rng = np.random.default_rng(7)
    cx = xyz[cells].mean(axis=1)
    cell_parts = ((cx[:, 0] + 0.8 * cx[:, 1] + rng.uniform(-1.5, 1.5, len(cells))) // 6.5).clip(0, 2).astype(np.int32)
    assert len(set(cell_parts.tolist())) == 3
    pt = abi.Partitioner(xyz, cells, cell_parts, 3, bnodes, bgroup, 6)
    return pt, [pt.part_mesh(r) for r in range(3)], len(cells)


@pytest.mark.gpu
@pytest.mark.parametrize("scheme,degree", [(abi.SCHEME_FR, 2), (abi.SCHEME_DGSEM, 3), (abi.SCHEME_DG_MODAL, 2)])
def test_gpu_ragged_partition_residual_matches_oracle(lib, scheme, degree):
    """Row f1 through the CUDA path: the parts partition.cpp builds from a ragged 3-part cell -> part vector (ghost
    cells by neighbour then metis id, send / recv lists as mesh/part.hpp:964-986 orders them) go to one mcfd_solver
    each, ghosts fed by hand from the peers' columns through exactly those lists, against the oracle's 3-part residual
    on the same parts (same holder / sharer roles on both sides, so the 1e-10 bar applies)."""
    _pt, parts, _n = ragged_three_parts()
    views = [m.view for m in parts]
    gpu = [abi.Solver(v, scheme=scheme, degree=degree, rk_order=3, rank=r, world=3) for r, v in enumerate(views)]
    cpu = oracle.OracleSim(views, scheme, degree, 3)
    for sim in gpu + [cpu]:
        common.set_vortex_case(sim)
    cpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    cols = [common.perturbed(cpu.get_solution(r), seed=r + 1) for r in range(3)]
    cpu.set_solution(cols)
    want = cpu.get_residual(0.25)
    for r, g in enumerate(gpu):
        g.set_solution(cols[r])
    for r, m in enumerate(parts):   # Part::ShareGhostCellCoeffs / UpdateGhostCellCoeffs by hand (part.hpp:1258-1302)
        a = m.arrays()
        fields = gpu[r].fields_per_cell
        chunks = []
        for peer in a["nbr_rank"]:
            pa = parts[peer].arrays()
            j = list(pa["nbr_rank"]).index(r)
            chunks.append(cols[peer].reshape(-1, fields)[pa["send_cells"][pa["send_ptr"][j]:pa["send_ptr"][j + 1]]].ravel())
        assert a["recv_ptr"][-1] == sum(len(c) for c in chunks) // fields
        gpu[r].set_ghost_solution(np.concatenate(chunks))
    for r, g in enumerate(gpu):
        g.set_time(0.25)
        common.assert_parity(g.get_residual(), want[r], 1e-10, what=f"part {r}")


@pytest.mark.gpu
def test_gpu_runs_on_pipeline_built_part(lib):
    n, lo, hi = (6, 5, 4), (0, 0, 0), (10.0, 10.0, 10.0)
    xyz, cells, bnodes, bgroup = global_box(n, lo, hi)
    whole = abi.Partitioner(xyz, cells, np.zeros(len(cells), np.int32), 1, bnodes, bgroup, 6).part_mesh(0)
    box = abi.BoxMesh(n, lo, hi)
    cols = []
    for mesh in (whole, box):
        gpu = abi.Solver(mesh.view, scheme=abi.SCHEME_FR, degree=2, rk_order=3)
        common.set_vortex_case(gpu)
        gpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
        gpu.rk_update(0.0, 5e-3, 3)
        cols.append(gpu.get_solution())
    assert common.rel_err(cols[0], cols[1]) < 1e-12
