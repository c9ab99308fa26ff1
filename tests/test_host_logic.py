"""CPU tests of the host side: the C-ABI library loads and exports every symbol of
include/minicfd_b200.h, the box-mesh generator keeps the reference's per-part ordering contract
(mesh/part.hpp:818-835, 966-986, 1019-1078), and the product fails loudly without a GPU.
No compute entry point is exercised here (there is no GPU in this container).
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

from minicfd_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "minicfd_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(mcfd_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 25
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/minicfd_b200.h but not exported"
    assert declared == set(abi.SIGNATURES), declared ^ set(abi.SIGNATURES)


def test_struct_layouts_match_the_header(lib):
    assert C.sizeof(abi.Mesh) == lib.mcfd_struct_size(0)
    assert C.sizeof(abi.Config) == lib.mcfd_struct_size(1)
    assert C.sizeof(abi.Timing) == lib.mcfd_struct_size(2)


def test_no_gpu_means_an_error_not_a_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU-only container")
    mesh = abi.BoxMesh((2, 2, 2))
    with pytest.raises(RuntimeError, match="no CUDA device"):
        abi.Solver(mesh.view)
    z = np.zeros((1, 5)); fr = np.eye(3).reshape(1, 9).copy()
    with pytest.raises(RuntimeError):
        abi.riemann_flux(0, 1.4, fr, z + 1, z + 1)


@pytest.mark.parametrize("n,parts", [((4, 3, 2), (1, 1, 1)), ((5, 4, 3), (2, 1, 1)), ((4, 4, 4), (2, 2, 2)),
                                     ((7, 5, 3), (3, 2, 1)), ((2, 2, 2), (2, 2, 2))])
def test_box_mesh_part_ordering_contract(lib, n, parts):
    world = parts[0] * parts[1] * parts[2]
    meshes = [abi.BoxMesh(n, (0, 0, 0), (1, 2, 3), parts=parts, rank=r, distortion=0.1) for r in range(world)]
    owners = {}
    total_faces = 0
    for r, m in enumerate(meshes):
        v, a, gid = m.view, m.arrays(), m.cell_global_ids()
        nl, ng = v.n_local_cells, v.n_ghost_cells
        # every global cell is local on exactly one part
        for g in gid[:nl]:
            assert g not in owners
            owners[int(g)] = r
        # inner cells first, each block ascending global id (part.hpp:818-835)
        assert (np.diff(gid[:v.n_inner_cells]) > 0).all() and (np.diff(gid[v.n_inner_cells:nl]) > 0).all()
        nlf, ngf, nbf = v.n_local_faces, v.n_ghost_faces, v.n_boundary_faces
        h, s = a["face_holder"], a["face_sharer"]
        assert ((h >= 0) & (h < nl)).all()
        assert ((s[:nlf] >= 0) & (s[:nlf] < nl)).all()                   # local faces: both cells local
        assert ((s[nlf:nlf + ngf] >= nl) & (s[nlf:nlf + ngf] < nl + ng)).all()  # ghost faces: held by the local cell
        assert (s[nlf + ngf:] == -1).all()
        # inner cells touch no ghost face (part.hpp:822)
        assert (h[nlf:nlf + ngf] >= v.n_inner_cells).all()
        touched = np.zeros(nl, bool); touched[h[nlf:nlf + ngf]] = True
        assert touched[v.n_inner_cells:].all()
        # face normals point holder -> sharer / outward (part.hpp:1521-1522)
        xyz, cn, fn = a["node_xyz"], a["cell_nodes"], a["face_nodes"]
        cc = xyz[cn].mean(axis=1)
        fc = xyz[fn].mean(axis=1)
        # holder's sorted face (geometry/hexahedron.hpp:371-380): outward orientation
        table = np.array([[0, 3, 2, 1], [0, 1, 5, 4], [1, 2, 6, 5], [2, 3, 7, 6], [0, 4, 7, 3], [4, 5, 6, 7]])
        for f in range(nlf + ngf + nbf):
            hn = cn[h[f]]
            side = [i for i in range(6) if set(hn[table[i]]) == set(fn[f])]
            assert len(side) == 1
            q = xyz[hn[table[side[0]]]]
            normal = np.cross(q[1] - q[0], q[3] - q[0])
            assert normal @ (fc[f] - cc[h[f]]) > 0
            if s[f] >= 0:
                assert normal @ (cc[s[f]] - cc[h[f]]) > 0
        total_faces += nlf + nbf
        total_faces += ngf / 2.0
        # halo lists: ghosts grouped by neighbour part, ascending global id (part.hpp:966-986)
        for i, peer in enumerate(a["nbr_rank"]):
            lo, hi = a["recv_ptr"][i], a["recv_ptr"][i + 1]
            g = gid[nl + lo:nl + hi]
            assert (np.diff(g) > 0).all()
            pa, pg = meshes[peer].arrays(), meshes[peer].cell_global_ids()
            j = list(pa["nbr_rank"]).index(r)
            sent = pg[pa["send_cells"][pa["send_ptr"][j]:pa["send_ptr"][j + 1]]]
            assert np.array_equal(sent, g)           # what the peer packs is what I unpack, in order
    assert len(owners) == n[0] * n[1] * n[2]
    nx, ny, nz = n
    assert total_faces == (nx + 1) * ny * nz + nx * (ny + 1) * nz + nx * ny * (nz + 1)


def test_distortion_is_partition_independent(lib):
    whole = abi.BoxMesh((4, 4, 4), distortion=0.2)
    part = abi.BoxMesh((4, 4, 4), parts=(2, 2, 1), rank=3, distortion=0.2)
    wa, pa = whole.arrays(), part.arrays()
    wc = {int(g): wa["node_xyz"][wa["cell_nodes"][i]] for i, g in enumerate(whole.cell_global_ids())}
    for i, g in enumerate(part.cell_global_ids()):
        assert np.array_equal(pa["node_xyz"][pa["cell_nodes"][i]], wc[int(g)])
