"""Row f2: what Main() does with the column between steps -- Section::GatherFields /
ScatterFields (mesh/part.hpp:370-387) and mesh::vtk::Writer<Part>::WriteSolutions (mesh/vtk.hpp:385-487).

Fixtures tests/golden/vtu_*.npz hold, for a given column, the files the reference's own unmodified writer
produced, its full-precision samples (LocalToGlobalAndValue at the 27 nodes of vtk::Hexahedron27) and its
field-major arrays (tests/golden/make_golden_vtu.py).
CPU: the host writer reproduces the reference's files byte for byte from the fixture's samples.
GPU: the device sampling matches the reference's to 1e-10 (points bit for bit), the files written through
mcfd_write_vtu equal the reference's (same bytes outside the 6-digit value tokens, and those within a unit
of the last printed digit), gather/scatter are exact.
"""
import base64
import glob
import os
import re

import numpy as np
import pytest

from minicfd_b200 import abi
from tests import common

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(HERE, "golden", "vtu_*.npz")))
NAMES = ["Density", "MomentumX", "MomentumY", "MomentumZ", "EnergyStagnationDensity"]


def _load(name):
    return np.load(os.path.join(HERE, "golden", name + ".npz"))


def _meshes(d):
    return common.make_parts(tuple(int(x) for x in d["cells"]), tuple(d["lo"]), tuple(d["hi"]),
                             tuple(int(x) for x in d["parts"]), float(d["distortion"]))


def _split_vtu(text):
    """(skeleton with the five value arrays blanked, [values per field])"""
    fields = []

    def take(m):
        fields.append(np.array(m.group(2).split(), dtype=np.float64))
        return m.group(1) + "@\n" + m.group(3)
    pattern = r'(<DataArray type="Float64" Name="(?:%s)" format="ascii">\n)([^<]*)\n(        </DataArray>)' % "|".join(NAMES)
    return re.sub(pattern, take, text), fields


def _points_of(text):
    block = re.search(r'Name="Points" NumberOfComponents="3" format="binary">\n([^\n]*)\n', text).group(1)
    n_bytes = int(np.frombuffer(base64.b64decode(block[:12]), dtype=np.uint64)[0])
    return np.frombuffer(base64.b64decode(block[12:]), dtype=np.float64).reshape(-1, 3), n_bytes


def test_fixtures_exist():
    assert len(CASES) >= 7


@pytest.mark.parametrize("name", CASES)
def test_reference_fields_are_the_transposed_column(name):
    """GatherFields stores coeff.reshaped()[i_field] of cell i_cell at fields_[i_field][i_cell]; the column is the
    cells' coefficient blocks one after the other (fem.hpp:153-172) -- so fields = column.reshape(cells, K N).T"""
    d = _load(name)
    for r in range(int(np.prod(d["parts"]))):
        f = d[f"fields_{r}"]
        assert np.array_equal(f, d[f"u_{r}"].reshape(-1, f.shape[0]).T)


@pytest.mark.parametrize("name", CASES)
def test_host_writer_reproduces_the_reference_files(lib, tmp_path, name):
    """mcfd_write_vtu_points (no device): same bytes as the reference's writer, given the reference's samples."""
    d = _load(name)
    world = int(np.prod(d["parts"]))
    for r in reversed(range(world)):
        abi.write_vtu_points(str(tmp_path), "Frame0", r, world, d[f"coords_{r}"], d[f"values_{r}"], lib=lib)
    assert (tmp_path / "Frame0.pvtu").read_bytes() == d["pvtu"].tobytes()
    for r in range(world):
        assert (tmp_path / "Frame0" / f"{r}.vtu").read_bytes() == d[f"vtu_{r}"].tobytes()


def test_host_writer_errors(lib, tmp_path):
    blocker = tmp_path / "file"
    blocker.write_text("x")
    with pytest.raises(RuntimeError):
        abi.write_vtu_points(str(blocker), "Frame0", 0, 1, np.zeros((27, 3)), np.zeros((27, 5)), lib=lib)
    with pytest.raises(RuntimeError):
        abi.write_vtu_points(str(tmp_path), "Frame0", 2, 2, np.zeros((27, 3)), np.zeros((27, 5)), lib=lib)
    abi.write_vtu_points(str(tmp_path), "Empty", 0, 1, np.zeros((0, 3)), np.zeros((0, 5)), lib=lib)  # an empty part
    assert 'NumberOfPoints="0" NumberOfCells="0"' in (tmp_path / "Empty" / "0.vtu").read_text()


def test_fixture_files_are_consistent():
    """the reference's printed values are its samples at 6 significant digits; its point block is its points"""
    for name in CASES:
        d = _load(name)
        text = d["vtu_0"].tobytes().decode()
        _, fields = _split_vtu(text)
        values = d["values_0"]
        for k in range(5):
            assert np.allclose(fields[k], values[:, k], rtol=1e-5, atol=1e-300)
        pts, n_bytes = _points_of(text)
        assert n_bytes == pts.size * 8 and np.array_equal(pts, d["coords_0"])


# ---------------------------------------------------------------- GPU ------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_sampling_and_files_match_the_reference(lib, tmp_path, name):
    d = _load(name)
    meshes = _meshes(d)
    world = len(meshes)
    for r in reversed(range(world)):
        sim = abi.Solver(meshes[r].view, scheme=int(d["scheme"]), degree=int(d["degree"]), rk_order=3, rank=r, world=world)
        sim.set_solution(d[f"u_{r}"])
        coords, values = sim.vtk_sample()
        assert np.array_equal(coords, d[f"coords_{r}"])              # LocalToGlobal, same operation order
        common.assert_parity(values, d[f"values_{r}"], 1e-10)
        assert np.array_equal(sim.get_solution(), d[f"u_{r}"])       # sampling used scratch columns only
        sim.write_vtu(str(tmp_path), "Frame0")
        got = (tmp_path / "Frame0" / f"{r}.vtu").read_text()
        want = d[f"vtu_{r}"].tobytes().decode()
        got_skel, got_fields = _split_vtu(got)
        want_skel, want_fields = _split_vtu(want)
        assert got_skel == want_skel                                   # XML, point block, connectivity, offsets, types
        for a, b in zip(got_fields, want_fields):
            assert a.shape == b.shape and np.allclose(a, b, rtol=2e-6, atol=1e-300)
    assert (tmp_path / "Frame0.pvtu").read_bytes() == d["pvtu"].tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["vtu_fr_p2", "vtu_dg_p2", "vtu_fr_p3"])
def test_gpu_gather_scatter_fields(lib, name):
    d = _load(name)
    meshes = _meshes(d)  # keeps the arrays behind .view alive
    sim = abi.Solver(meshes[0].view, scheme=int(d["scheme"]), degree=int(d["degree"]), rk_order=3)
    sim.set_solution(d["u_0"])
    fields = sim.gather_fields()
    assert np.array_equal(fields, d["fields_0"])
    sim.set_solution(np.zeros_like(d["u_0"]))
    sim.scatter_fields(fields)                                         # the restart direction
    assert np.array_equal(sim.get_solution(), d["u_0"])


@pytest.mark.gpu
def test_gpu_gather_fields_ragged_sizes(lib):
    """cell and field counts that are not multiples of the 32 x 32 transpose tile"""
    mesh = common.make_parts((7, 5, 3), (0, 0, 0), (10, 9, 8), (1, 1, 1), 0.05)[0]
    sim = abi.Solver(mesh.view, scheme=abi.SCHEME_DG_MODAL, degree=3, rk_order=1)
    u = np.random.default_rng(5).uniform(-1, 1, sim.n)
    sim.set_solution(u)
    f = sim.gather_fields()
    assert np.array_equal(f, u.reshape(-1, sim.fields_per_cell).T)
    sim.scatter_fields(f)
    assert np.array_equal(sim.get_solution(), u)
