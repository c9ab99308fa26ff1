"""Golden vectors: (1) tests/golden/*.npz, produced by the reference's own templates with
tests/golden/make_golden.py; (2) the known-answer cases of the reference's own unit tests
(test/riemann/euler/exact.cpp:54-127, Toro's tabulated star states, tolerance 1e-4 as there).
CPU: the oracle against both.  GPU (marked): the CUDA path, through the C-ABI, against both.
"""
import glob
import os

import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle
from tests import common

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(HERE, "golden", "*.npz"))
               if not os.path.basename(p).startswith(("riemann", "dg_", "dgtet_", "vtu_")))  # dg_*: tests/test_modal.py, dgtet_*: tests/test_tetra.py, vtu_*: tests/test_output.py
G = common.GAMMA


def _load(name):
    return np.load(os.path.join(HERE, "golden", name + ".npz"))


def _meshes(d):
    return common.make_parts(tuple(int(x) for x in d["cells"]), tuple(d["lo"]), tuple(d["hi"]),
                             tuple(int(x) for x in d["parts"]), float(d["distortion"]))


# ---- reference known-answer tests: exact.cpp TestSod .. TestRightVacuum --------------------------
def _cons(rho, u, p):
    return [rho, rho * u, 0.0, 0.0, p / (G - 1) + 0.5 * rho * u * u]


def _flux_of(rho, u, p):  # Exact::GetFlux of a primitive state (exact.cpp:36-41)
    return np.array([rho * u, rho * u * u + p, 0, 0, u * (p * G / (G - 1) + 0.5 * rho * u * u)])


TORO = [  # left (rho,u,p), right, expected state on the t-axis
    ((1.0, 0.0, 1.0), (0.125, 0.0, 0.1), (0.426319, +0.927453, 0.303130)),       # TestSod
    ((0.125, 0.0, 0.1), (1.0, 0.0, 1.0), (0.426319, -0.927453, 0.303130)),       # TestSod, mirrored
    ((5.99924, 19.5975, 460.894), (5.99242, 6.19633, 46.0950), (5.99924, 19.5975, 460.894)),  # TestShockCollision
    ((1.0, 0.0, 1000.0), (1.0, 0.0, 0.01), (0.575062, 19.59745, 460.8938)),      # TestBlastFromLeft
    ((1.0, 0.0, 0.01), (1.0, 0.0, 100.0), (0.575113, -6.196328, 46.09504)),      # TestBlastFromRight
    ((1.0, -2.0, 0.4), (1.0, +2.0, 0.4), (0.21852, 0.0, 0.001894)),              # TestAlmostVacuumed
    ((1.0, -4.0, 0.4), (1.0, +4.0, 0.4), (0.0, 0.0, 0.0)),                       # TestVacuumed
]


def _toro_arrays():
    ul = np.array([_cons(*l) for l, _, _ in TORO]); ur = np.array([_cons(*r) for _, r, _ in TORO])
    fr = np.tile(np.eye(3).reshape(1, 9), (len(TORO), 1))
    expect = np.array([_flux_of(*e) for _, _, e in TORO])
    return np.ascontiguousarray(fr), np.ascontiguousarray(ul), np.ascontiguousarray(ur), expect


def _check_toro(flux, expect):
    for k in (0, 1, 4):  # mass, momentumX, energy (exact.cpp:27-32)
        for a, b in zip(flux[:, k], expect[:, k]):
            if b == 0:
                assert abs(a) < 1e-3  # the tabulated states are rounded to 6 digits
            else:
                assert abs(a / b - 1) < 1e-4
    assert np.array_equal(flux[:, 2:4], np.zeros((len(TORO), 2)))


def _right_vacuum_cases():  # exact.cpp:87-127 TestRightVacuum
    out = []
    for u_left, expect in ((-6.0, (0, 0, 0)), (-5.0, (0, 0, 0)), (-4.0, (1.4 * 6.0 ** -5, 6.0 ** -1, 6.0 ** -7)),
                           (1.0, (1.4, 1.0, 1.0)), (2.0, (1.4, 2.0, 1.0))):
        out.append((_cons(1.4, u_left, 1.0), [0.0, 0.0, 0.0, 0.0, 0.0], _flux_of(*expect)))
    return out


def test_oracle_reference_known_answers(oracle_lib):
    fr, ul, ur, expect = _toro_arrays()
    _check_toro(oracle.riemann_flux(G, fr, ul, ur), expect)
    for l, r, e in _right_vacuum_cases():
        f = oracle.riemann_flux(G, np.eye(3).reshape(1, 9).copy(), np.array([l]), np.array([r]))[0]
        assert np.allclose(f, e, rtol=0, atol=1e-14)


def test_oracle_riemann_golden(oracle_lib):
    d = _load("riemann")
    f = oracle.riemann_flux(G, d["frames"], d["u_left"], d["u_right"])
    scale = np.maximum(np.abs(d["flux"]).max(axis=1, keepdims=True), 1e-30)
    assert (np.abs(f - d["flux"]) / scale).max() < 1e-12
    for kind in range(6):
        b = oracle.boundary_flux(kind, G, d["frames"], d["u_left"], d["u_right"])
        assert common.rel_err(b, d[f"bflux{kind}"]) < 1e-13


@pytest.mark.parametrize("name", CASES)
def test_oracle_golden_case(lib, name):
    d = _load(name)
    meshes = _meshes(d)
    cpu = oracle.OracleSim([m.view for m in meshes], int(d["scheme"]), int(d["degree"]), int(d["rk_order"]))
    common.set_vortex_case(cpu)
    cpu.set_solution([d[f"u0_{r}"] for r in range(len(meshes))])
    res = cpu.get_residual(float(d["t0"]))
    for r in range(len(meshes)):
        assert common.rel_err(res[r], d[f"residual_{r}"]) < 1e-12
    cpu.rk_update(float(d["t0"]), float(d["dt"]), int(d["steps"]))
    for r in range(len(meshes)):
        assert common.rel_err(cpu.get_solution(r), d[f"u1_{r}"]) < 1e-12


# ---- the CUDA path against the same vectors -----------------------------------------------------
@pytest.mark.gpu
def test_gpu_reference_known_answers(lib):
    fr, ul, ur, expect = _toro_arrays()
    _check_toro(abi.riemann_flux(abi.RIEMANN_EXACT, G, fr, ul, ur), expect)
    for l, r, e in _right_vacuum_cases():
        f = abi.riemann_flux(abi.RIEMANN_EXACT, G, np.eye(3).reshape(1, 9).copy(), np.array([l]), np.array([r]))[0]
        assert np.allclose(f, e, rtol=0, atol=1e-14)


@pytest.mark.gpu
def test_gpu_riemann_golden(lib):
    d = _load("riemann")
    f = abi.riemann_flux(abi.RIEMANN_EXACT, G, d["frames"], d["u_left"], d["u_right"])
    scale = np.maximum(np.abs(d["flux"]).max(axis=1, keepdims=True), 1e-30)
    err = np.abs(f - d["flux"]) / scale
    assert np.median(err) < 1e-14 and err.max() < 1e-9   # Newton stops at |f| < 1e-8 (exact.hpp:163)
    for kind in range(6):
        b = abi.boundary_flux(kind, G, d["frames"], d["u_left"], d["u_right"])
        assert common.rel_err(b, d[f"bflux{kind}"]) < 1e-13


@pytest.mark.gpu
@pytest.mark.parametrize("name", [c for c in CASES if "parts" not in c])
def test_gpu_golden_case(lib, name):
    d = _load(name)
    meshes = _meshes(d)
    gpu = abi.Solver(meshes[0].view, scheme=int(d["scheme"]), degree=int(d["degree"]), rk_order=int(d["rk_order"]))
    common.set_vortex_case(gpu)
    gpu.set_solution(d["u0_0"])
    gpu.set_time(float(d["t0"]))
    common.assert_parity(gpu.get_residual(), d["residual_0"], 1e-10)
    gpu.rk_update(float(d["t0"]), float(d["dt"]), int(d["steps"]))
    common.assert_parity(gpu.get_solution(), d["u1_0"], 1e-10)
