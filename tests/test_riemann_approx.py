"""the approximate Riemann solvers (Roe, HLLC, AUSM) in the rotated frame.

What pins them (the reference cannot run them through rotated::Euler in 3-D):
  AUSM : the reference's own Ausm<Gas,3> (riemann/euler/ausm.hpp:153-217), compiled, all five components;
  HLLC : the reference's Hllc<Gas,2> (hllc.hpp:100-180) with w = 0, and -- for the component the reference lacks --
         the property that exchanging the two tangential axes exchanges the two tangential fluxes;
  Roe  : /root/reference/python/riemann.py:397-443 (1-D) through the committed fixture, the same exchange property,
         and consistency F(u, u) = F(u) as test/riemann/euler/{ausm,hllc}.cpp assert it.
CPU: the plain-C oracle and the device source compiled for the host against those.  GPU: the CUDA path through
mcfd_riemann_flux and through mcfd_create (residual + RK steps) against the oracle.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import oracle, refharness
from tests import common

HERE = os.path.dirname(os.path.abspath(__file__))
G = common.GAMMA
KINDS = {"roe": abi.RIEMANN_ROE, "hllc": abi.RIEMANN_HLLC, "ausm": abi.RIEMANN_AUSM}
IDENTITY = np.array([1.0, 0, 0, 0, 1.0, 0, 0, 0, 1.0])
SWAPPED = np.array([1.0, 0, 0, 0, 0, 1.0, 0, 1.0, 0])   # n = x, t = z, b = y


def golden():
    return np.load(os.path.join(HERE, "golden", "riemann_approx.npz"))


def to_cons(prim):
    rho, u, v, w, p = prim.T
    return np.ascontiguousarray(np.column_stack([rho, rho * u, rho * v, rho * w,
                                                 p / (G - 1) + 0.5 * rho * (u * u + v * v + w * w)]))


def euler_flux(prim):
    cons = to_cons(prim)
    f = cons * prim[:, 1:2]
    f[:, 1] += prim[:, 4]
    f[:, 4] += prim[:, 4] * prim[:, 1]
    return f


def frames_of(frame, n):
    return np.ascontiguousarray(np.tile(frame, (n, 1)))


def random_frames(rng, n):
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    q[:, :, 2] *= np.sign(np.linalg.det(q))[:, None]
    return np.ascontiguousarray(np.transpose(q, (0, 2, 1)).reshape(n, 9))


def check_against_reference(flux_fn):
    """flux_fn(kind, frames, cons_left, cons_right) -> [n][5]; the checks every implementation must pass"""
    d = golden()
    left, right = d["left"], d["right"]
    n = left.shape[0]
    ident = frames_of(IDENTITY, n)
    # AUSM: all five components of Ausm<Gas,3>
    got = flux_fn(abi.RIEMANN_AUSM, ident, to_cons(left), to_cons(right))
    assert common.rel_err(got, d["ausm_flux"]) < 1e-13
    # HLLC: Hllc<Gas,2> with w = 0 ...
    l0, r0 = left.copy(), right.copy()
    l0[:, 3] = 0; r0[:, 3] = 0
    got = flux_fn(abi.RIEMANN_HLLC, ident, to_cons(l0), to_cons(r0))
    assert common.rel_err(got, d["hllc_flux_w0"]) < 1e-13
    assert np.abs(got[:, 3]).max() == 0.0
    # Roe: the reference's 1-D flux with v = w = 0
    l1, r1 = l0.copy(), r0.copy()
    l1[:, 2] = 0; r1[:, 2] = 0
    got = flux_fn(abi.RIEMANN_ROE, ident, to_cons(l1), to_cons(r1))
    assert common.col_err(got[:, [0, 1, 4]], d["roe_flux_1d"]) < 1e-12
    assert np.abs(got[:, 2:4]).max() == 0.0
    for kind in KINDS.values():
        # ... and the third component: a frame with t and b exchanged sees (v, w) exchanged, so the two tangential
        # fluxes come back exchanged in the frame and identical in global coordinates
        a = flux_fn(kind, ident, to_cons(left), to_cons(right))
        b = flux_fn(kind, frames_of(SWAPPED, n), to_cons(left), to_cons(right))
        assert common.rel_err(b, a) < 1e-13, kind
        # consistency, test/riemann/euler/ausm.cpp:41-46, hllc.cpp (F(u, u) - F(u) = 0 to rounding)
        same = flux_fn(kind, ident, to_cons(left), to_cons(left))
        assert common.rel_err(same, euler_flux(left)) < 2e-15, kind


def test_reference_known_answers():
    """the loose known-answer bounds of test/riemann/euler/ausm.cpp:30-34,54-79 on the committed reference output"""
    d = golden()
    exact = np.array([[0.426319, 0.927453, 1.5, -0.7, 0.303130], [0.426319, -0.927453, 1.5, -0.7, 0.303130],
                      [5.99924, 19.5975, 1.5, -0.7, 460.894], [0.575062, 19.59745, 1.5, -0.7, 460.8938],
                      [0.575113, -6.196328, 2.5, 0.3, 46.09504]])
    want = euler_flux(exact)
    for name in ("ausm_flux", "hllc_flux_w0"):
        got = d[name][:5]
        for comp, bound in ((0, 0.34), (1, 0.27)):  # (the energy bounds there are for v = w = 0)
            assert (np.abs(got[:, comp] - want[:, comp]) / np.abs(want[:, comp])).max() <= bound, (name, comp)


def test_oracle_approx_vs_reference(oracle_lib):
    check_against_reference(lambda kind, fr, l, r: oracle.riemann_flux(G, fr, l, r, riemann=kind))


def test_oracle_approx_vs_compiled_reference_random(oracle_lib):
    """fresh random states against the compiled Ausm<Gas,3> / Hllc<Gas,2> (only where oracle/_ref exists)"""
    if not refharness.available(abi.SCHEME_FR, 2):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(7)
    n = 5000
    def side():
        return np.ascontiguousarray(np.column_stack([rng.uniform(0.05, 4, n), rng.uniform(-3, 3, (n, 3)), rng.uniform(0.05, 4, n)]))
    left, right = side(), side()
    ident = frames_of(IDENTITY, n)
    got = oracle.riemann_flux(G, ident, to_cons(left), to_cons(right), riemann=abi.RIEMANN_AUSM)
    assert common.rel_err(got, refharness.ausm3_flux(left, right)) < 1e-13
    left[:, 3] = 0; right[:, 3] = 0
    got = oracle.riemann_flux(G, ident, to_cons(left), to_cons(right), riemann=abi.RIEMANN_HLLC)
    assert common.rel_err(got, refharness.hllc2_flux(left, right)) < 1e-13


@pytest.fixture(scope="module")
def host():
    subprocess.check_call(["make", "-C", os.path.join(HERE, "native")], stdout=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(HERE, "native", "_build", "libriemann_host.so"))
    c_f64p = C.POINTER(C.c_double)
    lib.host_riemann_flux.argtypes = [C.c_int, C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    return lib


def _host_flux(host):
    def fn(kind, fr, l, r):
        out = np.empty_like(l)
        p = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
        assert host.host_riemann_flux(kind, G, l.shape[0], p(fr), p(l), p(r), p(out)) == 0
        return out
    return fn


def test_device_source_on_host_vs_reference(host):
    """minicfd_b200/csrc/riemann.cuh (every function is __host__ __device__) compiled for the CPU: a development
    aid that catches formula errors without a GPU; the GPU tests below are the tests of the product path"""
    check_against_reference(_host_flux(host))


@pytest.mark.parametrize("name", KINDS)
def test_device_source_on_host_vs_oracle_rotated(host, oracle_lib, name):
    rng = np.random.default_rng(11)
    n = 4000
    prim = lambda: np.column_stack([rng.uniform(0.1, 3, n), rng.uniform(-2.5, 2.5, (n, 3)), rng.uniform(0.1, 3, n)])
    l, r, fr = to_cons(prim()), to_cons(prim()), random_frames(rng, n)
    assert common.rel_err(_host_flux(host)(KINDS[name], fr, l, r), oracle.riemann_flux(G, fr, l, r, riemann=KINDS[name])) < 1e-12


# ---- the CUDA path -------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_approx_vs_reference(lib):
    check_against_reference(lambda kind, fr, l, r: abi.riemann_flux(kind, G, fr, l, r))


@pytest.mark.gpu
@pytest.mark.parametrize("name", KINDS)
def test_gpu_approx_vs_oracle_rotated(lib, oracle_lib, name):
    rng = np.random.default_rng(12)
    n = 20000
    prim = lambda: np.column_stack([rng.uniform(0.1, 3, n), rng.uniform(-2.5, 2.5, (n, 3)), rng.uniform(0.1, 3, n)])
    l, r, fr = to_cons(prim()), to_cons(prim()), random_frames(rng, n)
    assert common.rel_err(abi.riemann_flux(KINDS[name], G, fr, l, r), oracle.riemann_flux(G, fr, l, r, riemann=KINDS[name])) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("name", KINDS)
@pytest.mark.parametrize("scheme,degree", [(abi.SCHEME_FR, 2), (abi.SCHEME_DGSEM, 3), (abi.SCHEME_DG_MODAL, 2)])
def test_gpu_residual_and_steps_per_solver(lib, oracle_lib, name, scheme, degree):
    """one residual + 10 RK3 steps per solver through mcfd_create, against the oracle running the same solver"""
    mesh = abi.BoxMesh((5, 4, 3), (0, 0, 0), (10, 10, 10), distortion=0.1)
    gpu = abi.Solver(mesh.view, scheme=scheme, degree=degree, riemann=KINDS[name], rk_order=3)
    cpu = oracle.OracleSim([mesh.view], scheme, degree, 3, riemann=KINDS[name])
    for sim in (gpu, cpu):
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    u = common.perturbed(cpu.get_solution(0))
    gpu.set_solution(u); cpu.set_solution([u])
    gpu.set_time(0.2)
    common.assert_parity(gpu.get_residual(), cpu.get_residual(0.2)[0], 1e-10)
    gpu.rk_update(0.2, 5e-3, 10); cpu.rk_update(0.2, 5e-3, 10)
    common.assert_parity(gpu.get_solution(), cpu.get_solution(0), 1e-10)
