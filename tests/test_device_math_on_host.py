"""The device math of minicfd_b200/csrc/riemann.cuh (every function there is __host__ __device__), run on
the CPU through tests/native/riemann_host.cu and compared with the oracle.  A development aid for the
`-m "not gpu"` suite: it catches formula errors before a GPU is involved.  The GPU parity tests
(tests/test_gpu_parity.py) remain the tests of the product path.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle
from tests import common

HERE = os.path.dirname(os.path.abspath(__file__))
c_f64p = C.POINTER(C.c_double)


@pytest.fixture(scope="module")
def host():
    subprocess.check_call(["make", "-C", os.path.join(HERE, "native")], stdout=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(HERE, "native", "_build", "libriemann_host.so"))
    lib.host_riemann_flux.argtypes = [C.c_int, C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    lib.host_boundary_flux.argtypes = [C.c_int, C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    lib.host_expansion_powers.argtypes = [C.c_int, C.c_int64, c_f64p, C.c_double, c_f64p, c_f64p]
    return lib


def _p(a):
    return a.ctypes.data_as(c_f64p)


def _states(rng, n, lo, hi, vmax):
    rho = rng.uniform(lo, hi, n); p = rng.uniform(lo, hi, n)
    vel = rng.uniform(-vmax, vmax, (n, 3))
    return np.ascontiguousarray(np.column_stack([rho, rho[:, None] * vel, p / 0.4 + 0.5 * rho * (vel ** 2).sum(1)]))


def _frames(rng, n):
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    q[:, :, 2] *= np.sign(np.linalg.det(q))[:, None]
    return np.ascontiguousarray(np.transpose(q, (0, 2, 1)).reshape(n, 9))


def test_expansion_powers(host):
    rng = np.random.default_rng(1)
    r = np.concatenate([rng.uniform(1e-6, 1.0, 20000), 10.0 ** rng.uniform(-29, 0, 20000), [1.0 - 1e-16, 1e-30]])
    t = np.empty_like(r); s = np.empty_like(r)
    host.host_expansion_powers(1, r.size, _p(r), 1.0 / 7, _p(t), _p(s))
    assert np.abs(t / r ** (1.0 / 7) - 1).max() < 1e-14
    assert np.abs(s / r ** (-6.0 / 7) - 1).max() < 1e-14
    host.host_expansion_powers(0, r.size, _p(r), 0.2 / 1.4, _p(t), _p(s))
    assert np.abs(t / r ** (1.0 / 7) - 1).max() < 1e-14
    assert np.abs(s / r ** (-6.0 / 7) - 1).max() < 1e-14


@pytest.mark.parametrize("gamma", [1.4, 5.0 / 3.0])
@pytest.mark.parametrize("regime", ["smooth", "moderate", "strong"])
def test_exact_flux_matches_oracle(host, gamma, regime):
    rng = np.random.default_rng(17)
    n = 20000
    fr = _frames(rng, n)
    if regime == "smooth":      # neighbouring states differ by 1e-3, as across a face in a resolved flow
        ul = _states(rng, n, 0.5, 1.5, 1.5)
        ur = ul * (1 + 1e-3 * rng.uniform(-1, 1, ul.shape))
    elif regime == "moderate":
        ul, ur = _states(rng, n, 0.1, 2.0, 1.5), _states(rng, n, 0.1, 2.0, 1.5)
    else:                        # strong shocks, sonic rarefactions, near-vacuum and vacuum
        ul, ur = _states(rng, n, 0.05, 1.0, 0.5), _states(rng, n, 0.05, 1.0, 0.5)
        push = rng.uniform(-6.0, 6.0, n)
        for u, sgn in ((ul, -1.0), (ur, +1.0)):
            rho = u[:, 0].copy(); v = u[:, 1:4] / rho[:, None]
            p = 0.4 * (u[:, 4] - 0.5 * rho * (v ** 2).sum(1))
            v = v + sgn * push[:, None] * fr[:, 0:3]
            u[:, 1:4] = rho[:, None] * v
            u[:, 4] = p / (gamma - 1) + 0.5 * rho * (v ** 2).sum(1)
    a = np.empty_like(ul)
    assert host.host_riemann_flux(0, gamma, n, _p(fr), _p(ul), _p(ur), _p(a)) == 0
    b = oracle.riemann_flux(gamma, fr, ul, ur)
    scale = np.maximum(np.abs(b).max(axis=1, keepdims=True), 1e-30)
    err = np.abs(a - b) / scale
    # same iteration path, roundings differ: 1e-12; a solve whose f lands within rounding of a loop
    # threshold may take one Newton step more or fewer -> bounded by the solver's own 1e-8 tolerance
    assert np.median(err) < 1e-14
    assert (err > 1e-11).mean() < 1e-3
    assert err.max() < 1e-7


def test_boundary_flux_matches_oracle(host):
    rng = np.random.default_rng(23)
    n = 5000
    ui, uo, fr = _states(rng, n, 0.5, 1.5, 2.0), _states(rng, n, 0.5, 1.5, 2.0), _frames(rng, n)
    for kind in range(6):
        a = np.empty_like(ui)
        host.host_boundary_flux(kind, common.GAMMA, n, _p(fr), _p(ui), _p(uo), _p(a))
        b = oracle.boundary_flux(kind, common.GAMMA, fr, ui, uo)
        assert common.rel_err(a, b) < 1e-13, kind


def test_exact_flux_degenerate_data(host):
    """Data continuous across the face up to a few ulp (a smooth IC sampled at shared nodes): the sign
    of f(p_L), pure rounding noise, selects FindRoot's halving branch (exact.hpp:163-166).  The device
    math must take the reference's branch, so the results agree to rounding, not to Newton's 1e-8."""
    rng = np.random.default_rng(29)
    n = 50000
    fr = _frames(rng, n)
    ul = _states(rng, n, 0.5, 1.5, 1.5)
    ur = ul.copy()
    for k in range(5):  # +-3 ulp on every component
        ur[:, k] = ul[:, k] * (1.0 + rng.integers(-3, 4, n) * 2.0 ** -53)
    a = np.empty_like(ul)
    host.host_riemann_flux(0, 1.4, n, _p(fr), _p(ul), _p(ur), _p(a))
    b = oracle.riemann_flux(1.4, fr, ul, ur)
    err = np.abs(a - b).max(axis=1) / np.abs(b).max(axis=1)
    assert err.max() < 1e-13


def test_division_helpers(host):
    """rn_div must reproduce IEEE division (it feeds noise-level sign tests); fast_rcp / fast_rsqrt ~1 ulp.
    (Host seeds are float 1/x, 1/sqrt(x): at least as coarse as the 20-bit MUFU seeds of the device.)"""
    host.host_rn_div.argtypes = [C.c_int64, c_f64p, c_f64p, c_f64p]
    host.host_fast_ops.argtypes = [C.c_int64, c_f64p, c_f64p, c_f64p]
    rng = np.random.default_rng(31)
    n = 2_000_000
    a = rng.uniform(-3, 3, n) * 10.0 ** rng.integers(-3, 4, n)
    b = rng.uniform(0.01, 30, n)
    q = np.empty(n)
    host.host_rn_div(n, _p(a), _p(b), _p(q))
    assert (q != a / b).mean() < 1e-6
    x = rng.uniform(1e-3, 1e3, n)
    r = np.empty(n); s = np.empty(n)
    host.host_fast_ops(n, _p(x), _p(r), _p(s))
    assert np.abs(r * x - 1).max() < 5e-16
    assert np.abs(s * np.sqrt(x) - 1).max() < 5e-16
