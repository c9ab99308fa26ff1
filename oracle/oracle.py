"""TEST INFRASTRUCTURE ONLY -- ctypes loader of oracle/_build/liboracle.so, the plain-C
restatement of the path (oracle/euler_oracle.c).  Imported by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs; never by the product.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
c_f64p = C.POINTER(C.c_double)
_lib = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    lib.oracle_create.restype = C.c_void_p
    lib.oracle_create.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double]
    lib.oracle_destroy.argtypes = [C.c_void_p]
    lib.oracle_column_size.restype = C.c_int64
    lib.oracle_column_size.argtypes = [C.c_void_p, C.c_int]
    lib.oracle_set_boundary.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, c_f64p, C.c_int]
    lib.oracle_approximate.argtypes = [C.c_void_p, C.c_int, c_f64p, C.c_int, C.c_double]
    lib.oracle_get_solution.argtypes = [C.c_void_p, C.c_int, c_f64p]
    lib.oracle_set_solution.argtypes = [C.c_void_p, C.c_int, c_f64p]
    lib.oracle_get_residual.argtypes = [C.c_void_p, C.c_double, C.POINTER(c_f64p)]
    lib.oracle_rk_update.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int]
    lib.oracle_set_limiter.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double]
    lib.oracle_limit.argtypes = [C.c_void_p]
    lib.oracle_troubled_count.restype = C.c_int64
    lib.oracle_troubled_count.argtypes = [C.c_void_p]
    lib.oracle_riemann_flux.argtypes = [C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    lib.oracle_riemann_flux_kind.argtypes = [C.c_int, C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    lib.oracle_set_riemann.argtypes = [C.c_void_p, C.c_int]
    lib.oracle_measure_l1_error.argtypes = [C.c_void_p, C.c_int, C.c_int, c_f64p, C.c_int, C.c_double, c_f64p]
    lib.oracle_boundary_flux.argtypes = [C.c_int, C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    _lib = lib
    return lib


def _p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_f64p)


class OracleSim:
    """The C restatement on one or more mesh parts (mcfd_mesh structs), all parts in lock-step."""

    def __init__(self, mesh_views, scheme=2, degree=2, rk_order=3, gamma=1.4, limiter=0, weno_w0=1e-3, weno_eps=1e-6,
                 riemann=0):
        self.lib = load()
        self.n_parts = len(mesh_views)
        arr_t = type(mesh_views[0]) * self.n_parts
        self._meshes = arr_t(*mesh_views)
        h = self.lib.oracle_create(self.n_parts, C.cast(self._meshes, C.c_void_p), scheme, degree, rk_order, gamma)
        if not h:
            raise RuntimeError("oracle_create failed")
        self._h = C.c_void_p(h)
        if self.lib.oracle_set_limiter(self._h, limiter, weno_w0, weno_eps) != 0:
            raise RuntimeError("limiters need the modal scheme")
        if self.lib.oracle_set_riemann(self._h, riemann) != 0:
            raise RuntimeError("unknown Riemann solver kind")
        self.sizes = [self.lib.oracle_column_size(self._h, r) for r in range(self.n_parts)]

    def set_boundary(self, group, kind, func=0, params=()):
        p = np.ascontiguousarray(params, dtype=np.float64)
        rc = self.lib.oracle_set_boundary(self._h, group, kind, func, _p(p) if p.size else None, p.size)
        assert rc == 0

    def approximate(self, func, params, t=0.0):
        p = np.ascontiguousarray(params, dtype=np.float64)
        self.lib.oracle_approximate(self._h, func, _p(p), p.size, t)

    def get_solution(self, part=0):
        out = np.empty(self.sizes[part])
        self.lib.oracle_get_solution(self._h, part, _p(out))
        return out

    def set_solution(self, columns):
        for r, c in enumerate(columns):
            c = np.ascontiguousarray(c, dtype=np.float64)
            self.lib.oracle_set_solution(self._h, r, _p(c))
        self.lib.oracle_limit(self._h)  # SetSolutionColumn of the limited scheme limits (dg/general.hpp:247-250)

    def troubled_count(self):
        return self.lib.oracle_troubled_count(self._h)

    def measure_l1_error(self, func, params, t, part=0):
        """Part::MeasureL1Error (mesh/part.hpp:1082-1096) of one part"""
        p = np.ascontiguousarray(params, dtype=np.float64)
        out = np.zeros(5)
        self.lib.oracle_measure_l1_error(self._h, part, func, _p(p), p.size, t, _p(out))
        return out

    def get_residual(self, t=0.0):
        outs = [np.empty(n) for n in self.sizes]
        ptrs = (c_f64p * self.n_parts)(*[_p(o) for o in outs])
        self.lib.oracle_get_residual(self._h, t, ptrs)
        return outs

    def rk_update(self, t, dt, n_steps):
        self.lib.oracle_rk_update(self._h, t, dt, n_steps)

    def close(self):
        if self._h:
            self.lib.oracle_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def riemann_flux(gamma, frames, u_left, u_right, riemann=0):
    """riemann: mcfd_riemann kind (0 exact, 1 Roe, 2 HLLC, 3 AUSM); a fresh Riemann object per point"""
    out = np.empty_like(u_left)
    load().oracle_riemann_flux_kind(riemann, gamma, u_left.shape[0], _p(frames), _p(u_left), _p(u_right), _p(out))
    return out


def boundary_flux(kind, gamma, frames, u_inner, u_given):
    out = np.empty_like(u_inner)
    load().oracle_boundary_flux(kind, gamma, u_inner.shape[0], _p(frames), _p(u_inner), _p(u_given), _p(out))
    return out
