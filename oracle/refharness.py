"""TEST INFRASTRUCTURE ONLY -- ctypes loader of oracle/_ref/libref_*.so (the reference's own
templates compiled by oracle/Makefile from /root/reference, see oracle/ref/harness.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (minicfd_b200/) never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
c_f64p = C.POINTER(C.c_double)

_NAMES = {  # (scheme, limiter) -> file stem
    (2, 0): "fr", (1, 0): "dgsem", (0, 0): "dg", (0, 1): "dg_lazy", (0, 2): "dg_eigen",
}


def lib_path(scheme, degree, limiter=0, cell=0, variant=""):
    """cell: mcfd_cell_type of the mesh (1 = TETRA_4: the libref_dgtet_* builds, modal DG only);
    variant: "" the -O2 build every test compares with, ".O0" / ".O3fma" the same templates compiled otherwise
    (oracle/Makefile NOISE_LIBS, tests/test_reference_noise_floor.py)"""
    stem = _NAMES[(scheme, limiter)]
    if cell == 1:
        stem = stem.replace("dg", "dgtet", 1)
    return os.path.join(REF_DIR, f"libref_{stem}_p{degree}{variant}.so")


def available(scheme, degree, limiter=0, cell=0, variant=""):
    return os.path.exists(lib_path(scheme, degree, limiter, cell, variant))


_cache = {}


def _load(scheme, degree, limiter, cell=0, variant=""):
    key = (scheme, degree, limiter, cell, variant)
    if key in _cache:
        return _cache[key]
    lib = C.CDLL(lib_path(scheme, degree, limiter, cell, variant))
    lib.ref_create.restype = C.c_void_p
    lib.ref_create.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_double]
    lib.ref_destroy.argtypes = [C.c_void_p]
    lib.ref_column_size.restype = C.c_int64
    lib.ref_column_size.argtypes = [C.c_void_p, C.c_int]
    lib.ref_set_boundary.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, c_f64p, C.c_int]
    lib.ref_approximate.argtypes = [C.c_void_p, C.c_int, c_f64p, C.c_int, C.c_double]
    lib.ref_get_solution.argtypes = [C.c_void_p, C.c_int, c_f64p]
    lib.ref_set_solution.argtypes = [C.c_void_p, C.POINTER(c_f64p)]
    lib.ref_get_residual.argtypes = [C.c_void_p, C.c_double, C.POINTER(c_f64p)]
    lib.ref_rk_update.restype = C.c_double
    lib.ref_rk_update.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_int]
    lib.ref_troubled_count.restype = C.c_int64
    lib.ref_troubled_count.argtypes = [C.c_void_p]
    if hasattr(lib, "ref_write_vtu"):
        lib.ref_write_vtu.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        lib.ref_gather_fields.argtypes = [C.c_void_p, C.c_int, c_f64p]
        lib.ref_vtk_sample.argtypes = [C.c_void_p, C.c_int, c_f64p, c_f64p]
    lib.ref_riemann_flux.argtypes = [C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    lib.ref_boundary_flux.argtypes = [C.c_int, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]
    if hasattr(lib, "ref_measure_l1_error"):
        lib.ref_measure_l1_error.argtypes = [C.c_void_p, C.c_int, C.c_int, c_f64p, C.c_int, C.c_double, c_f64p]
    if hasattr(lib, "ref_ausm3_flux"):
        lib.ref_ausm3_flux.argtypes = [C.c_int64, c_f64p, c_f64p, c_f64p]
        lib.ref_hllc2_flux.argtypes = [C.c_int64, c_f64p, c_f64p, c_f64p]
    assert lib.ref_scheme() == scheme and lib.ref_degree() == degree and lib.ref_limiter() == limiter
    _cache[key] = lib
    return lib


def _p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_f64p)


class RefSim:
    """The reference operator (+ RungeKutta) on one or more mesh parts given as mcfd_mesh structs."""

    def __init__(self, mesh_views, scheme=2, degree=2, limiter=0, rk_order=3, weno_w0=1e-3, weno_eps=1e-6, variant=""):
        self.lib = _load(scheme, degree, limiter, getattr(mesh_views[0], "cell_type", 0), variant)
        self.n_parts = len(mesh_views)
        arr_t = type(mesh_views[0]) * self.n_parts
        self._meshes = arr_t(*mesh_views)
        self._h = C.c_void_p(self.lib.ref_create(self.n_parts, C.cast(self._meshes, C.c_void_p), rk_order,
                                                 weno_w0, weno_eps))
        self.sizes = [self.lib.ref_column_size(self._h, r) for r in range(self.n_parts)]

    def set_boundary(self, group, kind, func=0, params=()):
        p = np.ascontiguousarray(params, dtype=np.float64)
        rc = self.lib.ref_set_boundary(self._h, group, kind, func, _p(p) if p.size else None, p.size)
        assert rc == 0

    def approximate(self, func, params, t=0.0):
        p = np.ascontiguousarray(params, dtype=np.float64)
        self.lib.ref_approximate(self._h, func, _p(p), p.size, t)

    def get_solution(self, part=0):
        out = np.empty(self.sizes[part])
        self.lib.ref_get_solution(self._h, part, _p(out))
        return out

    def set_solution(self, columns):
        cols = [np.ascontiguousarray(c, dtype=np.float64) for c in columns]
        ptrs = (c_f64p * self.n_parts)(*[_p(c) for c in cols])
        self.lib.ref_set_solution(self._h, ptrs)

    def get_residual(self, t=0.0):
        outs = [np.empty(n) for n in self.sizes]
        ptrs = (c_f64p * self.n_parts)(*[_p(o) for o in outs])
        self.lib.ref_get_residual(self._h, t, ptrs)
        return outs

    def rk_update(self, t, dt, n_steps):
        """Returns wall seconds spent in the reference's time loop."""
        return self.lib.ref_rk_update(self._h, t, dt, n_steps)

    def troubled_count(self):
        return self.lib.ref_troubled_count(self._h)

    def measure_l1_error(self, func, params, t, part=0):
        """Part::MeasureL1Error (mesh/part.hpp:1082-1096) over the reference's own cells of one part"""
        p = np.ascontiguousarray(params, dtype=np.float64)
        out = np.zeros(5)
        assert self.lib.ref_measure_l1_error(self._h, part, func, _p(p), p.size, t, _p(out)) == 0
        return out

    def write_vtu(self, directory, soln_name):
        """mesh::vtk::Writer<Part>::WriteSolutions for every part: <directory>/<soln_name>.pvtu + <soln_name>/<r>.vtu"""
        assert self.lib.ref_write_vtu(self._h, os.fsencode(directory), soln_name.encode()) == 0

    def vtk_sample(self, part=0):
        """LocalToGlobalAndValue at the 27 vtk::Hexahedron27 nodes of every local cell, full precision"""
        n_points = 27 * (self.sizes[part] // self.lib.ref_fields_per_cell())
        coords, values = np.empty((n_points, 3)), np.empty((n_points, 5))
        self.lib.ref_vtk_sample(self._h, part, _p(coords), _p(values))
        return coords, values

    def gather_fields(self, part=0):
        """Section::GatherFields: [fields_per_cell][n_local_cells]"""
        out = np.empty(self.sizes[part])
        self.lib.ref_gather_fields(self._h, part, _p(out))
        return out.reshape(self.lib.ref_fields_per_cell(), -1)

    def close(self):
        if self._h:
            self.lib.ref_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def riemann_flux(frames, u_left, u_right):
    lib = _load(2, 2, 0)
    out = np.empty_like(u_left)
    lib.ref_riemann_flux(u_left.shape[0], _p(frames), _p(u_left), _p(u_right), _p(out))
    return out


def boundary_flux(kind, frames, u_inner, u_given):
    lib = _load(2, 2, 0)
    out = np.empty_like(u_inner)
    rc = lib.ref_boundary_flux(kind, u_inner.shape[0], _p(frames), _p(u_inner), _p(u_given), _p(out))
    assert rc == 0
    return out


def ausm3_flux(prim_left, prim_right):
    """Ausm<Gas,3>::GetFluxUpwind (riemann/euler/ausm.hpp:153-217) on normal-frame primitives [n][5] = rho,u,v,w,p"""
    lib = _load(2, 2, 0)
    out = np.empty_like(prim_left)
    assert lib.ref_ausm3_flux(prim_left.shape[0], _p(prim_left), _p(prim_right), _p(out)) == 0
    return out


def hllc2_flux(prim_left, prim_right):
    """Hllc<Gas,2>::GetFluxUpwind (riemann/euler/hllc.hpp:100-180) on (rho,u,v,p) of [n][5] primitives; out[:,3] = 0"""
    lib = _load(2, 2, 0)
    out = np.empty_like(prim_left)
    assert lib.ref_hllc2_flux(prim_left.shape[0], _p(prim_left), _p(prim_right), _p(out)) == 0
    return out
