"""ctypes mirror of include/minicfd_b200.h -- glue for tests and bench.py only.

The product is the C-ABI shared library `minicfd_b200/lib/libminicfd_b200.so` (host C++ + sm_100a
CUDA).  This module declares every entry point of the header with plain pointers and sizes, and
fails loudly when the library is missing: there is no Python or CPU fallback.
"""
import ctypes as C
import os

import numpy as np

K = 5
_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libminicfd_b200.so")

OK, ERR_ARG, ERR_CUDA, ERR_UNSUPPORTED, ERR_NCCL, ERR_STATE = range(6)
SCHEME_DG_MODAL, SCHEME_DGSEM, SCHEME_FR = 0, 1, 2
CELL_HEXA_8, CELL_TETRA_4 = 0, 1
RIEMANN_EXACT, RIEMANN_ROE, RIEMANN_HLLC, RIEMANN_AUSM = 0, 1, 2, 3
LIMITER_NONE, LIMITER_LAZY, LIMITER_EIGEN, LIMITER_DUMMY = 0, 1, 2, 3
BC_SOLID_WALL, BC_SUPERSONIC_INLET, BC_SUPERSONIC_OUTLET, BC_SUBSONIC_INLET, BC_SUBSONIC_OUTLET, BC_SMART = range(6)
FUNC_HOST_VALUES, FUNC_CONSTANT, FUNC_VORTEX, FUNC_PLANE_SPLIT = 0, 1, 2, 3
FLAG_FORCE_PER_NODE_METRICS, FLAG_KEEP_FACE_ORDER, FLAG_NO_SHAPE_CLASSES = 1, 4, 8

c_i64p = C.POINTER(C.c_int64)
c_i32p = C.POINTER(C.c_int32)
c_f64p = C.POINTER(C.c_double)


class Mesh(C.Structure):
    """struct mcfd_mesh"""
    _fields_ = [
        ("n_nodes", C.c_int64), ("node_xyz", c_f64p),
        ("n_local_cells", C.c_int64), ("n_inner_cells", C.c_int64), ("n_ghost_cells", C.c_int64),
        ("cell_nodes", c_i64p),
        ("n_local_faces", C.c_int64), ("n_ghost_faces", C.c_int64), ("n_boundary_faces", C.c_int64),
        ("face_holder", c_i64p), ("face_sharer", c_i64p), ("face_nodes", c_i64p),
        ("bface_group", c_i32p), ("n_bc_groups", C.c_int32),
        ("n_neighbours", C.c_int32), ("nbr_rank", c_i32p),
        ("send_ptr", c_i64p), ("send_cells", c_i64p), ("recv_ptr", c_i64p),
        ("cell_type", C.c_int32), ("reserved", C.c_int32),
    ]


class Config(C.Structure):
    """struct mcfd_config"""
    _fields_ = [
        ("scheme", C.c_int32), ("degree", C.c_int32), ("riemann", C.c_int32), ("limiter", C.c_int32),
        ("rk_order", C.c_int32), ("device", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
        ("gamma", C.c_double), ("weno_w0", C.c_double), ("weno_eps", C.c_double),
        ("flags", C.c_int32), ("reserved", C.c_int32),
    ]


class Timing(C.Structure):
    """struct mcfd_timing"""
    _fields_ = [
        ("ms_total", C.c_double), ("ms_face", C.c_double), ("ms_cell", C.c_double),
        ("n_face_launches", C.c_int64), ("n_cell_launches", C.c_int64), ("n_other_launches", C.c_int64),
    ]


class GlobalMesh(C.Structure):
    """mirror of mcfd_global_mesh"""
    _fields_ = [
        ("n_nodes", C.c_int64), ("node_xyz", c_f64p), ("n_cells", C.c_int64), ("cell_nodes", c_i64p),
        ("n_boundary_faces", C.c_int64), ("bface_nodes", c_i64p), ("bface_group", c_i32p), ("n_bc_groups", C.c_int32),
    ]


# name -> (restype, argtypes): every symbol include/minicfd_b200.h declares
SIGNATURES = {
    "mcfd_partitioner_create": (C.c_int, [C.POINTER(GlobalMesh), c_i32p, C.c_int32, C.POINTER(C.c_void_p)]),
    "mcfd_partitioner_destroy": (None, [C.c_void_p]),
    "mcfd_partitioner_dual_range": (c_i64p, [C.c_void_p]),
    "mcfd_partitioner_dual_index": (c_i64p, [C.c_void_p]),
    "mcfd_partitioner_node_parts": (c_i32p, [C.c_void_p]),
    "mcfd_partitioner_face_parts": (c_i32p, [C.c_void_p]),
    "mcfd_partitioner_cells_new_to_old": (c_i32p, [C.c_void_p]),
    "mcfd_partitioner_nodes_new_to_old": (c_i32p, [C.c_void_p]),
    "mcfd_partitioner_info_text": (C.c_int, [C.c_void_p, C.c_int32, C.c_char_p, C.c_int64, C.POINTER(C.c_int64)]),
    "mcfd_partitioner_part_mesh": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]),
    "mcfd_get_new_order": (C.c_int, [c_i32p, C.c_int32, c_i32p, c_i32p]),
    "mcfd_shuffle_connectivity": (C.c_int, [c_i32p, C.c_int32, c_i32p, C.c_int32, C.c_int32, c_i32p]),
    "mcfd_get_node_parts": (C.c_int, [C.c_int32, C.c_int32, c_i32p, c_i32p, c_i32p, C.c_int32, c_i32p]),
    "mcfd_box_mesh_create": (C.c_int, [c_i64p, c_f64p, c_f64p, c_i32p, C.c_int32, C.c_double, C.POINTER(C.c_void_p)]),
    "mcfd_box_mesh_create_ex": (C.c_int, [c_i64p, c_f64p, c_f64p, c_i32p, C.c_int32, C.c_double, c_i64p, c_i64p,
                                          C.POINTER(C.c_void_p)]),
    "mcfd_box_mesh_split_tetra": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "mcfd_box_mesh_view": (C.POINTER(Mesh), [C.c_void_p]),
    "mcfd_box_mesh_cell_global_ids": (c_i64p, [C.c_void_p]),
    "mcfd_box_mesh_destroy": (None, [C.c_void_p]),
    "mcfd_create": (C.c_int, [C.POINTER(Config), C.POINTER(Mesh), C.POINTER(C.c_void_p)]),
    "mcfd_destroy": (C.c_int, [C.c_void_p]),
    "mcfd_last_error": (C.c_char_p, [C.c_void_p]),
    "mcfd_column_size": (C.c_int64, [C.c_void_p]),
    "mcfd_fields_per_cell": (C.c_int32, [C.c_void_p]),
    "mcfd_set_time": (C.c_int, [C.c_void_p, C.c_double]),
    "mcfd_set_solution": (C.c_int, [C.c_void_p, c_f64p, C.c_int64]),
    "mcfd_get_solution": (C.c_int, [C.c_void_p, c_f64p, C.c_int64]),
    "mcfd_get_residual": (C.c_int, [C.c_void_p, c_f64p, C.c_int64]),
    "mcfd_gather_fields": (C.c_int, [C.c_void_p, c_f64p, C.c_int64]),
    "mcfd_scatter_fields": (C.c_int, [C.c_void_p, c_f64p, C.c_int64]),
    "mcfd_vtk_sample": (C.c_int, [C.c_void_p, c_f64p, c_f64p, C.c_int64]),
    "mcfd_write_vtu": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p)]),
    "mcfd_write_vtu_points": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(C.c_char_p), C.c_int32, C.c_int32, C.c_int64,
                                        c_f64p, c_f64p]),
    "mcfd_halo_transport": (C.c_int, [C.c_void_p]),
    "mcfd_approximate": (C.c_int, [C.c_void_p, C.c_int32, c_f64p, C.c_int32, C.c_double]),
    "mcfd_set_boundary": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_f64p, C.c_int32]),
    "mcfd_bc_point_count": (C.c_int64, [C.c_void_p, C.c_int32]),
    "mcfd_bc_points": (C.c_int, [C.c_void_p, C.c_int32, c_f64p, C.c_int64]),
    "mcfd_set_bc_values": (C.c_int, [C.c_void_p, C.c_int32, c_f64p, C.c_int64]),
    "mcfd_rk_update": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_int32]),
    "mcfd_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "mcfd_attach_nccl": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]),
    "mcfd_detach_nccl": (C.c_int, [C.c_void_p]),
    "mcfd_measure_l1_error": (C.c_int, [C.c_void_p, C.c_int32, c_f64p, C.c_int32, C.c_double, c_f64p]),
    "mcfd_measure_l1_error_static": (C.c_int, [C.c_void_p, C.c_int32, c_f64p, C.c_int32, C.c_double, c_f64p]),
    "mcfd_set_ghost_solution": (C.c_int, [C.c_void_p, c_f64p, C.c_int64]),
    "mcfd_get_timing": (C.c_int, [C.c_void_p, C.POINTER(Timing)]),
    "mcfd_troubled_cells": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]),
    "mcfd_set_profiling": (C.c_int, [C.c_void_p, C.c_int32]),
    "mcfd_shape_classes": (C.c_int32, [C.c_void_p, C.POINTER(C.c_int32)]),
    "mcfd_mesh_shape_classes": (C.c_int, [C.POINTER(Config), C.POINTER(Mesh), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                          C.POINTER(C.c_int32)]),
    "mcfd_struct_size": (C.c_int64, [C.c_int32]),
    "mcfd_device_solution": (C.c_void_p, [C.c_void_p]),
    "mcfd_stream": (C.c_void_p, [C.c_void_p]),
    "mcfd_riemann_flux": (C.c_int, [C.c_int32, C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]),
    "mcfd_boundary_flux": (C.c_int, [C.c_int32, C.c_double, C.c_int64, c_f64p, c_f64p, c_f64p, c_f64p]),
}

_lib = None


def load(path=None):
    """dlopen the product library and type every entry point.  Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(path)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def f64p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_f64p)


class BoxMesh:
    """Owner of a mcfd_box_mesh; `.view` is the mcfd_mesh struct (pointers into C++ memory)."""

    def __init__(self, n, lo=(0.0, 0.0, 0.0), hi=(1.0, 1.0, 1.0), parts=(1, 1, 1), rank=0, distortion=0.0, lib=None,
                 hole=None, tetra=False):
        """hole = ((i0, j0, k0), (i1, j1, k1)): cells i0 <= i < i1, ... are left out (boundary group 6 around them);
        tetra = True: every hexahedron cut into six TETRA_4 (mcfd_box_mesh_split_tetra)."""
        self._lib = lib or load()
        self.n, self.lo, self.hi, self.parts, self.rank = tuple(n), tuple(lo), tuple(hi), tuple(parts), rank
        n_a = (C.c_int64 * 3)(*n)
        lo_a = (C.c_double * 3)(*lo)
        hi_a = (C.c_double * 3)(*hi)
        p_a = (C.c_int32 * 3)(*parts)
        self._h = C.c_void_p()
        if hole is None:
            rc = self._lib.mcfd_box_mesh_create(n_a, lo_a, hi_a, p_a, rank, distortion, C.byref(self._h))
        else:
            rc = self._lib.mcfd_box_mesh_create_ex(n_a, lo_a, hi_a, p_a, rank, distortion, (C.c_int64 * 3)(*hole[0]),
                                                   (C.c_int64 * 3)(*hole[1]), C.byref(self._h))
        if rc != OK:
            raise ValueError(f"mcfd_box_mesh_create failed: {rc}")
        if tetra:
            hexa, self._h = self._h, C.c_void_p()
            rc = self._lib.mcfd_box_mesh_split_tetra(hexa, C.byref(self._h))
            self._lib.mcfd_box_mesh_destroy(hexa)
            if rc != OK:
                raise ValueError(f"mcfd_box_mesh_split_tetra failed: {rc}")
        self.view = self._lib.mcfd_box_mesh_view(self._h).contents
        self.view._owner = self   # the struct points into this object's C++ arrays: `BoxMesh(...).view` alone keeps them alive

    @property
    def n_local_cells(self):
        return self.view.n_local_cells

    def cell_global_ids(self):
        n = self.view.n_local_cells + self.view.n_ghost_cells
        return np.ctypeslib.as_array(self._lib.mcfd_box_mesh_cell_global_ids(self._h), shape=(n,)).copy()

    def arrays(self):
        """Copies of the mesh arrays as numpy (for the oracle restatement and inspection)."""
        v = self.view
        nc = v.n_local_cells + v.n_ghost_cells
        nf = v.n_local_faces + v.n_ghost_faces + v.n_boundary_faces
        as_arr = np.ctypeslib.as_array
        npe, fnpe = (4, 3) if v.cell_type == 1 else (8, 4)
        out = dict(
            node_xyz=as_arr(v.node_xyz, shape=(v.n_nodes, 3)).copy(),
            cell_nodes=as_arr(v.cell_nodes, shape=(nc, npe)).copy(),
            face_holder=as_arr(v.face_holder, shape=(nf,)).copy(),
            face_sharer=as_arr(v.face_sharer, shape=(nf,)).copy(),
            face_nodes=as_arr(v.face_nodes, shape=(nf, fnpe)).copy(),
            bface_group=(as_arr(v.bface_group, shape=(v.n_boundary_faces,)).copy()
                         if v.n_boundary_faces else np.zeros(0, np.int32)),
            nbr_rank=(as_arr(v.nbr_rank, shape=(v.n_neighbours,)).copy() if v.n_neighbours else np.zeros(0, np.int32)),
            send_ptr=as_arr(v.send_ptr, shape=(v.n_neighbours + 1,)).copy(),
            recv_ptr=as_arr(v.recv_ptr, shape=(v.n_neighbours + 1,)).copy(),
        )
        ns = int(out["send_ptr"][-1])
        out["send_cells"] = as_arr(v.send_cells, shape=(ns,)).copy() if ns else np.zeros(0, np.int64)
        return out

    def close(self):
        if self._h:
            self._lib.mcfd_box_mesh_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PartMesh(BoxMesh):
    """One part built by the general partition pipeline (same owner type and accessors as BoxMesh)."""

    def __init__(self, lib, handle):
        self._lib, self._h = lib, handle
        self.view = self._lib.mcfd_box_mesh_view(self._h).contents
        self.view._owner = self


def i32p(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_i32p)


class Partitioner:
    """mcfd_partitioner: dual graph, node / face parts, new orders, partition text and part meshes of a global
    hexahedral mesh for a given cell -> part vector."""

    def __init__(self, node_xyz, cell_nodes, cell_parts, n_parts, bface_nodes=None, bface_group=None, n_bc_groups=0, lib=None):
        self._lib = lib or load()
        self._keep = dict(
            xyz=np.ascontiguousarray(node_xyz, dtype=np.float64), cells=np.ascontiguousarray(cell_nodes, dtype=np.int64),
            parts=np.ascontiguousarray(cell_parts, dtype=np.int32),
            bnodes=np.ascontiguousarray(bface_nodes if bface_nodes is not None else np.zeros((0, 4)), dtype=np.int64),
            bgroup=np.ascontiguousarray(bface_group if bface_group is not None else np.zeros(0), dtype=np.int32))
        k = self._keep
        self.n_nodes, self.n_cells, self.n_bfaces, self.n_parts = len(k["xyz"]), len(k["cells"]), len(k["bnodes"]), n_parts
        gm = GlobalMesh(self.n_nodes, f64p(k["xyz"]), self.n_cells, k["cells"].ctypes.data_as(c_i64p), self.n_bfaces,
                        k["bnodes"].ctypes.data_as(c_i64p), i32p(k["bgroup"]), n_bc_groups)
        self._h = C.c_void_p()
        rc = self._lib.mcfd_partitioner_create(C.byref(gm), i32p(k["parts"]), n_parts, C.byref(self._h))
        if rc != OK:
            raise ValueError(f"mcfd_partitioner_create failed: {rc}")

    def dual_graph(self):
        r = np.ctypeslib.as_array(self._lib.mcfd_partitioner_dual_range(self._h), shape=(self.n_cells + 1,)).copy()
        i = np.ctypeslib.as_array(self._lib.mcfd_partitioner_dual_index(self._h), shape=(int(r[-1]),)).copy()
        return r, i

    def _i32(self, fn, n):
        return np.ctypeslib.as_array(fn(self._h), shape=(n,)).copy() if n else np.zeros(0, np.int32)

    def node_parts(self):
        return self._i32(self._lib.mcfd_partitioner_node_parts, self.n_nodes)

    def face_parts(self):
        return self._i32(self._lib.mcfd_partitioner_face_parts, self.n_bfaces)

    def cells_new_to_old(self):
        return self._i32(self._lib.mcfd_partitioner_cells_new_to_old, self.n_cells)

    def nodes_new_to_old(self):
        return self._i32(self._lib.mcfd_partitioner_nodes_new_to_old, self.n_nodes)

    def info_text(self, part):
        need = C.c_int64(0)
        assert self._lib.mcfd_partitioner_info_text(self._h, part, None, 0, C.byref(need)) == OK
        buf = C.create_string_buffer(need.value)
        assert self._lib.mcfd_partitioner_info_text(self._h, part, buf, need.value, C.byref(need)) == OK
        return buf.raw.decode()

    def part_mesh(self, part):
        h = C.c_void_p()
        rc = self._lib.mcfd_partitioner_part_mesh(self._h, part, C.byref(h))
        if rc != OK:
            raise ValueError(f"mcfd_partitioner_part_mesh failed: {rc}")
        return PartMesh(self._lib, h)

    def close(self):
        if self._h:
            self._lib.mcfd_partitioner_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def mesh_shape_classes(mesh_view, scheme=SCHEME_DG_MODAL, degree=2, flags=0, lib=None):
    """(shapes, classes, per_point_frames) of a mesh from the host tables alone -- no GPU involved"""
    lib = lib or load()
    cfg = Config(scheme, degree, RIEMANN_EXACT, LIMITER_NONE, 3, 0, 0, 1, 1.4, 1e-3, 1e-6, flags, 0)
    a, b, c = C.c_int32(0), C.c_int32(0), C.c_int32(0)
    rc = lib.mcfd_mesh_shape_classes(C.byref(cfg), C.byref(mesh_view), C.byref(a), C.byref(b), C.byref(c))
    if rc != OK:
        raise RuntimeError(f"mcfd_mesh_shape_classes failed ({rc}): {lib.mcfd_last_error(None).decode()}")
    return int(a.value), int(b.value), bool(c.value)


class Solver:
    """Thin RAII wrapper over an mcfd_solver handle.  Every call raises on a non-zero status."""

    def __init__(self, mesh_view, scheme=SCHEME_FR, degree=2, riemann=RIEMANN_EXACT, limiter=LIMITER_NONE, rk_order=3,
                 device=0, rank=0, world=1, gamma=1.4, weno_w0=1e-3, weno_eps=1e-6, flags=0, lib=None):
        self._lib = lib or load()
        self.cfg = Config(scheme, degree, riemann, limiter, rk_order, device, rank, world, gamma, weno_w0, weno_eps,
                          flags, 0)
        self._h = C.c_void_p()
        rc = self._lib.mcfd_create(C.byref(self.cfg), C.byref(mesh_view), C.byref(self._h))
        if rc != OK:
            msg = self._lib.mcfd_last_error(None)
            raise RuntimeError(f"mcfd_create failed ({rc}): {msg.decode() if msg else ''}")
        self.n = self._lib.mcfd_column_size(self._h)
        self.fields_per_cell = self._lib.mcfd_fields_per_cell(self._h)

    def _check(self, rc):
        if rc != OK:
            msg = self._lib.mcfd_last_error(self._h)
            raise RuntimeError(f"mcfd call failed ({rc}): {msg.decode() if msg else ''}")

    def set_boundary(self, group, kind, func=FUNC_HOST_VALUES, params=()):
        p = np.ascontiguousarray(params, dtype=np.float64)
        self._check(self._lib.mcfd_set_boundary(self._h, group, kind, func, f64p(p) if p.size else None, p.size))

    def approximate(self, func, params, t=0.0):
        p = np.ascontiguousarray(params, dtype=np.float64)
        self._check(self._lib.mcfd_approximate(self._h, func, f64p(p), p.size, t))

    def measure_l1_error(self, func, params, t):
        p = np.ascontiguousarray(params, dtype=np.float64)
        out = np.zeros(5)
        self._check(self._lib.mcfd_measure_l1_error(self._h, func, f64p(p), p.size, t, f64p(out)))
        return out

    def measure_l1_error_static(self, func, params, t):
        """mcfd_measure_l1_error_static: the same sums without any allocation inside the library (partial sums in the
        library's fixed 2.6 MB device array; at most 16 777 216 quadrature points per part)."""
        p = np.ascontiguousarray(params, dtype=np.float64)
        out = np.zeros(5)
        self._check(self._lib.mcfd_measure_l1_error_static(self._h, func, f64p(p), p.size, t, f64p(out)))
        return out

    def set_time(self, t):
        self._check(self._lib.mcfd_set_time(self._h, t))

    def set_solution(self, col):
        col = np.ascontiguousarray(col, dtype=np.float64)
        self._check(self._lib.mcfd_set_solution(self._h, f64p(col), col.size))

    def get_solution(self, out=None):
        out = np.empty(self.n) if out is None else out
        self._check(self._lib.mcfd_get_solution(self._h, f64p(out), out.size))
        return out

    def get_residual(self, out=None):
        out = np.empty(self.n) if out is None else out
        self._check(self._lib.mcfd_get_residual(self._h, f64p(out), out.size))
        return out

    def gather_fields(self):
        """Section::GatherFields: [fields_per_cell][n_local_cells]"""
        out = np.empty(self.n)
        self._check(self._lib.mcfd_gather_fields(self._h, f64p(out), out.size))
        return out.reshape(self.fields_per_cell, -1)

    def scatter_fields(self, fields):
        fields = np.ascontiguousarray(fields, dtype=np.float64)
        self._check(self._lib.mcfd_scatter_fields(self._h, f64p(fields), fields.size))

    def vtk_sample(self):
        """vtk::Writer::Prepare: (coords [27 n_cells][3], values [27 n_cells][5])"""
        n_points = 27 * (self.n // self.fields_per_cell)
        coords, values = np.empty((n_points, 3)), np.empty((n_points, 5))
        self._check(self._lib.mcfd_vtk_sample(self._h, f64p(coords), f64p(values), n_points))
        return coords, values

    def write_vtu(self, directory, soln_name, field_names=None):
        names = None
        if field_names is not None:
            names = (C.c_char_p * 5)(*[n.encode() for n in field_names])
        self._check(self._lib.mcfd_write_vtu(self._h, os.fsencode(directory), soln_name.encode(), names))

    def set_ghost_solution(self, col):
        col = np.ascontiguousarray(col, dtype=np.float64)
        self._check(self._lib.mcfd_set_ghost_solution(self._h, f64p(col), col.size))

    def bc_points(self, group):
        n = self._lib.mcfd_bc_point_count(self._h, group)
        xyz = np.empty((n, 3))
        if n:
            self._check(self._lib.mcfd_bc_points(self._h, group, f64p(xyz), n))
        return xyz

    def set_bc_values(self, group, values):
        values = np.ascontiguousarray(values, dtype=np.float64)
        self._check(self._lib.mcfd_set_bc_values(self._h, group, f64p(values), values.shape[0]))

    def rk_update(self, t, dt, n_steps):
        self._check(self._lib.mcfd_rk_update(self._h, t, dt, n_steps))

    def attach_nccl(self, unique_id_bytes, rank, world):
        buf = C.create_string_buffer(bytes(unique_id_bytes), 128)
        self._check(self._lib.mcfd_attach_nccl(self._h, buf, rank, world))

    def detach_nccl(self):
        """collective: call on every rank before the solver goes (see include/minicfd_b200.h)"""
        self._check(self._lib.mcfd_detach_nccl(self._h))

    def halo_transport(self):
        """0 no communicator, 1 NCCL send/recv, 2 peer-memory stores + arrival counters"""
        return int(self._lib.mcfd_halo_transport(self._h))

    def timing(self):
        t = Timing()
        self._check(self._lib.mcfd_get_timing(self._h, C.byref(t)))
        return t

    def troubled_cells(self):
        """running number of cell rebuilds over every limiter pass since the solver was created"""
        n = C.c_int64(0)
        self._check(self._lib.mcfd_troubled_cells(self._h, None, C.byref(n)))
        return int(n.value)

    def troubled_flags(self):
        """int32[n_local]: 1 for the cells the most recent limiter pass rebuilt"""
        f = np.zeros(self.n // self.fields_per_cell, dtype=np.int32)
        self._check(self._lib.mcfd_troubled_cells(self._h, f.ctypes.data_as(C.c_void_p), None))
        return f

    def troubled_count(self):
        """cells rebuilt by all limiter passes so far (what the reference harness counts in Part::Reconstruct)"""
        n = C.c_int64(0)
        self._check(self._lib.mcfd_troubled_cells(self._h, None, C.byref(n)))
        return int(n.value)

    def shape_classes(self):
        """(shapes, classes) of a repetitive modal mesh; (0, 0): the general kernels run"""
        n = C.c_int32(0)
        return int(self._lib.mcfd_shape_classes(self._h, C.byref(n))), int(n.value)

    def set_profiling(self, on):
        self._check(self._lib.mcfd_set_profiling(self._h, int(on)))

    def close(self):
        if self._h:
            leaks = self.halo_transport() == 2   # peer links still open: mcfd_destroy has to keep the exported buffers
            self._lib.mcfd_destroy(self._h)
            self._h = C.c_void_p()
            if leaks:   # close() cannot detach by itself: mcfd_detach_nccl is collective
                import warnings
                msg = self._lib.mcfd_last_error(None)
                if msg and msg.startswith(b"mcfd_destroy"):
                    warnings.warn(msg.decode(), ResourceWarning)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id(lib=None):
    lib = lib or load()
    buf = C.create_string_buffer(128)
    rc = lib.mcfd_nccl_unique_id(buf)
    if rc != OK:
        raise RuntimeError(f"mcfd_nccl_unique_id failed ({rc})")
    return buf.raw


def riemann_flux(kind, gamma, frames, u_left, u_right, lib=None):
    lib = lib or load()
    n = u_left.shape[0]
    out = np.empty((n, 5))
    rc = lib.mcfd_riemann_flux(kind, gamma, n, f64p(frames), f64p(u_left), f64p(u_right), f64p(out))
    if rc != OK:
        raise RuntimeError(f"mcfd_riemann_flux failed ({rc})")
    return out


def boundary_flux(kind, gamma, frames, u_inner, u_given, lib=None):
    lib = lib or load()
    n = u_inner.shape[0]
    out = np.empty((n, 5))
    rc = lib.mcfd_boundary_flux(kind, gamma, n, f64p(frames), f64p(u_inner), f64p(u_given), f64p(out))
    if rc != OK:
        raise RuntimeError(f"mcfd_boundary_flux failed ({rc})")
    return out


def write_vtu_points(directory, soln_name, rank, world, coords, values, field_names=None, lib=None):
    """host half of mcfd_write_vtu: <directory>/<soln_name>/<rank>.vtu (+ .pvtu on rank 0) from sampled points"""
    lib = lib or load()
    coords = np.ascontiguousarray(coords, dtype=np.float64)
    values = np.ascontiguousarray(values, dtype=np.float64)
    names = (C.c_char_p * 5)(*[n.encode() for n in field_names]) if field_names is not None else None
    rc = lib.mcfd_write_vtu_points(os.fsencode(directory), soln_name.encode(), names, rank, world,
                                   coords.shape[0] // 27, f64p(coords), f64p(values))
    if rc != OK:
        raise RuntimeError(f"mcfd_write_vtu_points failed ({rc})")
