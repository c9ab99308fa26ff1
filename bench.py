#!/usr/bin/env python
"""bench.py -- DOF-RK-stage updates/s of the Euler residual + Runge-Kutta hot path.

  python bench.py --gpus N --steps K --warmup W          (N > 1: launched under torchrun, one rank per GPU)
  python bench.py --impl reference ...                   (the reference's own CPU templates, oracle/_ref)

Workload (BASELINE.json configs[1]): Euler, FR p=2 on Lobatto nodes (what demo/euler ships),
exact Riemann solver, RK3, synthetic 64^3 hexahedral box [0,10]^3, isentropic vortex, smart boundaries
carrying the exact state; FP64.  One "step" = one RungeKutta<3>::Update = 3 residual evaluations + stage
updates over the whole mesh.  metric = column.size() * stages * steps / time.
N > 1 is STRONG scaling: the same 64^3 mesh block-partitioned over N ranks, halo exchange per stage over peer memory (NCCL as the fallback).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

VORTEX = [5.0, 5.0, 5.0, 1.0, 1.0, 1.4]
STAGES = {1: 1, 2: 2, 3: 3}
BYTES_PER_UPDATE_RK3 = (40.0 + 48.0 + 48.0) / 3.0  # 45.33 B / DOF-stage-update


def set_boundaries(sim, abi, ic="vortex"):
    """vortex: smart boundaries carrying the exact vortex state on the x/y faces, slip walls on z (w = 0 there).
    sod (--ic sod, the limiter workloads): walls all round."""
    if ic == "sod":
        for g in range(6):
            sim.set_boundary(g, abi.BC_SOLID_WALL)
        return
    for g in range(4):
        sim.set_boundary(g, abi.BC_SMART, abi.FUNC_VORTEX, VORTEX)
    for g in (4, 5):
        sim.set_boundary(g, abi.BC_SOLID_WALL)


def set_initial(sim, abi, args):
    if getattr(args, "workload", "box") == "double_mach":
        double_mach_case(sim, abi)   # boundaries and the initial projection, double_mach.cpp:22-52
        return
    if getattr(args, "workload", "box") == "forward_step":
        forward_step_case(sim, abi)  # forward_step.cpp:5-40
        return
    if args.ic == "sod":  # an oblique Sod discontinuity through the box centre (Primitive(1,0,0,0,1) | (0.125,0,0,0,0.1))
        left = [1.0, 0.0, 0.0, 0.0, 2.5]; right = [0.125, 0.0, 0.0, 0.0, 0.25]
        sim.approximate(abi.FUNC_PLANE_SPLIT, [1.0, 0.3, 0.2, 0.5 * args.size, 0.0] + left + right, 0.0)
    else:
        sim.approximate(abi.FUNC_VORTEX, VORTEX, 0.0)


# ---- BASELINE.json configs[2]: double Mach reflection (demo/euler/double_mach.cpp:5-52) ----
DM_X_GAP = 1.0 / 6.0


def double_mach_states():
    """double_mach.cpp:5-20 : the states either side of the Mach-10 shock, tan 60, the speed of its foot"""
    import math
    gp, g, gm = 2.4, 1.4, 0.4
    rho_b, p_b, mach, a_b = 1.4, 1.0, 10.0, 1.0
    u_nu = mach * a_b
    rho_a = rho_b * (mach * mach * gp / 2) / (1 + mach * mach * gm / 2)
    p_a = p_b * (mach * mach * g - gm / 2) / (gp / 2)
    u_n_a = u_nu * (rho_a - rho_b) / rho_a
    tan60 = math.sqrt(3.0)
    cos30, sin30 = tan60 * 0.5, 0.5
    cons = lambda rho, u, v, w, p: [rho, rho * u, rho * v, rho * w, p / 0.4 + 0.5 * rho * (u * u + v * v + w * w)]
    return cons(rho_a, u_n_a * cos30, u_n_a * (-sin30), 0.0, p_a), cons(rho_b, 0.0, 0.0, 0.0, p_b), tan60, u_nu / cos30


def double_mach_mesh(abi, n, parts, rank):
    """4 x 1 x 0.2 box of cubes with h = 1 / n[1]; boundary group 6 = the bottom faces left of x = 1/6 ("Gap")"""
    import ctypes as C
    mesh = abi.BoxMesh(n, (0, 0, 0), (4.0, 1.0, n[2] / n[1]), parts=parts, rank=rank)
    a, v = mesh.arrays(), mesh.view
    nb = v.n_boundary_faces
    if nb:
        first = v.n_local_faces + v.n_ghost_faces
        centre = a["node_xyz"][a["face_nodes"][first:first + nb]].mean(axis=1)
        group = a["bface_group"].astype(np.int32).copy()
        group[(group == 2) & (centre[:, 0] < DM_X_GAP)] = 6
        mesh._gap_groups = np.ascontiguousarray(group)
        v.bface_group = mesh._gap_groups.ctypes.data_as(C.POINTER(C.c_int32))
    v.n_bc_groups = 7
    return mesh


def double_mach_case(sim, abi):
    after, before, tan60, u_x = double_mach_states()
    shock = [tan60, -1.0, 0.0, tan60 * DM_X_GAP, tan60 * u_x] + after + before   # FUNC_PLANE_SPLIT: n, offset, speed, L, R
    sim.set_boundary(3, abi.BC_SUPERSONIC_INLET, abi.FUNC_PLANE_SPLIT, shock)   # Top
    sim.set_boundary(0, abi.BC_SUPERSONIC_INLET, abi.FUNC_PLANE_SPLIT, shock)   # Left
    for g in (4, 5, 2):
        sim.set_boundary(g, abi.BC_SOLID_WALL)                                  # Back, Front, Bottom
    sim.set_boundary(1, abi.BC_SUPERSONIC_OUTLET)                               # Right
    sim.set_boundary(6, abi.BC_SUPERSONIC_OUTLET)                               # Gap
    sim.approximate(abi.FUNC_PLANE_SPLIT, shock, 0.0)


# ---- BASELINE.json configs[3]: forward-facing step (demo/euler/forward_step.cpp, input/forward_step.geo) on tetrahedra ----
# A 3 x 1 channel, 0.2 deep, with a 0.2-high step from x = 0.6: a box of cubes (h = 1 / n, n a multiple of 5) with
# the step's cells left out, every cube cut into six tetrahedra (gmsh, which meshes the .geo, is absent here).
# The scheme is the reference-backed form of the config: modal DG + weno::Eigen + exact Riemann; the
# "FR + HLLC" wording of BASELINE.json cannot run in the reference (FR is hexahedral, Hllc<Gas,3> does not exist).
def forward_step_shape(n):
    assert n % 5 == 0, "--cells must be a multiple of 5 for the forward step (the step is 0.2 high and starts at x = 0.6)"
    return (3 * n, n, max(2, n // 5)), ((3 * n // 5, 0, 0), (3 * n, n // 5, max(2, n // 5)))


def forward_step_mesh(abi, n, parts, rank):
    shape, hole = forward_step_shape(n)
    return abi.BoxMesh(shape, (0, 0, 0), (3.0, 1.0, shape[2] / n), parts=parts, rank=rank, hole=hole, tetra=True)


def forward_step_case(sim, abi):
    given = [1.4, 4.2, 0.0, 0.0, 1.0 / 0.4 + 0.5 * 1.4 * 9.0]         # Primitive(1.4, 3, 0, 0, 1): Mach 3
    sim.set_boundary(0, abi.BC_SUPERSONIC_INLET, abi.FUNC_CONSTANT, given)  # Left-Upper, Left-Lower
    sim.set_boundary(1, abi.BC_SUPERSONIC_OUTLET)                            # Right
    for g in (2, 3, 4, 5, 6):
        sim.set_boundary(g, abi.BC_SOLID_WALL)                               # floor ahead of the step, Top, Back, Front, Step
    sim.approximate(abi.FUNC_CONSTANT, given, 0.0)


def mesh_shape(args):
    """cells per direction of the workload's box"""
    if args.workload == "double_mach":
        return (4 * args.n, args.n, max(2, args.n // 5))
    if args.workload == "forward_step":
        return forward_step_shape(args.n)[0]
    return (args.n,) * 3


def n_cells(args):
    nx, ny, nz = mesh_shape(args)
    if args.workload == "forward_step":
        (i0, j0, k0), (i1, j1, k1) = forward_step_shape(args.n)[1]
        return 6 * (nx * ny * nz - (i1 - i0) * (j1 - j0) * (k1 - k0))
    return nx * ny * nz


def fields_per_cell(args):
    if args.scheme == "modal":
        return 5 * {1: 4, 2: 10, 3: 20}[args.degree]
    return 5 * (args.degree + 1) ** 3


def parts_for(world):
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}.get(world) or (world, 1, 1)


class ClockSampler:
    """SM clock / throttle reasons DURING the timed region (B200_PROFILING.md's clocks line).  NVML is polled from a
    thread of this process every 2 ms (the timed call sits in C code with the GIL released), because at N = 8 the whole
    timed region lasts a few milliseconds -- less than `nvidia-smi -lms` needs to produce its first line.  Falls back
    to the nvidia-smi recipe when pynvml is missing."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASON_BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.proc = None
        self.path = None
        self.thread = None
        self.samples = []
        self.bits = 0
        self.max_mhz = None
        self.running = False

    def _poll(self, nvml, handle):
        while self.running:
            try:
                self.samples.append(float(nvml.nvmlDeviceGetClockInfo(handle, nvml.NVML_CLOCK_SM)))
                self.bits |= int(nvml.nvmlDeviceGetCurrentClocksThrottleReasons(handle))
            except Exception:
                pass
            time.sleep(0.002)

    def prepare(self):
        """NVML initialisation takes milliseconds and varies from rank to rank: do it BEFORE the barrier that opens the
        timed region, so that start() only spawns the polling thread."""
        try:
            import pynvml as nvml
            nvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            index = int(visible.split(",")[self.gpu_index]) if visible and visible.replace(",", "").isdigit() else self.gpu_index
            self.handle = nvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(nvml.nvmlDeviceGetMaxClockInfo(self.handle, nvml.NVML_CLOCK_SM))
            nvml.nvmlDeviceGetClockInfo(self.handle, nvml.NVML_CLOCK_SM)
            self.nvml = nvml
        except Exception:
            self.nvml = None

    def start(self):
        if getattr(self, "nvml", None) is None and not hasattr(self, "nvml"):
            self.prepare()
        if self.nvml is not None:
            import threading
            self.running = True
            self.thread = threading.Thread(target=self._poll, args=(self.nvml, self.handle), daemon=True)
            self.thread.start()
            return
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu_index)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.thread:
            self.running = False
            self.thread.join(timeout=2)
            if self.samples:
                out["sm_mhz"] = float(np.median(self.samples)); out["sm_max_mhz"] = self.max_mhz
                out["samples"] = len(self.samples)
            out["reasons"] = sorted(nm for bit, nm in self.REASON_BITS.items() if self.bits & bit)
            out["source"] = "nvml, 2 ms polling during the timed region"
            return out
        if not self.proc:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                c = [x.strip() for x in line.split(",")]
                if len(c) < 9:
                    continue
                try:
                    sm.append(float(c[1])); mx.append(float(c[2]))
                except ValueError:
                    continue
                for nm, val in zip(names, c[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out["sm_mhz"] = float(np.median(sm)); out["sm_max_mhz"] = float(max(mx)); out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_model():
    """the silicon the CPU figures were taken on"""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def split_parts(p):
    px = py = pz = 1
    while px * py * pz < p:
        if px <= py and px <= pz: px *= 2
        elif py <= pz: py *= 2
        else: pz *= 2
    return px, py, pz


FULL_MESH_DOF_LIMIT = 6.0e7   # the reference arm runs the WHOLE mesh up to this size (64^3 p=2: 3.5e7 DOF, ~1 s per step
                              # on 16 threads); beyond it (128^3, 256^3 p=3) a sub-box of the same h, and says so


def reference_sample(threads, degree, scheme, args=None, full=False):
    """The reference's own templates (oracle/_ref), one mesh part per host thread (the reference's MPI layout, run
    as threads).  full=False: a bounded sample of the workload -- a sub-box of the mesh with the same h, dt, IC and BC
    kinds, 16^3 cells per part.  full=True (the reference arm): the whole n^3 mesh cut into one block per thread when
    it is small enough.  Falls back to the plain-C oracle port if oracle/_ref did not travel.
    Returns (sim, dof, threads, kind, description, meshes, cells)."""
    from minicfd_b200 import abi
    from oracle import refharness
    p = 1
    while p * 2 <= threads:
        p *= 2
    px, py, pz = split_parts(p)
    if args is not None and args.workload == "forward_step":
        # a coarser forward-step mesh stands in (same domain, boundary sections, scheme and limiter): the reference's
        # own templates on tetrahedra (oracle/_ref/libref_dgtet_*)
        n_ref = min(args.n, 20 if full else 10)
        shape = forward_step_shape(n_ref)[0]
        while shape[0] % px or shape[1] % py or shape[2] % pz:
            p //= 2; px, py, pz = split_parts(p)
        meshes = [forward_step_mesh(abi, n_ref, (px, py, pz), r) for r in range(p)]
        views = [m.view for m in meshes]
        from oracle import oracle as _oracle
        if refharness.available(scheme, degree, args.limiter_id, cell=1):
            sim, kind = refharness.RefSim(views, scheme, degree, args.limiter_id, 3), "reference"
        else:
            sim, kind, p = _oracle.OracleSim(views, scheme, degree, 3, limiter=args.limiter_id), "port", 1
        forward_step_case(sim, abi)
        if args.limiter_id:
            sim.set_solution([sim.get_solution(r) for r in range(len(views))])
        cells = sum(v.n_local_cells for v in views)
        desc = (("the whole " if n_ref == args.n else f"a coarser (h = 1/{n_ref} instead of 1/{args.n}) ") +
                f"forward-step mesh of {cells} tetrahedra, {len(views)} parts ({px}x{py}x{pz}) on {p} host threads")
        return sim, sum(sim.sizes), p, kind, desc, meshes, cells
    if args is not None and args.workload == "double_mach":
        # a coarser double-Mach mesh stands in (same domain, same boundary sections, same scheme and limiter)
        n_ref = min(args.n, 30 if full else 20)
        n = (4 * n_ref, n_ref, max(2, n_ref // 5))
        while n[0] % px or n[1] % py or n[2] % pz:   # fewer parts if the blocks do not divide
            p //= 2; px, py, pz = split_parts(p)
        meshes = [double_mach_mesh(abi, n, (px, py, pz), r) for r in range(p)]
        views = [m.view for m in meshes]
        from oracle import oracle as _oracle
        if refharness.available(scheme, degree, args.limiter_id):
            sim, kind = refharness.RefSim(views, scheme, degree, args.limiter_id, 3), "reference"
        else:
            sim, kind, p = _oracle.OracleSim(views, scheme, degree, 3, limiter=args.limiter_id), "port", 1
        double_mach_case(sim, abi)
        if args.limiter_id:
            sim.set_solution([sim.get_solution(r) for r in range(len(views))])
        whole = n_ref == args.n
        desc = (("the whole " if whole else f"a coarser (h = 1/{n_ref} instead of 1/{args.n}) ") +
                f"{n[0]}x{n[1]}x{n[2]}-cell double-Mach mesh, {len(views)} parts ({px}x{py}x{pz}) on {p} host threads")
        return sim, sum(sim.sizes), p, kind, desc, meshes, n[0] * n[1] * n[2]
    size, cells = (args.size, args.n) if args is not None else (10.0, 64)
    h = size / cells
    dof_full = cells ** 3 * (fields_per_cell(args) if args is not None else 135)
    whole = full and dof_full <= FULL_MESH_DOF_LIMIT and cells % max(px, py, pz) == 0
    n = (cells, cells, cells) if whole else (16 * px, 16 * py, 16 * pz)
    hi = tuple(h * x for x in n)
    meshes = [abi.BoxMesh(n, (0, 0, 0), hi, parts=(px, py, pz), rank=r) for r in range(p)]
    views = [m.view for m in meshes]
    limiter = args.limiter_id if args is not None else 0
    if refharness.available(scheme, degree, limiter):
        sim = refharness.RefSim(views, scheme, degree, limiter, 3)
        kind = "reference"
    else:
        from oracle import oracle
        sim = oracle.OracleSim(views, scheme, degree, 3, limiter=limiter)
        kind = "port"
        p = 1  # the C port steps its parts sequentially on one thread
    ic = args.ic if args is not None else "vortex"
    set_boundaries(sim, abi, ic)
    if args is not None:
        set_initial(sim, abi, args)
    else:
        sim.approximate(abi.FUNC_VORTEX, VORTEX, 0.0)
    dof = sum(sim.sizes)
    what = (f"the whole {cells}^3 mesh" if whole else
            f"{n[0]}x{n[1]}x{n[2]}-cell sub-box of the {cells}^3 mesh (same h={size:g}/{cells}, same dt, IC and BC kinds)")
    desc = f"{what}, {len(views)} parts ({px}x{py}x{pz}) on {p} host threads"
    return sim, dof, p, kind, desc, meshes, n[0] * n[1] * n[2]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    sim, dof, threads, kind, desc, _keep, cells = reference_sample(cores, args.degree, args.scheme_id, args, full=True)
    dt = args.dt
    t = 0.0
    for _ in range(args.warmup):
        sim.rk_update(t, dt, 1); t += dt
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sim.rk_update(t, dt, 1); t += dt
    wall = time.perf_counter() - t0
    value = dof * 3 * args.steps / wall
    config = workload_config(args)
    if cells != n_cells(args):   # a sub-box stood in for the mesh: say what actually ran, not what was asked for
        config.update(cells=cells, dof=dof, partition=desc,
                      workload=config["workload"] + f" -- SAMPLED on {desc}; DOF-updates/s is size-normalised")
    else:
        config.update(partition=desc)
    line = {
        "impl": "reference", "metric": "DOF-RK-stage updates/sec", "value": value, "unit": "DOF-updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config,
        "cpu_baseline": {"value": value, "unit": "DOF-updates/s", "cores": threads, "kind": kind, "sample": desc,
                         "cpu": cpu_model()},
        "e2e": {"value": value, "unit": "DOF-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_cores_available": cores,
    }
    print(json.dumps(line))


def workload_config(args):
    cells, fields = n_cells(args), fields_per_cell(args)
    if args.workload == "forward_step":
        what = (f"Euler {args.scheme.upper()} p={args.degree} RK3 exact-Riemann, forward-facing step (demo/euler/forward_step.cpp): "
                f"3 x 1 channel with a 0.2 step from x = 0.6, {cells} tetrahedra (cubes of h = 1/{args.n} cut in six), Mach-3 inlet, "
                "outlet right, walls elsewhere")
    elif args.workload == "double_mach":
        nx, ny, nz = mesh_shape(args)
        what = (f"Euler {args.scheme.upper()} p={args.degree} RK3 exact-Riemann, double Mach reflection (demo/euler/double_mach.cpp): "
                f"4 x 1 x {nz / ny:g} box of {nx}x{ny}x{nz} hexahedra (h = 1/{ny}), Mach-10 shock at 60 degrees, moving-shock inlets "
                "top/left, outlets right/gap, walls elsewhere")
    else:
        what = (f"Euler {args.scheme.upper()} p={args.degree} RK3 exact-Riemann, synthetic {args.n}^3 hexa box "
                f"[0,{args.size:g}]^3, " + ("isentropic vortex, smart x/y boundaries + z walls" if args.ic == "vortex"
                                             else "oblique Sod discontinuity, walls"))
    return {
        "workload": what + (f", weno::{args.limiter.capitalize()} limiter" if args.limiter != "none" else ""),
        "cells": cells, "dof": cells * fields, "dt": args.dt,
        "partition": "x".join(map(str, parts_for(args.gpus))),
        "cache": (f"state columns (3 x {cells * fields * 8 / 1e6:.0f} MB) + flux buffer + tables "
                  + ("exceed the 126 MB L2: inputs larger than L2, no flush" if 3 * cells * fields * 8 > 126e6
                     else "FIT in the 126 MB L2 and nothing is flushed: a parity-sized run, not a bench line")),
    }


def run_output(args):
    """SURVEY 8(f) row f2, measured: Section::GatherFields and vtk::Writer::Prepare on the device (kernel time from CUDA
    events inside the library, against the HBM peak) and through the C-ABI with host buffers; the .vtu text on a 32^3 mesh."""
    import tempfile as tf
    import torch
    from minicfd_b200 import abi
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    abi.load(args.lib)
    peak = 6650.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    out = {"metric": "solution output (row f2)", "config": workload_config(args), "peak_hbm_gbs": peak}
    mesh = abi.BoxMesh((args.n,) * 3, (0, 0, 0), (args.size,) * 3)
    solver = abi.Solver(mesh.view, scheme=args.scheme_id, degree=args.degree, rk_order=3, flags=args.flags)
    set_boundaries(solver, abi, args.ic)
    set_initial(solver, abi, args)
    n_cells, dof = args.n ** 3, solver.n
    for name, call, nbytes in (("gather_fields", solver.gather_fields, 16 * dof),
                               ("vtk_sample", solver.vtk_sample, 8 * dof + 27 * n_cells * 8 * 8 + n_cells * 24 * 8)):
        call()
        best_kernel, best_wall = 1e30, 1e30
        for _ in range(5):
            t0 = time.perf_counter()
            call()
            best_wall = min(best_wall, time.perf_counter() - t0)
            best_kernel = min(best_kernel, solver.timing().ms_total)
        out[name] = {"kernel_ms": best_kernel, "algorithmic_bytes": nbytes, "kernel_gbs": nbytes / best_kernel / 1e6,
                     "frac_of_hbm_peak": nbytes / best_kernel / 1e6 / peak, "c_abi_host_buffers_ms": best_wall * 1e3}
    small = abi.BoxMesh((32,) * 3, (0, 0, 0), (args.size,) * 3)
    s2 = abi.Solver(small.view, scheme=args.scheme_id, degree=args.degree, rk_order=3)
    set_boundaries(s2, abi, args.ic)
    set_initial(s2, abi, args)
    with tf.TemporaryDirectory() as tmp:
        t0 = time.perf_counter()
        s2.write_vtu(tmp, "Frame0")
        dt_w = time.perf_counter() - t0
        size = os.path.getsize(os.path.join(tmp, "Frame0", "0.vtu"))
    out["write_vtu_32cubed"] = {"seconds": dt_w, "file_bytes": size, "what": "sampling + the reference's ascii/base64 text, one host thread"}
    print(json.dumps(out))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from minicfd_b200 import abi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    abi.load(args.lib)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if args.workload == "double_mach":
        mesh = double_mach_mesh(abi, mesh_shape(args), parts_for(world), rank)
    elif args.workload == "forward_step":
        mesh = forward_step_mesh(abi, args.n, parts_for(world), rank)
    else:
        mesh = abi.BoxMesh((args.n,) * 3, (0, 0, 0), (args.size,) * 3, parts=parts_for(world), rank=rank)
    solver = abi.Solver(mesh.view, scheme=args.scheme_id, degree=args.degree, riemann=abi.RIEMANN_EXACT, rk_order=3,
                        limiter=args.limiter_id,
                        device=local_rank, rank=rank, world=world, flags=args.flags)
    if world > 1:
        ids = [abi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        solver.attach_nccl(ids[0], rank, world)
    set_boundaries(solver, abi, args.ic)
    set_initial(solver, abi, args)
    if args.limiter_id:   # SetSolutionColumn: the first limiter pass (dg/general.hpp:247-250)
        solver.set_solution(solver.get_solution())
    dof_local = solver.n
    dof_total = n_cells(args) * fields_per_cell(args)
    dt, t = args.dt, 0.0

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput: K steps in one mcfd_rk_update call, timed by CUDA events on the
    # library's stream (mcfd_timing.ms_total), max over ranks ----
    solver.rk_update(t, dt, args.warmup); t += dt * args.warmup
    sampler = ClockSampler(local_rank)
    sampler.prepare()
    barrier()
    sampler.start()
    solver.rk_update(t, dt, args.steps); t += dt * args.steps
    barrier()
    clocks = sampler.stop()
    tm = solver.timing()
    ms = torch.tensor([tm.ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = int(tm.n_face_launches + tm.n_cell_launches + tm.n_other_launches)
    value = dof_total * 3 * args.steps / (ms_total * 1e-3)
    # four more blocks of the same K steps, each timed the same way: how much the headline figure moves from block to
    # block (the contract's value above is the first block, exactly K steps)
    repeats = [ms_total / args.steps]
    for _ in range(4 if world == 1 else 2):
        barrier()
        solver.rk_update(t, dt, args.steps); t += dt * args.steps
        r_ms = torch.tensor([solver.timing().ms_total], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(r_ms, op=dist.ReduceOp.MAX)
        repeats.append(float(r_ms.item()) / args.steps)

    # ---- per-kernel durations (separate pass so the timed region above has no extra events) ----
    solver.set_profiling(True)
    solver.rk_update(t, dt, args.steps); t += dt * args.steps
    tp = solver.timing()
    solver.set_profiling(False)
    fq = {1: 3, 2: 6, 3: 12}[args.degree] if mesh.view.cell_type == abi.CELL_TETRA_4 else (args.degree + 1) ** 2
    n_face_pts = (mesh.view.n_local_faces + mesh.view.n_ghost_faces + mesh.view.n_boundary_faces) * fq
    stage_launches = 3 * args.steps
    ms_face_stage = tp.ms_face / stage_launches
    ms_cell_stage = tp.ms_cell / stage_launches
    face_bytes = dof_local * 8 + n_face_pts * 40            # read U once, write f* once
    cell_bytes = dof_local * (8 + 8 + 8 * 2.0 / 3.0) + n_face_pts * 40  # read U, write U', read U0 in 2 of 3 stages, read f*
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    dominant = "face_flux_kernel" if ms_face_stage >= ms_cell_stage else "cell_update_kernel"
    dom_ms = max(ms_face_stage, ms_cell_stage)
    dom_bytes = face_bytes if dominant == "face_flux_kernel" else cell_bytes
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
    # FP64 side of the roofline: the pipe's measured DFMA rate (minicfd_b200/csrc/tools/fp64_peak.cu on this pool's
    # B200, profiles/fp64_peak.json) and the FP64 operations the two kernels execute per face point / per node on this
    # workload (ncu source counters, profiles/fp64_work.json: DFMA = 2, DMUL/DADD = 1, active threads only)
    fp64_peak, fp64_work = None, {}
    try:
        fp64_peak = float(json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json")))["fp64_fma_tflops"])
        fp64_work = json.load(open(os.path.join(ROOT, "profiles", "fp64_work.json")))
    except Exception:
        pass
    work_key = f"{args.scheme}_p{args.degree}"
    flops = fp64_work.get(work_key, {})
    face_tflops = cell_tflops = None
    if fp64_peak and flops and ms_face_stage > 0 and ms_cell_stage > 0:
        face_tflops = flops["face_flop_per_point"] * n_face_pts / (ms_face_stage * 1e-3) / 1e12
        cell_tflops = flops["cell_flop_per_dof"] * dof_local / (ms_cell_stage * 1e-3) / 1e12
    kernels = {
        "face": {"ms": ms_face_stage, "hbm_gbs": face_bytes / max(ms_face_stage, 1e-9) / 1e6,
                 "hbm_frac": face_bytes / max(ms_face_stage, 1e-9) / 1e6 / peak,
                 "fp64_tflops": face_tflops, "fp64_frac": face_tflops / fp64_peak if face_tflops else None},
        "cell": {"ms": ms_cell_stage, "hbm_gbs": cell_bytes / max(ms_cell_stage, 1e-9) / 1e6,
                 "hbm_frac": cell_bytes / max(ms_cell_stage, 1e-9) / 1e6 / peak,
                 "fp64_tflops": cell_tflops, "fp64_frac": cell_tflops / fp64_peak if cell_tflops else None},
    }
    dom = kernels["face" if dominant == "face_flux_kernel" else "cell"]
    # the bound of the dominant kernel is the roofline it sits closer to
    fp64_bound = dom["fp64_frac"] is not None and dom["fp64_frac"] > dom["hbm_frac"]
    roofline = {"bound": "fp64" if fp64_bound else "hbm", "kernel": dominant,
                "achieved": dom["fp64_tflops"] if fp64_bound else achieved,
                "peak": fp64_peak if fp64_bound else peak, "unit": "TFLOP/s" if fp64_bound else "GB/s",
                "frac": dom["fp64_frac"] if fp64_bound else achieved / peak, "traffic": None,
                "peak_source": ("profiles/fp64_peak.json (DFMA microbenchmark on this pool's B200)" if fp64_bound else peak_src),
                "hbm_frac": achieved / peak, "fp64_frac": dom["fp64_frac"],
                "algorithmic_bytes_per_launch": dom_bytes, "ms_per_launch": dom_ms,
                "ms_face_per_stage": ms_face_stage, "ms_cell_per_stage": ms_cell_stage, "kernels": kernels,
                "stage_frac_of_hbm_roofline": BYTES_PER_UPDATE_RK3 * value / world / (peak * 1e9)}
    if args.scheme == "modal":
        roofline["kernel"] = "modal_face_kernel" if dominant == "face_flux_kernel" else "modal_cell_kernel"
    elif (dominant == "cell_update_kernel" and args.scheme == "fr" and args.degree == 2
          and os.environ.get("MCFD_CELL_LINES", "1") != "0"):
        roofline["kernel"] = "cell_update_line_kernel"   # LaunchCellUpdate's choice for FR p = 2 (cell_kernels.cu)
    # ncu DRAM bytes per launch were captured for the headline workload only (profiles/traffic.json says which)
    headline = args.n == 64 and args.degree == 2 and args.scheme == "fr" and world == 1 and args.limiter == "none"
    try:
        if headline:
            roofline["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(roofline["kernel"])
    except Exception:
        pass

    # ---- end to end through the C-ABI with HOST buffers: per step H2D of the column (pinned), one
    # RungeKutta update, D2H of the new column (pinned) ----
    host = torch.empty(dof_local, dtype=torch.float64).pin_memory()
    col = host.numpy()
    solver.get_solution(col)
    e2e_steps = max(1, min(args.steps, 5))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        solver.set_solution(col)
        solver.rk_update(t, dt, 1); t += dt
        solver.get_solution(col)
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = dof_total * 3 * e2e_steps / float(e2e_s.item())
    # The documented fast path: the column crosses PCIe once per FRAME, not once per step --
    # mcfd_set_solution, mcfd_rk_update(frame_steps), mcfd_get_solution.  Reported beside `e2e`, which pays the
    # reference's by-value Column interface every step and is PCIe-bound by construction.
    frame_steps = 10
    barrier()
    t0 = time.perf_counter()
    solver.set_solution(col)
    solver.rk_update(t, dt, frame_steps); t += dt * frame_steps
    solver.get_solution(col)
    barrier()
    frame_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(frame_s, op=dist.ReduceOp.MAX)
    e2e_frame_value = dof_total * 3 * frame_steps / float(frame_s.item())
    finite = bool(np.isfinite(col).all())
    troubled = int(solver.troubled_flags().sum()) if args.limiter_id else None

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sim, dof, threads, kind, desc, _keep, _cells = reference_sample(host_cores(), args.degree, args.scheme_id, args)
        sim.rk_update(0.0, dt, 1)
        n_cpu = 6
        wall = sim.rk_update(dt, dt, n_cpu) if kind == "reference" else None
        if wall is None:
            t0 = time.perf_counter(); sim.rk_update(dt, dt, n_cpu); wall = time.perf_counter() - t0
        cpu_baseline = {"value": dof * 3 * n_cpu / wall, "unit": "DOF-updates/s", "cores": threads, "kind": kind,
                        "sample": desc + f", {n_cpu} RK3 steps", "cpu": cpu_model()}
        if threads > 1:   # P = 1 beside P = all cores
            sim1, dof1, _t1, kind1, desc1, _keep1, _cells1 = reference_sample(1, args.degree, args.scheme_id, args)
            sim1.rk_update(0.0, dt, 1)
            t0 = time.perf_counter(); sim1.rk_update(dt, dt, 2); wall1 = time.perf_counter() - t0
            cpu_baseline["p1"] = {"value": dof1 * 3 * 2 / wall1, "cores": 1, "kind": kind1, "sample": desc1 + ", 2 RK3 steps"}

    if rank == 0:
        line = {
            "metric": "DOF-RK-stage updates/sec", "value": value, "unit": "DOF-updates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "repeat_ms_per_step": repeats, "median_ms_per_step": float(np.median(repeats)),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": dict(workload_config(args), halo=["none", "nccl send/recv", "peer-memory stores + arrival counters"][solver.halo_transport()]),
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": "DOF-updates/s", "h2d_bytes_per_step": dof_local * 8,
                    "d2h_bytes_per_step": dof_local * 8, "steps": e2e_steps,
                    "what": "per step: mcfd_set_solution(pinned host column) + mcfd_rk_update(1 step) + mcfd_get_solution"},
            "e2e_fast_path": {"value": e2e_frame_value, "unit": "DOF-updates/s", "steps_per_frame": frame_steps,
                              "h2d_bytes_per_frame": dof_local * 8, "d2h_bytes_per_frame": dof_local * 8,
                              "what": "per frame: mcfd_set_solution + mcfd_rk_update(10 steps) + mcfd_get_solution (INTEGRATION.md 2)"},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "solution_finite": finite,
        }
        if troubled is not None:
            line["troubled_cells_last_pass"] = troubled
        print(json.dumps(line))
    if world > 1:
        solver.detach_nccl()   # collective: peers close their mappings of this rank's columns before they are freed
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--cells", dest="n", type=int, default=64, help="cells per box edge (use --cells under torchrun: it reads --n as its own)")
    ap.add_argument("--degree", type=int, default=2)
    ap.add_argument("--scheme", default="fr", choices=["fr", "dgsem", "modal"])
    ap.add_argument("--limiter", default="none", choices=["none", "lazy", "eigen"], help="modal scheme only")
    ap.add_argument("--ic", default="vortex", choices=["vortex", "sod"])
    ap.add_argument("--size", type=float, default=10.0, help="box edge length")
    ap.add_argument("--workload", default="box", choices=["box", "double_mach", "forward_step"],
                    help="double_mach / forward_step: BASELINE configs[2] / configs[3] (use with --scheme modal --limiter eigen; "
                         "--cells = cells across the height)")
    ap.add_argument("--dt", type=float, default=1e-3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--what", default="step", choices=["step", "output"], help="output: the solution-output kernels of row f2")
    ap.add_argument("--flags", type=int, default=0, help="mcfd_config.flags (tuning experiments)")
    ap.add_argument("--lib", default=None, help="alternative build of libminicfd_b200.so (tuning experiments)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    args.scheme_id = {"fr": 2, "dgsem": 1, "modal": 0}[args.scheme]
    args.limiter_id = {"none": 0, "lazy": 1, "eigen": 2}[args.limiter]
    if args.what == "output":
        run_output(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
