#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the REFERENCE's own templates (oracle/_ref/libref_*.so, built
from /root/reference by oracle/Makefile) -- run in the dev container, commit the outputs.  The fixtures
travel with the repo, so the oracle and the CUDA path stay pinned on machines without /root/reference.

  python tests/golden/make_golden.py            (everything)
  python tests/golden/make_golden.py modal      (only the modal DG + limiter fixtures dg_*.npz)
  python tests/golden/make_golden.py tetra      (only the tetrahedral fixtures dgtet_*.npz)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from minicfd_b200 import abi          # noqa: E402
from oracle import refharness         # noqa: E402
from tests import common              # noqa: E402

CASES = [  # name, scheme, degree, rk_order, cells, parts, distortion, steps, dt
    ("fr_p2", abi.SCHEME_FR, 2, 3, (3, 2, 2), (1, 1, 1), 0.1, 3, 5e-3),
    ("fr_p1", abi.SCHEME_FR, 1, 2, (3, 3, 2), (1, 1, 1), 0.1, 3, 5e-3),
    ("fr_p3", abi.SCHEME_FR, 3, 3, (2, 2, 2), (1, 1, 1), 0.1, 2, 5e-3),
    ("dgsem_p2", abi.SCHEME_DGSEM, 2, 3, (3, 2, 2), (1, 1, 1), 0.1, 3, 5e-3),
    ("dgsem_p3", abi.SCHEME_DGSEM, 3, 2, (2, 2, 2), (1, 1, 1), 0.0, 2, 5e-3),
    ("fr_p2_2parts", abi.SCHEME_FR, 2, 3, (4, 2, 2), (2, 1, 1), 0.1, 3, 5e-3),
]


# modal DG (spatial::dg::General / WithLimiterAndSource): name, degree, limiter, rk_order, cells, parts, hi, steps, dt, ic
MODAL_CASES = [
    ("dg_p2", 2, 0, 3, (3, 2, 2), (1, 1, 1), (10.0, 9.0, 8.0), 3, 5e-3, "vortex"),
    ("dg_p1", 1, 0, 2, (3, 3, 2), (1, 1, 1), (10.0, 9.0, 8.0), 3, 5e-3, "vortex"),
    ("dg_p3", 3, 0, 3, (2, 2, 2), (1, 1, 1), (10.0, 9.0, 8.0), 2, 5e-3, "vortex"),
    ("dg_lazy_p2", 2, 1, 3, (4, 3, 3), (1, 1, 1), (0.8, 0.6, 0.6), 2, 2e-3, "sod"),
    ("dg_eigen_p2", 2, 2, 3, (6, 5, 4), (1, 1, 1), (1.2, 1.0, 0.8), 3, 2e-3, "sod"),
    ("dg_eigen_p1", 1, 2, 2, (6, 5, 4), (1, 1, 1), (1.2, 1.0, 0.8), 3, 2e-3, "sod"),
    ("dg_eigen_p3", 3, 2, 3, (5, 4, 3), (1, 1, 1), (1.0, 0.8, 0.6), 2, 2e-3, "sod"),
    ("dg_eigen_p2_2parts", 2, 2, 3, (6, 4, 4), (2, 1, 1), (1.2, 0.8, 0.8), 3, 2e-3, "sod"),
]


def set_modal_case(sim, ic, hi):
    """vortex: as the nodal cases; sod: an oblique Sod discontinuity through the box centre, walls all round."""
    if ic == "vortex":
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    else:
        for g in range(6):
            sim.set_boundary(g, abi.BC_SOLID_WALL)
        split = common.plane_split((1.0, 0.3, 0.2), 0.5 * hi[0], 0.0, common.SOD_LEFT, common.SOD_RIGHT)
        sim.approximate(abi.FUNC_PLANE_SPLIT, split, 0.0)


def modal_cases():
    for name, degree, limiter, rk, cells, parts, hi, steps, dt, ic in MODAL_CASES:
        meshes = common.make_parts(cells, (0, 0, 0), hi, parts, 0.1)
        ref = refharness.RefSim([m.view for m in meshes], abi.SCHEME_DG_MODAL, degree, limiter, rk)
        set_modal_case(ref, ic, hi)
        u0 = [common.perturbed(ref.get_solution(r), seed=r + 1, amp=1e-4) for r in range(len(meshes))]
        ref.set_solution(u0)                      # SetSolutionColumn: runs the limiter (dg/general.hpp:247-250)
        n_first = ref.troubled_count()
        ul = [ref.get_solution(r) for r in range(len(meshes))]
        res = ref.get_residual(0.25)
        ref.rk_update(0.25, dt, steps)
        u1 = [ref.get_solution(r) for r in range(len(meshes))]
        data = dict(scheme=abi.SCHEME_DG_MODAL, degree=degree, limiter=limiter, rk_order=rk, cells=np.array(cells),
                    parts=np.array(parts), lo=np.zeros(3), hi=np.array(hi), distortion=0.1, steps=steps, dt=dt, t0=0.25,
                    ic=ic, troubled_first=n_first, troubled_total=ref.troubled_count())
        for r in range(len(meshes)):
            data[f"u0_{r}"] = u0[r]; data[f"limited_{r}"] = ul[r]; data[f"residual_{r}"] = res[r]; data[f"u1_{r}"] = u1[r]
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **data)
        print(name, sum(u.size for u in u0), "DOF, troubled", n_first, ref.troubled_count())


# tetrahedral parts (oracle/_ref/libref_dgtet_*: Tetrahedron4 + gauss::Tetrahedron / Triangle, part.hpp:437-452):
# name, degree, limiter, rk_order, hexahedra (x 6 tetrahedra), parts, hi, steps, dt; an oblique Sod plane, walls all round
# (a 10 % jump where there is no limiter, as in tests/test_modal.py)
TETRA_CASES = [
    ("dgtet_p2", 2, 0, 3, (3, 2, 2), (1, 1, 1), (1.2, 1.0, 0.8), 3, 1e-3),
    ("dgtet_p1", 1, 0, 2, (3, 3, 2), (1, 1, 1), (1.2, 1.0, 0.8), 3, 1e-3),
    ("dgtet_p3", 3, 0, 3, (2, 2, 2), (1, 1, 1), (1.2, 1.0, 0.8), 2, 1e-3),
    ("dgtet_lazy_p2", 2, 1, 3, (3, 2, 2), (1, 1, 1), (0.8, 0.6, 0.6), 2, 1e-3),
    ("dgtet_eigen_p2", 2, 2, 3, (4, 3, 3), (1, 1, 1), (1.2, 1.0, 0.8), 3, 1e-3),
    ("dgtet_eigen_p3", 3, 2, 3, (3, 3, 2), (1, 1, 1), (1.0, 0.8, 0.6), 2, 1e-3),
    ("dgtet_eigen_p2_2parts", 2, 2, 3, (4, 3, 2), (2, 1, 1), (1.2, 0.8, 0.8), 3, 1e-3),
]


def tetra_cases():
    for name, degree, limiter, rk, cells, parts, hi, steps, dt in TETRA_CASES:
        world = parts[0] * parts[1] * parts[2]
        meshes = [abi.BoxMesh(cells, (0, 0, 0), hi, parts=parts, rank=r, distortion=0.1, tetra=True) for r in range(world)]
        ref = refharness.RefSim([m.view for m in meshes], abi.SCHEME_DG_MODAL, degree, limiter, rk)
        for g in range(6):
            ref.set_boundary(g, abi.BC_SOLID_WALL)
        right = common.SOD_RIGHT if limiter else common.prim_to_cons(0.9, 0.0, 0.0, 0.0, 0.9)
        ref.approximate(abi.FUNC_PLANE_SPLIT, common.plane_split((1.0, 0.3, 0.2), 0.5 * hi[0], 0.0, common.SOD_LEFT, right), 0.0)
        u0 = [common.perturbed(ref.get_solution(r), seed=r + 11, amp=1e-4) for r in range(world)]
        ref.set_solution(u0)
        n_first = ref.troubled_count()
        ul = [ref.get_solution(r) for r in range(world)]
        res = ref.get_residual(0.25)
        ref.rk_update(0.25, dt, steps)
        u1 = [ref.get_solution(r) for r in range(world)]
        data = dict(scheme=abi.SCHEME_DG_MODAL, degree=degree, limiter=limiter, rk_order=rk, cells=np.array(cells),
                    parts=np.array(parts), lo=np.zeros(3), hi=np.array(hi), distortion=0.1, steps=steps, dt=dt, t0=0.25,
                    ic="sod", cell_type=1, troubled_first=n_first, troubled_total=ref.troubled_count())
        for r in range(world):
            data[f"u0_{r}"] = u0[r]; data[f"limited_{r}"] = ul[r]; data[f"residual_{r}"] = res[r]; data[f"u1_{r}"] = u1[r]
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **data)
        print(name, sum(u.size for u in u0), "DOF, troubled", n_first, ref.troubled_count())


def riemann_cases(rng, n):
    def states(lo, hi, vmax):
        rho = rng.uniform(lo, hi, n); p = rng.uniform(lo, hi, n); vel = rng.uniform(-vmax, vmax, (n, 3))
        return np.ascontiguousarray(np.column_stack([rho, rho[:, None] * vel, p / 0.4 + 0.5 * rho * (vel ** 2).sum(1)]))
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    q[:, :, 2] *= np.sign(np.linalg.det(q))[:, None]
    fr = np.ascontiguousarray(np.transpose(q, (0, 2, 1)).reshape(n, 9))
    ul, ur = states(0.05, 2.0, 1.0), states(0.05, 2.0, 1.0)
    push = np.where(rng.uniform(size=n) < 0.5, 0.0, rng.uniform(-6.0, 6.0, n))  # half moderate, half strong
    for u, sgn in ((ul, -1.0), (ur, +1.0)):
        rho = u[:, 0].copy(); v = u[:, 1:4] / rho[:, None]
        p = 0.4 * (u[:, 4] - 0.5 * rho * (v ** 2).sum(1))
        v = v + sgn * push[:, None] * fr[:, 0:3]
        u[:, 1:4] = rho[:, None] * v
        u[:, 4] = p / 0.4 + 0.5 * rho * (v ** 2).sum(1)
    return fr, ul, ur


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "tetra":
        tetra_cases()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "modal":
        modal_cases()
        return
    modal_cases()
    rng = np.random.default_rng(20241104)
    fr, ul, ur = riemann_cases(rng, 600)
    out = dict(frames=fr, u_left=ul, u_right=ur, flux=refharness.riemann_flux(fr, ul, ur))
    for kind in range(6):
        out[f"bflux{kind}"] = refharness.boundary_flux(kind, fr, ul, ur)
    np.savez_compressed(os.path.join(HERE, "riemann.npz"), **out)
    for name, scheme, degree, rk, cells, parts, dist, steps, dt in CASES:
        meshes = common.make_parts(cells, (0, 0, 0), (10, 9, 8), parts, dist)
        ref = refharness.RefSim([m.view for m in meshes], scheme, degree, 0, rk)
        common.set_vortex_case(ref)
        ref.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
        u0 = [common.perturbed(ref.get_solution(r), seed=r + 1) for r in range(len(meshes))]
        ref.set_solution(u0)
        res = ref.get_residual(0.25)
        ref.rk_update(0.25, dt, steps)
        u1 = [ref.get_solution(r) for r in range(len(meshes))]
        data = dict(scheme=scheme, degree=degree, rk_order=rk, cells=np.array(cells), parts=np.array(parts),
                    lo=np.zeros(3), hi=np.array([10.0, 9.0, 8.0]), distortion=dist, steps=steps, dt=dt, t0=0.25)
        for r in range(len(meshes)):
            data[f"u0_{r}"] = u0[r]; data[f"residual_{r}"] = res[r]; data[f"u1_{r}"] = u1[r]
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **data)
        print(name, sum(u.size for u in u0), "DOF")


if __name__ == "__main__":
    main()
