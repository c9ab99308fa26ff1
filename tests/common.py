"""Shared helpers for the parity tests (test infrastructure)."""
import numpy as np

from minicfd_b200 import abi

VORTEX = [5.0, 5.0, 5.0, 1.0, 1.0, 1.4]  # beta, xc, yc, u0, v0, gamma
GAMMA = 1.4


def prim_to_cons(rho, u, v, w, p, gamma=GAMMA):
    return [rho, rho * u, rho * v, rho * w, p / (gamma - 1) + 0.5 * rho * (u * u + v * v + w * w)]


SOD_LEFT = prim_to_cons(1.0, 0.0, 0.0, 0.0, 1.0)
SOD_RIGHT = prim_to_cons(0.125, 0.0, 0.0, 0.0, 0.1)


def plane_split(normal, offset, speed, left, right):
    return list(normal) + [offset, speed] + list(left) + list(right)


def rel_err(a, b):
    """max |a-b| / max |b|  (the per-column relative measure used for the 1e-10 bar)"""
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def set_vortex_case(sim, wall_groups=(4, 5)):
    """Isentropic vortex: smart boundaries with the exact state on x/y faces, walls on z faces."""
    for g in range(6):
        if g in wall_groups:
            sim.set_boundary(g, abi.BC_SOLID_WALL)
        else:
            sim.set_boundary(g, abi.BC_SMART, abi.FUNC_VORTEX, VORTEX)


def make_parts(n, lo, hi, parts, distortion=0.0):
    world = parts[0] * parts[1] * parts[2]
    return [abi.BoxMesh(n, lo, hi, parts=parts, rank=r, distortion=distortion) for r in range(world)]


def perturbed(column, seed=20241104, amp=1e-3):
    rng = np.random.default_rng(seed)
    return column * (1.0 + amp * rng.uniform(-1, 1, column.shape))
