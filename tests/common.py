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


K = 5  # conservative components; every column is [cell][node or mode][K], K fastest (column layout)


def parity(a, b, floor=1e-12, zero_component=1e-9):
    """The BASELINE.md section 5 gate.  Returns (per_component, per_dof):
    per_component = max over the K components c of  max|a_c - b_c| / max|b_c|   (what the tests assert), where a
        component whose reference stays below `zero_component` x the column's scale everywhere (the transverse
        momenta of a planar flow: rounding noise, or the residue of the exact solver's 1e-8 stopping test) is measured
        against the column's scale instead of against noise;
    per_dof = the worst |a_i - b_i| / |b_i| over the entries with |b_i| > floor  (reported next to it)."""
    a = np.asarray(a, dtype=np.float64).reshape(-1, K)
    b = np.asarray(b, dtype=np.float64).reshape(-1, K)
    assert a.shape == b.shape
    if not (np.isfinite(a).all() and np.isfinite(b).all()):
        return float("inf"), float("inf")
    diff = np.abs(a - b)
    scale_all = max(float(np.abs(b).max()), 1e-300)
    per_comp = 0.0
    for c in range(K):
        scale = float(np.abs(b[:, c]).max())
        if scale <= zero_component * scale_all:
            scale = scale_all
        per_comp = max(per_comp, float(diff[:, c].max()) / scale)
    mask = np.abs(b) > floor
    per_dof = float((diff[mask] / np.abs(b[mask])).max()) if mask.any() else 0.0
    return per_comp, per_dof


_worst = {"per_component": 0.0, "per_dof": 0.0}

# The reference's exact solver starts FindRoot at p_L, halves while f > 0 and stops Newton at |f| < 1e-8
# (riemann/euler/exact.hpp:163-174).  Across a face whose two states agree to rounding, the SIGN of f(p_L) is rounding
# noise; when it comes out positive the returned p* is only the 1e-8 iterate, i.e. ~1e-9 off.  The same noise therefore
# reaches ~1e-12 absolute in every weak component (transverse momenta; x-momentum of a tube at early times), whatever the
# order of the floating-point sums: the plain-C oracle and the COMPILED REFERENCE differ by 1e-10 .. 8e-9 of such a
# component's own scale (tests/test_oracle_vs_reference.py, measured), while agreeing to 1e-12 on the strong ones.
# Tests whose data have such components assert NOISE_TOL per component and, beside it, the old column-scale figure.
NOISE_TOL = 1e-8


def strong_err(a, b, strong=1e-2):
    """per-component max|a_c - b_c| / max|b_c| over the components that carry the solution (scale at least `strong` x
    the column's): the 1e-10 bar of north_star applies to these without qualification"""
    a = np.asarray(a, dtype=np.float64).reshape(-1, K); b = np.asarray(b, dtype=np.float64).reshape(-1, K)
    scales = np.abs(b).max(0)
    worst = 0.0
    for c in range(K):
        if scales[c] >= strong * scales.max():
            worst = max(worst, float(np.abs(a[:, c] - b[:, c]).max() / scales[c]))
    return worst


_relaxed = []   # (test id, what, figure, bound): every assert_parity that passed only thanks to a bound above 1e-10


def assert_parity(a, b, tol=1e-10, what="", weak_tol=None, strong=1e-2, zero_component=1e-9):
    """The parity gate of the GPU tests, per component c with scale_c = max|b_c|:
      strong (scale_c >= 1e-2 x the column's scale): max|a_c - b_c| / scale_c < tol -- north_star's 1e-10;
      zero   (scale_c <= 1e-9 x the column's scale: the reference holds nothing there but rounding):
             max|a_c - b_c| / (column scale) < tol as well -- a spurious component of 1e-9 of the column must fail;
      weak   (in between: small but real, noise-level by the mechanism described at NOISE_TOL):
             max|a_c - b_c| / scale_c < NOISE_TOL (or weak_tol).
    Whatever passes with a figure above 1e-10 is recorded and listed in the session summary (tests/conftest.py); the
    per-DOF figure goes there too."""
    import os
    rel_err(a, b)   # session summary figures
    a = np.asarray(a, dtype=np.float64).reshape(-1, K); b = np.asarray(b, dtype=np.float64).reshape(-1, K)
    assert a.shape == b.shape and np.isfinite(a).all() and np.isfinite(b).all(), f"{what}: shape or non-finite values"
    scales = np.abs(b).max(0)
    column = max(float(scales.max()), 1e-300)
    weak_bound = max(tol, NOISE_TOL if weak_tol is None else weak_tol)
    for c in range(K):
        d = float(np.abs(a[:, c] - b[:, c]).max())
        if scales[c] >= strong * column:
            kind, err, bound = "strong", d / scales[c], tol
        elif scales[c] <= zero_component * column:
            kind, err, bound = "zero", d / column, tol
        else:
            kind, err, bound = "weak", d / scales[c], weak_bound
        assert err < bound, f"{what} {kind} component {c}: {err:.3e} >= {bound:g} (scale {scales[c]:.3e} of {column:.3e})"
        if err >= 1e-10:
            _relaxed.append((os.environ.get("PYTEST_CURRENT_TEST", "?").split(" ")[0], f"{what} {kind} component {c}".strip(), err, bound))


def col_err(a, b):
    """max|a-b| / max|b| over the whole column: a supplementary, tighter figure next to rel_err() where a weak
    component forces NOISE_TOL on the per-component gate"""
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(b).max(), 1e-300))


def rel_err(a, b):
    """Per-component relative error (see parity()); the per-DOF figure is kept for the session summary that
    tests/conftest.py prints.  Flat arrays whose size is not a multiple of K fall back to one component."""
    a = np.asarray(a); b = np.asarray(b)
    if a.size % K:
        return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
    pc, pd = parity(a, b)
    _worst["per_component"] = max(_worst["per_component"], pc if np.isfinite(pc) else 0.0)
    _worst["per_dof"] = max(_worst["per_dof"], pd if np.isfinite(pd) else 0.0)
    return pc


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


# ---- Exact::star_u is a MEMBER of the Riemann object (exact.hpp:64): a vacuum solve upwinds (v, w) by the contact speed
# of the last ordinary solve of the same face point ----------------------------------------------------------------
def cellwise_constant_column(mesh, scheme, degree, state_of_cell):
    """the column of a piecewise-constant field on a box of UNIT cubes: nodal schemes carry the state at every node
    (FR: times det J = 1/8), the modal scheme carries it in the first orthonormal mode (phi_0 = 1/sqrt(volume) = 1)"""
    a = mesh.arrays()
    n_local = mesh.view.n_local_cells
    centres = a["node_xyz"][a["cell_nodes"][:n_local]].mean(1)
    n = {1: 4, 2: 10, 3: 20}[degree] if scheme == abi.SCHEME_DG_MODAL else (degree + 1) ** 3
    col = np.zeros((n_local, n, K))
    for c in range(n_local):
        u = np.asarray(state_of_cell(centres[c]), dtype=np.float64)
        if scheme == abi.SCHEME_DG_MODAL:
            col[c, 0] = u
        else:
            col[c, :] = u * (0.125 if scheme == abi.SCHEME_FR else 1.0)
    return np.ascontiguousarray(col.reshape(-1))


def vacuum_sequence(sim, mesh, scheme, degree, set_solution, get_residual):
    """solve 1: a uniform stream with u > 0 (ordinary case everywhere: star_u = 0.5 on the x-faces); solve 2: the two
    halves of the box recede so fast across the plane x = 2 that the star region is a vacuum (f(0) >= 0,
    exact.hpp:137-156) while the t-axis lies inside the left expansion fan, whose density is positive: the flux then
    carries the passive (v, w) of the side PassiveScalars picks by the MEMBER star_u (exact.hpp:37-46, 64, 69-76).
    Returns (second residual, the column of the second solve)."""
    for g in range(6):
        sim.set_boundary(g, abi.BC_SOLID_WALL)
    calm = prim_to_cons(1.0, 0.5, 0.2, 0.1, 1.0)
    set_solution(cellwise_constant_column(mesh, scheme, degree, lambda x: calm))
    get_residual()
    left = prim_to_cons(1.0, 0.5, 0.9, -0.7, 0.4)
    right = prim_to_cons(1.0, 30.0, -0.8, 0.6, 0.4)
    col = cellwise_constant_column(mesh, scheme, degree, lambda x: left if x[0] < 2.0 else right)
    set_solution(col)
    return get_residual(), col


def transverse_momentum_err(a, b):
    """max difference in the (rho v, rho w) entries relative to their own largest entry"""
    a = np.asarray(a).reshape(-1, K)[:, 2:4]; b = np.asarray(b).reshape(-1, K)[:, 2:4]
    return float(np.abs(a - b).max() / np.abs(b).max())
