"""A small run of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck): nodal FR and DGSEM
stages (TMA bulk tiles + mbarrier in the cell kernels), modal DG with weno::Eigen on hexahedra (shape-class DMMA kernel
and the general one) and on tetrahedra, the output kernels and both L1-error entry points.  No torch import: the
sanitizer then instruments only this library.
    compute-sanitizer --tool racecheck python scripts/sanitize_case.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minicfd_b200 import abi   # noqa: E402
from tests import common       # noqa: E402

abi.load()
checks = []
for scheme, degree in ((abi.SCHEME_FR, 2), (abi.SCHEME_DGSEM, 2), (abi.SCHEME_FR, 3), (abi.SCHEME_FR, 1)):
    mesh = abi.BoxMesh((4, 4, 4), (0, 0, 0), (10, 10, 10), distortion=0.1)
    sim = abi.Solver(mesh.view, scheme=scheme, degree=degree, rk_order=3)
    common.set_vortex_case(sim)
    sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    sim.rk_update(0.0, 5e-3, 2)
    a = sim.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.01)
    b = sim.measure_l1_error_static(abi.FUNC_VORTEX, common.VORTEX, 0.01)
    checks.append(np.array_equal(a, b) and np.isfinite(sim.get_solution()).all())
    sim.close()
split = common.plane_split((1.0, 0.3, 0.2), 0.6, 0.0, common.SOD_LEFT, common.SOD_RIGHT)
for tetra, distortion, limiter in ((False, 0.0, abi.LIMITER_EIGEN), (False, 0.1, abi.LIMITER_EIGEN), (True, 0.0, abi.LIMITER_EIGEN),
                                   (False, 0.0, 0)):
    mesh = abi.BoxMesh((4, 4, 4), (0, 0, 0), (1.2, 1.2, 0.8), distortion=distortion, tetra=tetra)
    sim = abi.Solver(mesh.view, scheme=abi.SCHEME_DG_MODAL, degree=2, limiter=limiter, rk_order=3)
    for g in range(6):
        sim.set_boundary(g, abi.BC_SOLID_WALL)
    sim.approximate(abi.FUNC_PLANE_SPLIT, split, 0.0)
    sim.set_solution(sim.get_solution())
    sim.rk_update(0.0, 2e-3, 2)
    checks.append(bool(np.isfinite(sim.get_solution()).all()))
    sim.close()
print("sanitize_case:", "ok" if all(checks) else f"FAILED {checks}")
sys.exit(0 if all(checks) else 1)
