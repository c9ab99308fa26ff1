import sys, os
import numpy as np
sys.path.insert(0, "/root/repo")
from minicfd_b200 import abi
from tests import common
abi.load()
for scheme, degree, cells in [(2, 1, (16, 32, 64)), (2, 2, (12, 24, 48)), (2, 3, (8, 16, 32)), (0, 1, (16, 32, 64)), (0, 3, (8, 16, 32))]:
    errs = []
    for n in cells:
        h = 10.0 / n
        mesh = abi.BoxMesh((n, n, 2), (0, 0, 0), (10, 10, 2 * h))
        gpu = abi.Solver(mesh.view, scheme=scheme, degree=degree, rk_order=3)
        common.set_vortex_case(gpu); gpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
        steps = 8 * n // cells[0] * 5
        gpu.rk_update(0.0, 0.2 / steps, steps)
        errs.append(gpu.measure_l1_error(abi.FUNC_VORTEX, common.VORTEX, 0.2) / (2 * h))
    e = np.array(errs)[:, 0]
    print(scheme, degree, cells, e, np.log2(e[:-1] / e[1:]), flush=True)
