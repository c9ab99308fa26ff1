"""prints the parity figures of a few sensitive cases for a given build (diagnostic): python scripts/parity_probe.py [lib]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from minicfd_b200 import abi
from oracle import oracle
from tests import common, test_configs as tc
abi.load(sys.argv[1] if len(sys.argv) > 1 else None)
def report(name, a, b):
    pc, pd = common.parity(a, b)
    print(f"{name}: strong {common.strong_err(a, b):.2e} per-comp {pc:.2e} column {common.col_err(a, b):.2e}", flush=True)
mesh = abi.BoxMesh(tc.TUBE_N, (0, 0, 0), tc.TUBE_HI)
for steps in (5, 20):
    gpu = abi.Solver(mesh.view, scheme=abi.SCHEME_FR, degree=2, rk_order=3)
    cpu = oracle.OracleSim([mesh.view], abi.SCHEME_FR, 2, 3)
    for sim in (gpu, cpu):
        tc.sod_tube_case(sim); sim.rk_update(0.0, 1e-3, steps)
    report(f"sod tube FR p2 {steps} steps", gpu.get_solution(), cpu.get_solution(0))
m2 = abi.BoxMesh((12, 10, 8), (0, 0, 0), (10, 10, 10), distortion=0.1)
gpu = abi.Solver(m2.view, scheme=abi.SCHEME_FR, degree=2, rk_order=3); cpu = oracle.OracleSim([m2.view], abi.SCHEME_FR, 2, 3)
for sim in (gpu, cpu):
    common.set_vortex_case(sim); sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0); sim.rk_update(0.0, 5e-3, 10)
report("vortex distorted FR p2 10 steps", gpu.get_solution(), cpu.get_solution(0))
