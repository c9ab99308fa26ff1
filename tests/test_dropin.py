"""The drop-in boundary, executed on the GPU box: oracle/_ref/dropin_test_* (oracle/ref/dropin_test.cpp) lets
the reference's own mini::temporal::RungeKutta<3>::Update (temporal/rk.hpp:46-73) drive
mini_b200::CudaSystem<mini::temporal::System<double>> (minicfd_b200/host/cuda_system.hpp, over the C-ABI)
through the reference's four System virtuals with std::function boundary callbacks, next to the
reference's own spatial operator on the same mesh, and compares the solution columns (1e-10 relative).
The binaries are built in the dev container (they contain instantiated reference templates) and travel
to the GPU box; without them the test is skipped.
"""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,args", [("fr_p2", ["6", "5", "4", "0.1", "5", "5e-3"]),
                                       ("dgsem_p2", ["5", "4", "4", "0.1", "5", "5e-3"]),
                                       ("fr_p3", ["4", "3", "3", "0.0", "3", "5e-3"])])
def test_reference_rk_drives_the_gpu_system(lib, name, args):
    exe = os.path.join(ROOT, "oracle", "_ref", "dropin_test_" + name)
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/dropin_test_* not built (needs /root/reference at build time)")
    out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300)
    print(out.stdout, out.stderr)
    assert out.returncode == 0, out.stdout + out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["dropin_rel_err"] < 1e-10 and line["fast_rel_err"] < 1e-10
