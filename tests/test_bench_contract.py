"""bench.py's contract, as far as it can be checked without a GPU: the reference arm (the reference's own templates
on the host cores, oracle/_ref) prints one JSON line with the keys the driver reads, and the own arm refuses to run
without a CUDA device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_reference_arm_prints_the_contract_line():
    from oracle import refharness
    if not refharness.available(2, 2):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cells", "32"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "DOF-RK-stage updates/sec" and line["unit"] == "DOF-updates/s"
    assert line["value"] > 0 and line["higher_is_better"] is True and line["steps"] == 1 and line["n_gpus"] == 1
    assert line["config"]["workload"].startswith("Euler FR p=2 RK3 exact-Riemann, synthetic 32^3")
    assert line["config"]["cells"] == 32 ** 3 and line["config"]["dof"] == 32 ** 3 * 135   # what ran is what is printed
    cb = line["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == line["value"] and "the whole 32^3 mesh" in cb["sample"]
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_own_arm_needs_a_gpu():
    if _has_gpu():
        pytest.skip("a CUDA device is present")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"], capture_output=True, text=True,
                         timeout=600, cwd=ROOT)
    assert out.returncode != 0
    assert "no CPU fallback" in (out.stderr + out.stdout)
