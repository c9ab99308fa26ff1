"""How far does the reference sit from ITSELF?  (parity item 1.)

Several GPU parity bounds are wider than north_star's 1e-10: 1e-8 on residuals of Sod data (tests/test_modal.py,
tests/test_tetra.py), 2e-9 on the solution after modal RK steps on Sod data, NOISE_TOL = 1e-8 on weak components
(tests/common.py), 2e-7 on the weak components of the 8-part limiter run (tests/test_multi_gpu.py).  The reason given
there -- the reference's exact Riemann solver takes the SIGN of f(p_L) on rounding noise and stops Newton at 1e-8
(riemann/euler/exact.hpp:163-174), so no two evaluation orders agree better -- was so far shown only between two
different programs (the oracle port and the compiled reference).  Here it is shown on the reference alone: the same
unmodified templates (oracle/ref/harness.cpp) compiled three ways by oracle/Makefile,
    ""        -O2                                   (the build every other test compares with)
    ".O0"     -O0                                   (no optimisation)
    ".O3fma"  -O3 -mfma -ffp-contract=fast          (fused multiply-adds: another rounding of the same expressions)
run on the inputs of those GPU tests.  Findings asserted below:
  * -O0 and -O2 agree bit for bit: the reference is deterministic, the differences are not a harness artefact;
  * on smooth data (isentropic vortex) the FMA build agrees with the -O2 build to 1e-11 after 4 steps (measured 2e-12): contraction alone is harmless;
  * on Sod data the FMA build differs from the -O2 build by 1e-9 .. 4e-8 of a component's own scale -- the
    reference misses north_star's 1e-10 against itself -- and every widened bound of the GPU tests is within 10x of
    that self-noise.
CPU only; skipped where oracle/_ref was not built (no /root/reference).
"""
import numpy as np
import pytest

from minicfd_b200 import abi
from oracle import refharness
from tests import common
from tests.test_modal import set_case, sod_split

K = common.K
HI = (1.2, 1.0, 0.8)

# the widened bounds of the GPU tests, by name (the tests themselves hold the literals; test_bounds_are_the_tests_own checks them)
BOUND_RESIDUAL_SOD = 1e-8        # tests/test_modal.py, tests/test_tetra.py: residual of Sod data
BOUND_SOLUTION_SOD = 2e-9        # same files: solution after 4 modal RK steps on Sod data
BOUND_WEAK_MULTI = 2e-7          # tests/test_multi_gpu.py: weak components, limiter run on up to 8 parts


def _needs(scheme, limiter, cell, *variants):
    for v in ("",) + variants:
        if not refharness.available(scheme, 2, limiter, cell, v):
            pytest.skip("oracle/_ref (with the .O0 / .O3fma variants) not built: needs /root/reference")


def _component_figures(a, b):
    """per component: (max|a_c - b_c| / max|b_c|, is it a strong component) -- the measure of common.assert_parity"""
    a = np.asarray(a).reshape(-1, K); b = np.asarray(b).reshape(-1, K)
    scales = np.abs(b).max(0)
    return [(float(np.abs(a[:, c] - b[:, c]).max() / scales[c]), bool(scales[c] >= 1e-2 * scales.max())) for c in range(K)]


def _run(variant, meshes, scheme, limiter, ic, split, steps, dt, hi):
    sim = refharness.RefSim([m.view for m in meshes], scheme, 2, limiter=limiter, rk_order=3, variant=variant)
    if ic == "vortex":
        common.set_vortex_case(sim)
        sim.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    else:
        set_case(sim, "sod", hi)
        sim.approximate(abi.FUNC_PLANE_SPLIT, split, 0.0)
    n = len(meshes)
    sim.set_solution([sim.get_solution(r) for r in range(n)])     # SetSolutionColumn: the first limiter pass
    residual = [np.array(r) for r in sim.get_residual(0.0)]
    sim.rk_update(0.0, dt, steps)
    return residual, [sim.get_solution(r) for r in range(n)]


def test_unoptimised_build_agrees_bit_for_bit(lib):
    _needs(abi.SCHEME_DG_MODAL, 0, 0, ".O0")
    meshes = [abi.BoxMesh((5, 4, 3), (0, 0, 0), HI, distortion=0.12)]
    args = (meshes, abi.SCHEME_DG_MODAL, 0, "sod", sod_split(HI, weak=True), 2, 2e-3, HI)
    base, plain = _run("", *args), _run(".O0", *args)
    assert np.array_equal(base[0][0], plain[0][0]) and np.array_equal(base[1][0], plain[1][0])


@pytest.mark.parametrize("scheme", [abi.SCHEME_FR, abi.SCHEME_DG_MODAL])
def test_smooth_data_fma_build_agrees_to_rounding(lib, scheme):
    _needs(scheme, 0, 0, ".O3fma")
    meshes = [abi.BoxMesh((6, 5, 4), (0, 0, 0), (10, 10, 10), distortion=0.1)]
    args = (meshes, scheme, 0, "vortex", None, 4, 5e-3, None)
    base, fma = _run("", *args), _run(".O3fma", *args)
    assert base[1][0].tobytes() != fma[1][0].tobytes()            # it IS another rounding
    # (z-momentum of the planar vortex is noise in both builds: the column-scale figure covers it)
    # (the residual of this nearly steady flow is itself a small difference of O(1) fluxes: its own scale amplifies
    #  rounding, so the statement is about the solution)
    print(f"vortex, scheme {scheme}: residual {common.col_err(fma[0][0], base[0][0]):.1e}, after 4 steps {common.col_err(fma[1][0], base[1][0]):.1e}")
    assert common.col_err(fma[1][0], base[1][0]) < 1e-11
    for err, strong in _component_figures(fma[1][0], base[1][0]):
        assert not strong or err < 1e-11


@pytest.mark.parametrize("tetra,limiter", [(False, 0), (False, 2), (True, 0), (True, 2)])
def test_sod_data_reference_against_itself(lib, tetra, limiter):
    """The inputs of test_gpu_modal_matches_oracle / tests/test_tetra.py.  Measured here (profiles/r03_noise_floor.txt):
    residual momenta 3e-10 .. 3e-8, momenta after 4 steps 2e-10 .. 3e-9, density and energy 1e-12 .. 2e-11."""
    _needs(abi.SCHEME_DG_MODAL, limiter, int(tetra), ".O3fma")
    meshes = [abi.BoxMesh((4, 3, 3) if tetra else (7, 5, 4), (0, 0, 0), HI, distortion=0.12, tetra=tetra)]
    args = (meshes, abi.SCHEME_DG_MODAL, limiter, "sod", sod_split(HI, weak=(limiter == 0)), 4, 2e-3, HI)
    base, fma = _run("", *args), _run(".O3fma", *args)
    res = _component_figures(fma[0][0], base[0][0])
    sol = _component_figures(fma[1][0], base[1][0])
    print(f"reference(-O3 fma) vs reference(-O2), tetra={tetra} limiter={limiter}: residual "
          + " ".join(f"{e:.1e}" for e, _ in res) + " | after 4 steps " + " ".join(f"{e:.1e}" for e, _ in sol))
    worst_res, worst_sol = max(e for e, _ in res), max(e for e, _ in sol)
    # the reference does not meet 1e-10 against itself on these data ...
    assert worst_res > 1e-10 and worst_sol > 1e-10
    # ... density and energy, which carry the solution, stay far inside it ...
    assert max(sol[0][0], sol[4][0]) < 1e-10
    # ... and the widened GPU bounds are not generous next to that self-noise
    assert BOUND_RESIDUAL_SOD <= 10 * worst_res or limiter != 0     # (a limiter pass smooths the noisiest faces)
    assert BOUND_SOLUTION_SOD <= 10 * worst_sol
    assert worst_res < 1e-6 and worst_sol < 1e-7                   # noise, not a different solution


def test_eight_part_limiter_run_reference_against_itself(lib):
    """The inputs of tests/test_multi_gpu.py::test_nccl_limiter_halo_matches_oracle on 8 parts: the weak components
    (transverse momenta) of the reference move by 4e-8 of their own scale between two compilations -- the figure the
    8-GPU run shows against the oracle (profiles/r03_multi)."""
    _needs(abi.SCHEME_DG_MODAL, 2, 0, ".O3fma")
    hi, parts = (1.2, 1.2, 0.8), (2, 2, 2)
    meshes = [abi.BoxMesh((6, 6, 4), (0, 0, 0), hi, parts=parts, rank=r, distortion=0.1) for r in range(8)]
    split = common.plane_split((1.0, 0.3, 0.2), 0.6, 0.0, common.SOD_LEFT, common.SOD_RIGHT)
    args = (meshes, abi.SCHEME_DG_MODAL, 2, "sod", split, 3, 2e-3, hi)
    base, fma = _run("", *args), _run(".O3fma", *args)
    weak, strong = 0.0, 0.0
    for r in range(8):
        for err, is_strong in _component_figures(fma[1][r], base[1][r]):
            if is_strong:
                strong = max(strong, err)
            else:
                weak = max(weak, err)
    print(f"reference(-O3 fma) vs reference(-O2), 8 parts + weno::Eigen: strong {strong:.1e}, weak {weak:.1e}")
    assert weak > common.NOISE_TOL / 10 and BOUND_WEAK_MULTI <= 10 * weak
    assert strong < 1e-9 and weak < 1e-6


def test_bounds_are_the_tests_own():
    """The literals above are the ones the GPU tests assert (so this file cannot drift from them)."""
    import inspect
    from tests import test_modal, test_multi_gpu, test_tetra
    modal, tetra, multi = (inspect.getsource(m) for m in (test_modal, test_tetra, test_multi_gpu))
    assert "cpu.get_solution(0), 2e-9)" in modal and "2e-9)" in tetra
    assert "1e-8" in modal and 'd["residual_0"], 1e-8)' in tetra
    assert "weak_tol=2e-7" in multi
