#!/usr/bin/env python
"""Generates tests/golden/vtu_*.npz: the files the REFERENCE's own mesh::vtk::Writer<Part>::WriteSolutions
(/root/reference/include/mini/mesh/vtk.hpp:385-487, compiled unmodified into oracle/_ref/libref_*.so by
oracle/Makefile) writes for a given solution column, together with that column and Section::GatherFields'
field-major arrays (mesh/part.hpp:370-378).  Run in the dev container, commit the outputs; the GPU tests
(tests/test_output.py) feed the same column to mcfd_write_vtu / mcfd_gather_fields and compare.

  python tests/golden/make_golden_vtu.py
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from minicfd_b200 import abi          # noqa: E402
from oracle import refharness         # noqa: E402
from tests import common              # noqa: E402

CASES = [  # name, scheme, degree, cells, parts, distortion
    ("vtu_fr_p2", abi.SCHEME_FR, 2, (3, 2, 2), (1, 1, 1), 0.1),
    ("vtu_fr_p1", abi.SCHEME_FR, 1, (3, 2, 2), (1, 1, 1), 0.1),
    ("vtu_fr_p3", abi.SCHEME_FR, 3, (2, 2, 2), (1, 1, 1), 0.1),
    ("vtu_dgsem_p2", abi.SCHEME_DGSEM, 2, (3, 2, 2), (1, 1, 1), 0.1),
    ("vtu_dg_p2", abi.SCHEME_DG_MODAL, 2, (3, 2, 2), (1, 1, 1), 0.1),
    ("vtu_dg_p3", abi.SCHEME_DG_MODAL, 3, (2, 2, 2), (1, 1, 1), 0.1),
    ("vtu_fr_p2_2parts", abi.SCHEME_FR, 2, (4, 2, 2), (2, 1, 1), 0.1),
]


def main():
    for name, scheme, degree, cells, parts, distortion in CASES:
        meshes = common.make_parts(cells, (0, 0, 0), (10.0, 9.0, 8.0), parts, distortion)
        ref = refharness.RefSim([m.view for m in meshes], scheme, degree, 0, 3)
        common.set_vortex_case(ref)
        ref.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
        u = [common.perturbed(ref.get_solution(r), seed=r + 7, amp=1e-3) for r in range(len(meshes))]
        ref.set_solution(u)
        data = dict(scheme=scheme, degree=degree, cells=np.array(cells), parts=np.array(parts), lo=np.zeros(3),
                    hi=np.array((10.0, 9.0, 8.0)), distortion=distortion)
        with tempfile.TemporaryDirectory() as tmp:
            ref.write_vtu(tmp, "Frame0")
            data["pvtu"] = np.frombuffer(open(os.path.join(tmp, "Frame0.pvtu"), "rb").read(), dtype=np.uint8)
            for r in range(len(meshes)):
                data[f"u_{r}"] = u[r]
                data[f"fields_{r}"] = ref.gather_fields(r)
                data[f"coords_{r}"], data[f"values_{r}"] = ref.vtk_sample(r)
                data[f"vtu_{r}"] = np.frombuffer(open(os.path.join(tmp, "Frame0", f"{r}.vtu"), "rb").read(), dtype=np.uint8)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **data)
        print(name, sum(x.size for x in u), "DOF,", sum(data[f"vtu_{r}"].size for r in range(len(meshes))), "bytes of vtu")


if __name__ == "__main__":
    main()
