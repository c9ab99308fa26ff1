"""Shape classes of a modal part (HostTables::n_classes, minicfd_b200/csrc/setup.cpp BuildShapeClasses) and the
per-face / per-point decision for the face frames, from the host tables alone (mcfd_mesh_shape_classes: no GPU).

The class search decides which cell kernel a modal run takes (the tensor-core kernel on repetitive meshes), the frame
decision how much a face kernel reads per point; both are comparisons of coordinates that differ by rounding, and both
went wrong once on meshes whose spacing is small against their distance from the origin (a rounding-based key split a
1.4 M-cell double-Mach mesh into thousands of classes; flat faces were taken for curved at a fixed 1e-14).
"""
import pytest

from minicfd_b200 import abi


def test_box_far_from_the_origin_is_one_shape_with_flat_faces(lib):
    # h = 0.01 at x ~ 1000: edge vectors carry ~1e-11 of relative rounding
    m = abi.BoxMesh((40, 10, 4), (1000.0, 500.0, 0.0), (1000.4, 500.1, 0.04))
    assert abi.mesh_shape_classes(m.view) == (1, 27, False)       # 27 = on which sides the boundary lies (3 x 3 x 3)
    assert abi.mesh_shape_classes(m.view, scheme=abi.SCHEME_FR) == (0, 0, False)   # nodal schemes have no classes


def test_double_mach_sized_spacing(lib):
    # the mesh of bench.py --workload double_mach at h = 1/60 (172 800 cells): 4 / 240 is not a binary fraction
    m = abi.BoxMesh((240, 60, 12), (0, 0, 0), (4.0, 1.0, 0.2))
    assert abi.mesh_shape_classes(m.view) == (1, 27, False)


@pytest.mark.parametrize("degree", [1, 2, 3])
def test_kuhn_tetrahedra_have_six_shapes(lib, degree):
    m = abi.BoxMesh((5, 4, 3), (0, 0, 0), (1.0, 0.8, 0.6), tetra=True)
    shapes, classes, per_point = abi.mesh_shape_classes(m.view, degree=degree)
    assert shapes == 6 and 6 <= classes <= 256 and not per_point


def test_any_distortion_means_no_classes(lib):
    assert abi.mesh_shape_classes(abi.BoxMesh((6, 5, 4), distortion=0.1).view) == (0, 0, True)
    assert abi.mesh_shape_classes(abi.BoxMesh((6, 5, 4), distortion=1e-6).view) == (0, 0, True)    # curved faces, however slightly
    assert abi.mesh_shape_classes(abi.BoxMesh((5, 4, 3), distortion=0.05, tetra=True).view)[:2] == (0, 0)  # flat triangles, no repetition
    assert abi.mesh_shape_classes(abi.BoxMesh((6, 5, 4)).view, flags=abi.FLAG_NO_SHAPE_CLASSES) == (0, 0, False)
    # a displacement of 1e-10 h moves phi(x_q) and w grad phi(x_q) by as much: sharing one cell's tables would leave the
    # 1e-10 parity bar, so such a mesh must fall back to the general kernels (the match is at 1e-12 h, not 1e-9 h)
    assert abi.mesh_shape_classes(abi.BoxMesh((5, 4, 3), distortion=1e-10, tetra=True).view)[:2] == (0, 0)


def test_parts_of_a_partitioned_box_keep_their_classes(lib):
    for r in range(4):
        m = abi.BoxMesh((6, 6, 4), (0, 0, 0), (1.2, 1.2, 0.8), parts=(2, 2, 1), rank=r)
        shapes, classes, per_point = abi.mesh_shape_classes(m.view)
        assert shapes == 1 and 8 <= classes <= 64 and not per_point   # ghost faces are held by the local cell: more arrangements
