"""N > 1 on real GPUs: one process per GPU, NCCL halo exchange inside mcfd_rk_update
(replaces Part::ShareGhostCellCoeffs / UpdateGhostCellCoeffs, mesh/part.hpp:1258-1302), compared with
the oracle run on the same parts.  Skipped (not failed) on a single-GPU box.
"""
import os
import socket

import numpy as np
import pytest
import torch

from tests import common

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, parts, steps, out):
    import torch.distributed as dist
    from minicfd_b200 import abi
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        mesh = abi.BoxMesh(n, (0, 0, 0), (10, 10, 10), parts=parts, rank=rank, distortion=0.1)
        solver = abi.Solver(mesh.view, scheme=abi.SCHEME_FR, degree=2, rk_order=3, device=rank, rank=rank, world=world)
        ids = [abi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        solver.attach_nccl(ids[0], rank, world)
        common.set_vortex_case(solver)
        solver.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
        solver.rk_update(0.0, 5e-3, steps)
        out.put((rank, solver.get_solution()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,parts", [(2, (2, 1, 1)), (4, (2, 2, 1)), (8, (2, 2, 2))])
def test_nccl_halo_matches_oracle(lib, world, parts):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    from minicfd_b200 import abi
    from oracle import oracle
    n, steps = (6, 6, 4), 4
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, parts, steps, out)) for r in range(world)]
    for p in procs:
        p.start()
    cols = dict(out.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    meshes = common.make_parts(n, (0, 0, 0), (10, 10, 10), parts, 0.1)
    cpu = oracle.OracleSim([m.view for m in meshes], abi.SCHEME_FR, 2, 3)
    common.set_vortex_case(cpu)
    cpu.approximate(abi.FUNC_VORTEX, common.VORTEX, 0.0)
    cpu.rk_update(0.0, 5e-3, steps)
    for r in range(world):
        assert common.rel_err(cols[r], cpu.get_solution(r)) < 1e-10
