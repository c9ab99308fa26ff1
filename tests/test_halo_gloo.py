"""World-size-2 test of the N > 1 host path on CPU (`gloo`): each rank builds ITS part of a box mesh,
packs the coefficient blocks of its inter cells in send-list order (Part::ShareGhostCellCoeffs,
mesh/part.hpp:1258-1284), exchanges them with its neighbours and unpacks them behind its local cells
(UpdateGhostCellCoeffs, part.hpp:1285-1302) -- the same lists mcfd_attach_nccl drives with ncclSend/ncclRecv.
The payload encodes (global cell id, field), so ghost indexing is checked bit for bit.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from minicfd_b200 import abi

FIELDS = 135  # K * N for p = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, parts, out, tetra=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mesh = abi.BoxMesh(n, parts=parts, rank=rank, distortion=0.1, tetra=tetra)   # tetra: six TETRA_4 per hexahedron (row f4)
        v, a, gid = mesh.view, mesh.arrays(), mesh.cell_global_ids()
        nl, ng = v.n_local_cells, v.n_ghost_cells
        column = np.zeros((nl + ng, FIELDS))
        column[:nl] = gid[:nl, None] * 1000.0 + np.arange(FIELDS)[None, :]
        reqs, recv_bufs = [], []
        for i, peer in enumerate(a["nbr_rank"]):
            cells = a["send_cells"][a["send_ptr"][i]:a["send_ptr"][i + 1]]
            send = torch.from_numpy(np.ascontiguousarray(column[cells]))   # pack
            recv = torch.empty((int(a["recv_ptr"][i + 1] - a["recv_ptr"][i]), FIELDS), dtype=torch.float64)
            recv_bufs.append(recv)
            reqs.append(dist.isend(send, int(peer)))
            reqs.append(dist.irecv(recv, int(peer)))
        for r in reqs:
            r.wait()
        for i in range(len(recv_bufs)):                                    # unpack behind the local cells
            column[nl + a["recv_ptr"][i]:nl + a["recv_ptr"][i + 1]] = recv_bufs[i].numpy()
        expect = gid[:, None] * 1000.0 + np.arange(FIELDS)[None, :]
        ok = bool(np.array_equal(column, expect))
        flag = torch.tensor([1 if ok else 0])
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            out.put(int(flag.item()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,parts,tetra", [((6, 4, 4), (2, 1, 1), False), ((4, 6, 5), (1, 2, 1), False), ((4, 3, 2), (2, 1, 1), True)])
def test_two_rank_halo_exchange(lib, n, parts, tetra):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, parts, out, tetra)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) == 1
