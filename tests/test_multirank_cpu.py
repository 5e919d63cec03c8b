"""CPU, world_size 2 over gloo: the host-side logic of the multi-GPU path -- rendezvous plumbing (NCCL id
broadcast), the i-body partition, and that "each rank advances its slice from the shared snapshot, then
all-gather" is exactly the synchronous step.  The per-rank force rows come from the oracle here (test
infrastructure); on GPUs the same partition drives libgravsim (tests/test_gpu_parity.py::test_two_gpus_when_present)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from conftest import G_SI, Oracle, load_package


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, steps, out_dir):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    gs = load_package()
    from gravsim_b200 import dist as gdist
    assert gdist.env_ranks() == (rank, world, rank)
    # 1. the NCCL-id courier: 128 opaque bytes from rank 0 reach every rank unchanged
    secret = bytes((7 * i + 3) % 256 for i in range(128))
    got = gdist.broadcast_bytes(secret if rank == 0 else None, 128)
    assert got == secret
    # 2. partition: contiguous, disjoint, covering, tile-aligned in padded space
    ranges = [gdist.owned_bodies(gs.plan_slice, n, world, r) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    b0, b1 = ranges[rank]
    # 3. synchronous step by slices + all-gather == single-rank synchronous step (bit-exact in fp64)
    oracle = Oracle()
    state = gs.ic_reference_uniform(n, seed=77, dtype=np.float64)
    dt = 3600.0
    for _ in range(steps):
        idx = np.arange(b0, b1, dtype=np.int32)
        acc = oracle.accel_subset(state, idx) if idx.size else np.zeros((0, 3))
        mine = state[b0:b1].copy()
        mine[:, 3:6] = mine[:, 3:6] + acc * dt          # mpi.c:207-210
        mine[:, 0:3] = mine[:, 0:3] + mine[:, 3:6] * dt  # mpi.c:212-215
        state = gdist.gather_slices(mine, [e - b for b, e in ranges])   # mpi.c:227-231
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), state)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [3000, 5])
def test_two_ranks_partition_and_exchange(tmp_path, n):
    port = _free_port()
    steps = 3
    mp.spawn(_worker, args=(2, port, n, steps, str(tmp_path)), nprocs=2, join=True)
    gs = load_package()
    ref = Oracle().run_sync(gs.ic_reference_uniform(n, seed=77, dtype=np.float64), 3600.0, steps, G=G_SI)
    r0, r1 = np.load(tmp_path / "rank0.npy"), np.load(tmp_path / "rank1.npy")
    assert np.array_equal(r0, r1)            # every rank holds the full state, as after MPI_Allgatherv
    assert np.array_equal(r0, ref)           # and it is the synchronous step
