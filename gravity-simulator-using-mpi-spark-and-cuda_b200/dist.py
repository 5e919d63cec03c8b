"""Rendezvous plumbing for one-process-per-GPU runs (torchrun): rank/world from the environment and a
byte broadcast for the NCCL id that gs_create() needs on every rank.  Replaces what `mpirun` +
MPI_Init/MPI_Bcast do for the reference (mpi.c:142-144,160).  torch.distributed is only the courier:
the per-step position exchange itself is NCCL inside libgravsim (gs_api.cu)."""
import os

import numpy as np


def env_ranks():
    """(rank, world_size, local_rank) as exported by torchrun; (0, 1, 0) when run directly."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def broadcast_bytes(payload, nbytes, src=0, device=None):
    """Every rank returns rank `src`'s `payload` (bytes of length nbytes).  Works on gloo (CPU) and nccl."""
    import torch
    import torch.distributed as dist
    buf = torch.zeros(nbytes, dtype=torch.uint8)
    if dist.get_rank() == src:
        if payload is None or len(payload) != nbytes:
            raise ValueError("source rank must provide exactly nbytes")
        buf.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    if device is not None:
        buf = buf.to(device)
    dist.broadcast(buf, src)
    return bytes(buf.cpu().numpy().tobytes())


def owned_bodies(plan_slice, n, world_size, rank, variant=0):
    """Real (unpadded) body range [begin, end) of `rank`: the block partition of mpi.c:184-187 on tile
    granularity.  `plan_slice` is gravsim_b200.plan_slice (gs_plan_slice)."""
    n_pad, b, e = plan_slice(n, world_size, rank, variant)
    return min(b, n), min(e, n)


def gather_slices(local_rows, counts):
    """All-gather of ragged row blocks (the MPI_Allgatherv of mpi.c:227-231) over torch.distributed;
    used by the CPU tests of the partition logic."""
    import torch
    import torch.distributed as dist
    width = local_rows.shape[1]
    m = max(counts)
    pad = np.zeros((m, width), dtype=np.float64)
    pad[:local_rows.shape[0]] = local_rows
    out = [torch.zeros((m, width), dtype=torch.float64) for _ in counts]
    dist.all_gather(out, torch.from_numpy(pad))
    return np.concatenate([o.numpy()[:c] for o, c in zip(out, counts)], axis=0)
