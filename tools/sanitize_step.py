"""Small run for compute-sanitizer (racecheck / memcheck): the default pair-symmetric pass, tracked cutoff with a
redo (coincident bodies), on one GPU and, when two are present, on two GPUs (one host thread each: peer access,
system-scope atomics into the peer's j-side sums, flag protocol).  Prints max errors against nothing: the sanitizer's
report is the result; numerical parity is tests/'s job."""
import os
import sys
import threading

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

gs = load_package()
W = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n = int(sys.argv[2]) if len(sys.argv) > 2 else 9001
b = gs.ic_reference_uniform(n, seed=3)
b[n - 5] = b[7]                                   # coincident pair in different tiles: exercises the redo
uid = gs.nccl_unique_id() if W > 1 else None
out = [None] * W


def work(rank):
    with gs.GravSim(n, device=rank, rank=rank, world_size=W, nccl_id=uid, variant=3) as sim:
        assert sim.info.symmetric == 1
        lo, hi = sim.my_range
        sim.upload_slice(b[lo:hi])
        sim.step(2)
        out[rank] = sim.download()


ts = [threading.Thread(target=work, args=(r,)) for r in range(W)]
[t.start() for t in ts]
[t.join() for t in ts]
print("finite:", all(np.isfinite(o).all() for o in out), "ranks agree:", all(np.array_equal(o, out[0]) for o in out))
