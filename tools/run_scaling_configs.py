"""BASELINE.json configs[3] (two-galaxy collision, energy drift) and configs[4] (N sweep x GPU count).

    python tools/run_scaling_configs.py sweep  [N,N,...]                 # 1 GPU
    torchrun --nproc-per-node P tools/run_scaling_configs.py sweep ...   # P GPUs (strong scaling per N)
    torchrun --nproc-per-node 8 tools/run_scaling_configs.py galaxy N STEPS

Rank 0 prints one JSON object.  Device-timed with CUDA events on the engine's stream, max over ranks.
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from __graft_entry__ import load_package  # noqa: E402

gs = load_package()
from gravsim_b200 import dist as gdist  # noqa: E402

PEAK = 148 * 128 * 2 * 1.965e9
rank, world, local = gdist.env_ranks()
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def fresh_id():
    """One NCCL id per engine context (an id initialises exactly one communicator)."""
    if world == 1:
        return None
    return gdist.broadcast_bytes(gs.nccl_unique_id() if rank == 0 else None, 128, device="cuda")


stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x):
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed_steps(sim, steps):
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    sim.step(steps)
    b.record(stream)
    barrier()
    return max_over_ranks(a.elapsed_time(b)) * 1e-3


def sweep(ns):
    rows = []
    for n in ns:
        bodies = gs.ic_plummer(n, seed=7)
        per_step = n * (n - 1.0) / (2.6e12 * world)
        steps = int(max(2, min(400, round(1.5 / max(per_step, 1e-5)))))
        with gs.GravSim(n, dt=1.0 / 256, G=1.0, eps=0.01, device=local, rank=rank, world_size=world,
                        nccl_id=fresh_id(), stream=stream.cuda_stream) as sim:
            sim.upload(bodies)
            sim.step(2 if per_step < 0.5 else 1)
            secs = timed_steps(sim, steps)
            info = sim.info
        rate = n * (n - 1.0) * steps / secs
        rows.append({"n": n, "n_gpus": world, "steps": steps, "ms_per_step": secs / steps * 1e3, "rate": rate,
                     "pct_fp32_peak_per_gpu": 100 * rate * 20 / PEAK / world, "variant": info.variant,
                     "grid": info.grid, "n_pad": info.n_pad})
        if rank == 0:
            print(f"# N={n} P={world} {secs / steps * 1e3:.3f} ms/step {rate:.3e} int/s "
                  f"{rows[-1]['pct_fp32_peak_per_gpu']:.1f}%/GPU variant {info.variant}", file=sys.stderr, flush=True)
    return {"config": "sweep", "workload": "plummer eps=0.01 euler", "rows": rows}


def galaxy(n, steps):
    bodies = gs.ic_two_galaxy(n, seed=11, sep=8.0, impact=2.0, vrel=0.5)
    dt, eps = 1.0 / 128, 0.01
    with gs.GravSim(n, dt=dt, G=1.0, eps=eps, integrator=gs.INTEGRATOR_LEAPFROG, device=local, rank=rank,
                    world_size=world, nccl_id=fresh_id(), stream=stream.cuda_stream) as sim:
        sim.upload(bodies)
        t0 = time.perf_counter()
        ke0, pe0, mom0 = sim.energy()
        t_energy = time.perf_counter() - t0
        checks = [{"step": 0, "ke": ke0, "pe": pe0, "E": ke0 + pe0}]
        total = 0.0
        done = 0
        chunk = max(1, steps // 5)
        while done < steps:
            k = min(chunk, steps - done)
            total += timed_steps(sim, k)
            done += k
            ke, pe, mom = sim.energy()
            checks.append({"step": done, "ke": ke, "pe": pe, "E": ke + pe, "rel_drift": (ke + pe - checks[0]["E"]) / abs(checks[0]["E"])})
        info = sim.info
    rate = n * (n - 1.0) * steps / total
    return {"config": "two_galaxy", "n": n, "n_gpus": world, "steps": steps, "dt": dt, "eps": eps,
            "integrator": "leapfrog-kdk", "ms_per_step": total / steps * 1e3, "rate": rate,
            "pct_fp32_peak_per_gpu": 100 * rate * 20 / PEAK / world, "energy_pass_seconds": t_energy,
            "energy": checks, "max_abs_rel_drift": max(abs(c.get("rel_drift", 0.0)) for c in checks),
            "momentum_initial": list(mom0), "momentum_final": list(mom), "variant": info.variant}


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "sweep":
        ns = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [16384, 65536, 262144, 1048576]
        res = sweep(ns)
    else:
        res = galaxy(int(sys.argv[2]), int(sys.argv[3]))
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
