#!/usr/bin/env python
"""bench.py -- body-body interactions/sec of the all-pairs gravity hot path on N B200s.

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): N = 1,048,576-body
Plummer sphere in N-body units (G = 1, M = 1, a = 1, equal masses), softening eps = 0.01,
dt = 1/256, the reference's own integrator (v += a dt; x += v dt: cuda.cu:73-76, mpi.c:206-215,
i.e. symplectic Euler == leapfrog with velocities read half a step apart); one "step" = one full time step = N*(N-1) ordered
interactions (force pass + fused integration + position exchange when N_gpus > 1).

    python bench.py [--gpus N] [--steps K] [--warmup W]          # our CUDA path
    python bench.py --impl reference [--steps K] [--warmup W]    # the reference's own mpi.c on host cores
    torchrun --nproc-per-node N bench.py --gpus N ...            # one rank per GPU (strong scaling)

Prints ONE JSON line on rank 0.  `value` is device-timed (CUDA events on the stream the kernels
run on) with inputs resident in HBM; `e2e` is the same metric through the C-ABI with host buffers
(pinned H2D of the reference-layout records + step + D2H inside the timed region).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
from __graft_entry__ import load_package  # noqa: E402

N_BODIES = 1 << 20
EPS = 0.01
DT = 1.0 / 256
FLOP_PER_INTERACTION = 20            # direct-sum convention
REF_SAMPLE_N = 32768                 # bounded sample for the CPU reference arm
METRIC = "body-body interactions/sec at N=1M (Plummer sphere), 1/2/4/8 B200"
UNIT = "interactions/s"


def workload_config(n_gpus):
    return {"workload": "plummer_n1048576_euler", "n_bodies": N_BODIES, "eps": EPS, "dt": DT, "G": 1.0,
            "integrator": "symplectic-euler (reference: cuda.cu:73-76, mpi.c:206-215)", "interactions_per_step": N_BODIES * (N_BODIES - 1),
            "partition": f"{n_gpus} x N/{n_gpus} i-bodies, per-step exchange of 16 B/body positions over NVLink"
            if n_gpus > 1 else "single GPU",
            "l2": "flushed between timed steps (256 MiB write); positions 16 MiB"}


def traffic_bytes(n, world):
    """dram__bytes_read.sum + dram__bytes_write.sum of one force-kernel launch, from the committed ncu
    --set full capture of this same workload (profiles/force_kernel_traffic.json); None otherwise."""
    try:
        t = json.load(open(os.path.join(REPO, "profiles", "force_kernel_traffic.json")))
        return t["dram_traffic_bytes_per_launch"] if (n == N_BODIES and world == 1) else None
    except Exception:       # noqa: BLE001
        return None


def measured_peaks():
    try:
        return json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
    except Exception:       # noqa: BLE001
        return {}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except Exception:   # noqa: BLE001
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(2)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s, p in zip(sm, pw) if p >= 0.5 * max(pw)] or sm     # samples under load
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# reference arm: the reference's own mpi.c (oracle/_ref, compiled from /root/reference) on host cores
# ---------------------------------------------------------------------------------------------------
def run_reference_mpi(steps, warmup, n_sample=REF_SAMPLE_N):
    ref_dir = os.path.join(REPO, "oracle", "_ref")
    exe = os.path.join(ref_dir, "mpi_ref_o3")          # -O3 -march=native of the BUILD host: probe it first
    probe_env = dict(os.environ, GS_MPI_NP="1", GS_REF_N="8", GS_REF_STEPS="1")
    try:
        with tempfile.TemporaryDirectory() as pd:
            ok = subprocess.run([exe], env=probe_env, cwd=pd, capture_output=True, timeout=60).returncode == 0
    except Exception:       # noqa: BLE001
        ok = False
    if not ok:
        exe = os.path.join(ref_dir, "mpi_ref")         # plain -O2 build
    cores = os.cpu_count() or 1
    gs = load_package()
    bodies = gs.ic_reference_uniform(n_sample, seed=1234, dtype=np.float64)
    if not os.path.exists(exe):
        return None
    with tempfile.TemporaryDirectory() as wd:
        sin = os.path.join(wd, "state.bin")
        bodies.tofile(sin)

        def once(nsteps):
            env = dict(os.environ, GS_MPI_NP=str(cores), GS_REF_N=str(n_sample), GS_REF_STEPS=str(nsteps),
                       GS_REF_DT="3600.0", GS_REF_STATE=sin)
            r = subprocess.run([exe], env=env, cwd=wd, capture_output=True, text=True, timeout=1800)
            if r.returncode != 0:
                raise RuntimeError("reference mpi.c failed: " + r.stderr[-500:])
            span = [l for l in r.stdout.splitlines() if l.startswith("GS_REF_WTIME_SPAN")]
            return float(span[-1].split()[1])
        if warmup > 0:
            once(warmup)
        secs = once(steps)
    rate = n_sample * (n_sample - 1.0) * steps / secs
    return {"value": rate, "secs": secs, "cores": cores, "kind": "reference",
            "sample": f"reference mpi.c ({os.path.basename(exe)}: gcc -O3 -march=native, arithmetic untouched) over the "
                      f"fork-based mpi.h shim with {cores} ranks (no mpirun in the image), N={n_sample} reference-uniform "
                      f"bodies x {steps} steps = {n_sample * (n_sample - 1) * steps:.3e} interactions, timed by its own "
                      f"MPI_Wtime span (mpi.c:189,239); rate per interaction does not depend on N or on the body distribution"}


def main_reference(args, rank):
    if rank != 0:
        return 0
    res = run_reference_mpi(args.steps, args.warmup)
    if res is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/mpi_ref missing (built from /root/reference by oracle/Makefile)"}))
        return 0
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["secs"] / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {"value": res["value"], "unit": UNIT, "cores": res["cores"], "kind": res["kind"],
                             "sample": res["sample"]},
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=N_BODIES, help=argparse.SUPPRESS)      # exploration only
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return main_reference(args, rank)

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (libgravsim has no CPU path)")
    if world != args.gpus:
        raise SystemExit(f"bench.py: --gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={world})")
    torch.cuda.set_device(local_rank)
    gs = load_package()
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from gravsim_b200 import dist as gdist
        nccl_id = gdist.broadcast_bytes(gs.nccl_unique_id() if rank == 0 else None, 128, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n = args.n
    bodies = gs.ic_plummer(n, seed=2024)                      # identical on every rank (seeded)
    pinned_in = torch.from_numpy(bodies).pin_memory()
    pinned_out = torch.empty_like(pinned_in).pin_memory()
    stream = torch.cuda.Stream()               # the library launches on THIS stream; events are recorded on it
    torch.cuda.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sim = gs.GravSim(n, dt=DT, G=1.0, eps=EPS, integrator=gs.INTEGRATOR_EULER, device=local_rank,
                     rank=rank, world_size=world, nccl_id=nccl_id, stream=stream.cuda_stream)
    sim.upload_ptr(pinned_in.data_ptr(), n)

    # ---- device-resident timing -------------------------------------------------------------------
    for _ in range(args.warmup):
        sim.step(1)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = sim.kernel_launches
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.perf_counter()
    for a, b in evs:
        flush.fill_(1)                    # evict L2 between timed steps (not timed)
        a.record(stream)
        sim.step(1)
        b.record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = sim.kernel_launches - launches0
    clocks = sampler.stop() if rank == 0 else None
    ms_total = sum(a.elapsed_time(b) for a, b in evs)
    ms_t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_total = float(ms_t.item())
    inter_per_step = n * (n - 1.0)
    value = inter_per_step * args.steps / (ms_total * 1e-3)

    # ---- end-to-end through the C-ABI with host buffers ---------------------------------------------
    sim.upload_ptr(pinned_in.data_ptr(), n); sim.step(1); sim.download_ptr(pinned_out.data_ptr(), n)   # warm
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sim.upload_ptr(pinned_in.data_ptr(), n)
        sim.step(1)
        sim.download_ptr(pinned_out.data_ptr(), n)            # synchronises
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = inter_per_step * args.steps / float(t_e2e.item())
    assert np.isfinite(pinned_out.numpy()).all()
    info = sim.info
    sim.close()

    if rank == 0:
        peaks = measured_peaks()
        sm_max = peaks.get("sm_max_mhz", 1965.0)
        sms = info.sm_count
        peak_tf = sms * 128 * 2 * sm_max * 1e6 / 1e12                       # FP32 FMA peak, nominal boost clock
        per_gpu_rate = value / world
        ach_tf = per_gpu_rate * FLOP_PER_INTERACTION / 1e12
        roof = {"bound": "fp32", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf,
                "traffic": traffic_bytes(n, world),
                "note": f"CUDA-core FP32 FMA roofline ({sms} SMs x 128 lanes x 2 x {sm_max:.0f} MHz from MEASURED_PEAKS.json "
                        "sm_max_mhz; the file holds no FP32 figure): the interaction is not a dense contraction, so neither "
                        "the HBM nor the tensor roofline bounds it. achieved = 20 flop x N*(N-1)/n_gpus interactions per "
                        "force-kernel launch / CUDA-event step time (force kernel >99.9% of the step, profiles/).",
                "algorithmic_bytes_per_launch": 64.0 * n / world + 16.0 * n}
        # flops the kernel really executes per ordered interaction (FADD/FMUL = 1, FFMA = 2; MUFU.RSQ not counted):
        # row gather 18 (17 with equal masses); pair-symmetric 12.5 (11.5): d, r^2, rsqrt and rinv^3 are shared by
        # the two directions of a pair, so the 20-flop convention can exceed 100 % of the FMA peak.
        executed = (11.5 if info.uniform_mass else 12.5) if info.symmetric else (17.0 if info.uniform_mass else 18.0)
        roof["flop_per_interaction_convention"] = FLOP_PER_INTERACTION
        roof["flop_per_interaction_executed"] = executed
        roof["frac_executed"] = per_gpu_rate * executed / 1e12 / peak_tf
        if clocks and clocks.get("sm_mhz"):
            roof["frac_at_measured_clock"] = ach_tf / (sms * 128 * 2 * clocks["sm_mhz"] * 1e6 / 1e12)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": dict(workload_config(world), n_bodies=n, variant=info.variant, grid=info.grid,
                               i_tile=info.i_tile, j_tile=info.j_tile, regs=info.regs_per_thread,
                               force_pass="pair-symmetric (each unordered pair once, both bodies)" if info.symmetric
                               else "row gather (each ordered pair)",
                               mass_path="uniform (all masses equal: mass factor hoisted out of the inner loop)"
                               if info.uniform_mass else "general",
                               exchange=("peer-memory stores fused in the integrate epilogue + flags" if info.peer_exchange
                                         else "ncclAllGather") if world > 1 else "none"),
                "pct_fp32_peak": 100.0 * ach_tf / peak_tf,
                "wall_ms_per_step_incl_l2_flush": t_wall / args.steps * 1e3,
                "roofline": roof, "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 28 * n, "d2h_bytes_per_step": 28 * n,
                        "note": "per step: gs_upload_aos_f32 (pinned, reference 28-B records) + gs_step(1) + "
                                "gs_download_aos_f32; every rank moves the full array, as mpi.c's Bcast/Allgatherv do"},
                "gpu_launches": int(launches)}
        if world == 1 and not args.no_cpu_baseline:
            try:
                res = run_reference_mpi(steps=10, warmup=0)
                if res:
                    line["cpu_baseline"] = {"value": res["value"], "unit": UNIT, "cores": res["cores"],
                                            "kind": res["kind"], "sample": res["sample"]}
            except Exception as ex:     # noqa: BLE001
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                                        "sample": f"failed: {ex}"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
