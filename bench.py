#!/usr/bin/env python
"""bench.py -- body-body interactions/sec of the all-pairs gravity hot path on N B200s.

Default workload (BASELINE.json configs[2], the configuration the metric is quoted on): N = 1,048,576-body
Plummer sphere in N-body units (G = 1, M = 1, a = 1, equal masses), softening eps = 0.01, dt = 1/256, the
reference's own integrator (v += a dt; x += v dt: cuda.cu:73-76, mpi.c:206-215); one "step" = one full time
step = N*(N-1) ordered interactions (force pass + integration + position exchange when N_gpus > 1).

    python bench.py [--gpus N] [--steps K] [--warmup W]          # our CUDA path, headline workload
    python bench.py --impl reference [--steps K] [--warmup W]    # the reference's own mpi.c on host cores
    torchrun --nproc-per-node N bench.py --gpus N ...            # one rank per GPU (strong scaling)
    ... --workload ref65536 | galaxies4m | sweep                 # BASELINE.json configs[1], [3], [4]

Prints ONE JSON line on rank 0.  `value` is device-timed (CUDA events on the stream the kernels run on) with
inputs resident in HBM; `e2e` is the same metric through the C-ABI with host buffers (every rank uploads the
reference-layout records of its own bodies from pinned memory, steps, and downloads them again, inside the timed
region); `checks` are numerical self-checks of the run that was timed (library only: no oracle on this leg).
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

N_BODIES = 1 << 20
EPS = 0.01
DT = 1.0 / 256
FLOP_PER_INTERACTION = 20            # direct-sum convention
REF_SAMPLE_N = 32768                 # bounded sample for the CPU reference arm
CUDA_REF_N, CUDA_REF_STEPS = 65536, 20
METRIC = "body-body interactions/sec at N=1M (Plummer sphere), 1/2/4/8 B200"
UNIT = "interactions/s"
# bounds of the run's self-checks (`checks` in the JSON line)
CHECK_MOMENTUM = 1e-6      # |sum m a| / sum |m a| over all bodies, one step from rest
CHECK_DET_SAMPLE = 2e-5    # default pass vs GS_FLAG_DETERMINISTIC context, 256 bodies per rank, floor-normalised
CHECK_ENERGY = 2e-3        # |E_after - E_before| / |E_before| over the timed steps (symplectic Euler, dt = 1/256)


def load_gs():
    from __graft_entry__ import load_package
    return load_package()


def workload_config(n_gpus, n=N_BODIES):
    return {"workload": f"plummer_n{n}_euler", "n_bodies": n, "eps": EPS, "dt": DT, "G": 1.0,
            "integrator": "symplectic-euler (reference: cuda.cu:73-76, mpi.c:206-215)", "interactions_per_step": n * (n - 1),
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


def reference_uniform_numpy(n, seed):
    """Bodies in the reference's own distribution (cuda.cu:82-93,129-131 == mpi.c:76-104), generated WITHOUT the
    product library so that the reference arm's process never maps libgravsim.so: Sun/Earth/Mars, then uniform
    position +-3e11 m, velocity +-3e4 m/s, mass 1e23..1e25 kg.  fp32-representable values in f64 records."""
    rng = np.random.default_rng(seed)
    b = np.empty((n, 7), dtype=np.float64)
    b[:, 0:3] = rng.uniform(-3e11, 3e11, (n, 3))
    b[:, 3:6] = rng.uniform(-3e4, 3e4, (n, 3))
    b[:, 6] = rng.uniform(1e23, 1e25, n)
    head = np.array([[0, 0, 0, 0, 0, 0, 1.989e30],
                     [1.496e11, 0, 0, 0, 29780.0, 0, 5.972e24],
                     [2.279e11, 0, 0, 0, 24077.0, 0, 6.39e23]])
    b[:min(3, n)] = head[:min(3, n)]
    return b.astype(np.float32).astype(np.float64)


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
    if not os.path.exists(exe):
        return None
    cores = os.cpu_count() or 1
    bodies = reference_uniform_numpy(n_sample, seed=1234)
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
    return {"value": rate, "secs": secs, "cores": cores, "kind": "reference", "n_sample": n_sample,
            "sample": f"reference mpi.c ({os.path.basename(exe)}: gcc -O3 -march=native, arithmetic untouched) over the "
                      f"fork-based mpi.h shim with {cores} ranks (no mpirun in the image), N={n_sample} reference-uniform "
                      f"bodies x {steps} steps = {n_sample * (n_sample - 1) * steps:.3e} interactions, timed by its own "
                      f"MPI_Wtime span (mpi.c:189,239); rate per interaction does not depend on N or on the body distribution"}


def run_reference_cuda(n=CUDA_REF_N, steps=CUDA_REF_STEPS, bodies=None):
    """The reference's cuda.cu (oracle/_ref/cuda_ref: patched ONLY for N, step count and reading its input from a
    file; kernel, host update and timer untouched) on this box's GPU, timed by its own chrono span
    (cuda.cu:154,169-171).  Its output is not finite (cuda.cu:41 overflows fp32): a timing comparator only."""
    exe = os.path.join(REPO, "oracle", "_ref", "cuda_ref")
    if not os.path.exists(exe):
        return {"value": None, "unit": UNIT, "note": "oracle/_ref/cuda_ref missing (built from /root/reference by oracle/Makefile)"}
    b = reference_uniform_numpy(n, seed=4321).astype(np.float32) if bodies is None else np.ascontiguousarray(bodies, dtype=np.float32)
    with tempfile.TemporaryDirectory() as wd:
        sf = os.path.join(wd, "state.bin")
        b.tofile(sf)
        env = dict(os.environ, GS_REF_N=str(n), GS_REF_STEPS=str(steps), GS_REF_STATE=sf)
        r = subprocess.run([exe], env=env, cwd=wd, capture_output=True, text=True, timeout=900)
    m = re.search(r"Simulation took ([0-9.eE+-]+) seconds", r.stdout)
    if not m:
        return {"value": None, "unit": UNIT, "note": "cuda_ref failed: " + (r.stderr or r.stdout)[-200:]}
    pos = re.findall(r"Particle \d+: \((.*)\)", r.stdout)
    finite = [all(np.isfinite(float(x)) for x in p.replace("-nan", "nan").split(",")) for p in pos]
    secs = float(m.group(1))
    return {"value": n * (n - 1.0) * steps / secs, "unit": UNIT, "n_bodies": n, "steps": steps, "seconds": secs,
            "finite_printed_positions": f"{sum(finite)}/{len(finite)}", "kind": "reference cuda.cu",
            "note": "reference cuda.cu on this GPU, same distribution as configs[1] (N=65,536 reference-uniform), its own "
                    "chrono span cuda.cu:154,169-171 (kernel + sync + D2H + host update per step); N(N-1)/2 pair "
                    "evaluations credited as N(N-1) interactions; output non-finite as written (cuda.cu:41)"}


def main_reference(args, rank):
    if rank != 0:
        return 0
    res = run_reference_mpi(args.steps, args.warmup)
    if res is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/mpi_ref missing (built from /root/reference by oracle/Makefile)"}))
        return 0
    ns = res["n_sample"]
    cfg = {"workload": f"reference_uniform_n{ns}_eps0_sample", "stands_for": workload_config(args.gpus)["workload"],
           "extrapolated": True, "n_bodies": ns, "eps": 0.0, "dt": 3600.0, "G": 6.6743e-11,
           "integrator": "mpi.c:206-215 as written (in-place)", "interactions_per_step": ns * (ns - 1),
           "why": "mpi.c is O(N^2) fp64 at ~2.5e8 interactions/s/core: one step of the 1M-body workload would take minutes "
                  "on the host; its rate per interaction does not depend on N or on the distribution, so a bounded "
                  "sample of its own distribution is timed and the RATE is compared"}
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["secs"] / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": res["value"], "unit": UNIT, "cores": res["cores"], "kind": res["kind"],
                             "sample": res["sample"]},
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------
class Env:
    """torch is here for streams, events, pinned buffers and the torchrun rendezvous only."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (libgravsim has no CPU path)")
        if self.world != args.gpus:
            raise SystemExit(f"bench.py: --gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={self.world})")
        torch.cuda.set_device(self.local_rank)
        self.gs = load_gs()
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        self.stream = torch.cuda.Stream()      # the library launches on THIS stream; events are recorded on it
        torch.cuda.set_stream(self.stream)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def nccl_id(self):
        if self.world == 1:
            return None
        from gravsim_b200 import dist as gdist
        return gdist.broadcast_bytes(self.gs.nccl_unique_id() if self.rank == 0 else None, 128, device="cuda")

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, values, op="max"):
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return [float(x) for x in t.tolist()]

    def sim(self, n, **kw):
        return self.gs.GravSim(n, device=self.local_rank, rank=self.rank, world_size=self.world, nccl_id=self.nccl_id(),
                               stream=self.stream.cuda_stream, **kw)

    def time_steps(self, sim, steps, per_step_flush=True, batch=1):
        """CUDA-event time of `steps` x gs_step(batch) on the engine's stream, max over ranks [ms]."""
        tc = self.torch.cuda
        evs = [(tc.Event(enable_timing=True), tc.Event(enable_timing=True)) for _ in range(steps)]
        self.barrier()
        t0 = time.perf_counter()
        for a, b in evs:
            if per_step_flush:
                self.flush.fill_(1)               # evict L2 between timed steps (not timed)
            a.record(self.stream)
            sim.step(batch)
            b.record(self.stream)
        self.barrier()
        wall = time.perf_counter() - t0
        ms = sum(a.elapsed_time(b) for a, b in evs)
        return self.reduce([ms])[0], wall

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def fp32_peak_tflops(sm_count):
    sm_max = measured_peaks().get("sm_max_mhz", 1965.0)
    return sm_count * 128 * 2 * sm_max * 1e6 / 1e12, sm_max


def executed_flops(info):
    # flops the kernel really executes per ordered interaction (FADD/FMUL = 1, FFMA = 2; MUFU.RSQ not counted):
    # row gather 18 (17 with equal masses); pair-symmetric 12.5 (11.5): d, r^2, rsqrt and rinv^3 are shared by
    # the two directions of a pair, so the 20-flop convention can exceed 100 % of the FMA peak.
    return (11.5 if info.uniform_mass else 12.5) if info.symmetric else (17.0 if info.uniform_mass else 18.0)


def describe(info, world):
    return dict(variant=info.variant, grid=info.grid, i_tile=info.i_tile, j_tile=info.j_tile, regs=info.regs_per_thread,
                force_pass="pair-symmetric (each unordered pair once, both bodies)" if info.symmetric
                else "row gather (each ordered pair)",
                mass_path="uniform (all masses equal: mass factor hoisted out of the inner loop)"
                if info.uniform_mass else "general",
                exchange=("peer-memory stores fused in the integrate epilogue + flags" if info.peer_exchange
                          else "ncclAllGather") if world > 1 else "none")


def self_checks(env, bodies, n, G, eps, dt, sim, e_before, e_after):
    """Numerical checks of the configuration that was just timed, from the library alone:
    (1) one step of the default pass from rest on every rank: a = v1/dt for ALL bodies; sum m a / sum |m a|
        (Newton's third law over every tile pair and peer atomic);
    (2) the same step in a second context built with GS_FLAG_DETERMINISTIC (row-gather kernel, other partition,
        no atomics): 256 bodies per rank, |a_sym - a_det| / max(|a_det|, 1 % of the median |a|);
    (3) |dE/E0| over the timed steps (gs_energy: fp64 accumulation, collective)."""
    gs = env.gs
    rest = bodies.copy()
    rest[:, 3:6] = 0.0
    lo, hi = sim.my_range
    acc = {}
    for name, ctx in (("default", sim), ("det", None)):
        own = ctx is None
        if own:
            ctx = env.sim(n, dt=dt, G=G, eps=eps, flags=gs.FLAG_DETERMINISTIC, variant=sim.info.variant)   # same i-tile: same slices
        ctx.upload_slice(rest[lo:hi])
        ctx.step(1)
        out = ctx.download_slice()
        acc[name] = out[:, 3:6].astype(np.float64) / dt
        if own:
            ctx.close()
    m = bodies[lo:hi, 6:7].astype(np.float64)
    f = m * acc["default"]
    tot = env.reduce(list(f.sum(0)) + [np.linalg.norm(f, axis=1).sum()], op="sum")
    momentum = float(np.linalg.norm(tot[:3]) / tot[3])
    idx = np.sort(np.random.default_rng(env.rank).choice(hi - lo, min(256, hi - lo), replace=False))
    an = np.linalg.norm(acc["det"], axis=1)
    floor = 0.01 * np.median(an)
    d = np.linalg.norm(acc["default"][idx] - acc["det"][idx], axis=1) / np.maximum(an[idx], floor)
    det = env.reduce([float(d.max())])[0]
    drift = abs(e_after - e_before) / abs(e_before)
    ok = bool(momentum <= CHECK_MOMENTUM and det <= CHECK_DET_SAMPLE and drift <= CHECK_ENERGY and np.isfinite(an).all())
    return {"ok": ok, "momentum_residual": momentum, "momentum_bound": CHECK_MOMENTUM,
            "default_vs_deterministic_sample_max": det, "sample_bound": CHECK_DET_SAMPLE, "sample_per_rank": int(idx.size),
            "energy_drift_over_timed_steps": drift, "energy_bound": CHECK_ENERGY,
            "how": "one step from rest on every rank (a = v1/dt), all bodies for the momentum sum; a second context with "
                   "GS_FLAG_DETERMINISTIC stepped from the same upload for the sample; gs_energy before/after the timed loop"}


def run_headline(env, args):
    gs, torch = env.gs, env.torch
    rank, world = env.rank, env.world
    n = args.n
    bodies = gs.ic_plummer(n, seed=2024)                      # identical on every rank (seeded)
    pinned_in = torch.from_numpy(bodies).pin_memory()
    pinned_out = torch.empty_like(pinned_in).pin_memory()
    sim = env.sim(n, dt=DT, G=1.0, eps=EPS, integrator=gs.INTEGRATOR_EULER)
    lo, hi = sim.my_range
    rec = pinned_in.element_size() * 7
    in_ptr, out_ptr = pinned_in.data_ptr() + lo * rec, pinned_out.data_ptr() + lo * rec
    sim.upload_slice_ptr(in_ptr, hi - lo)

    # ---- device-resident timing -------------------------------------------------------------------
    for _ in range(args.warmup):
        sim.step(1)
    env.barrier()
    e_before = e_after = None
    if not args.no_checks:
        ke, pe, _ = sim.energy()
        e_before = ke + pe
    sampler = ClockSampler(env.local_rank)
    if rank == 0:
        sampler.start()
    launches0 = sim.kernel_launches
    ms_total, t_wall = env.time_steps(sim, args.steps)
    launches = sim.kernel_launches - launches0
    clocks = sampler.stop() if rank == 0 else None
    if not args.no_checks:
        ke, pe, _ = sim.energy()
        e_after = ke + pe
    inter_per_step = n * (n - 1.0)
    value = inter_per_step * args.steps / (ms_total * 1e-3)

    # ---- end-to-end through the C-ABI with host buffers: every rank moves the records of its own bodies ----
    sim.upload_slice_ptr(in_ptr, hi - lo); sim.step(1); sim.download_slice_ptr(out_ptr, hi - lo)   # warm
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sim.upload_slice_ptr(in_ptr, hi - lo)
        sim.step(1)
        sim.download_slice_ptr(out_ptr, hi - lo)            # synchronises
    env.barrier()
    e2e_value = inter_per_step * args.steps / env.reduce([time.perf_counter() - t0])[0]
    assert np.isfinite(pinned_out.numpy()[lo:hi]).all()
    info = sim.info
    checks = None if args.no_checks else self_checks(env, bodies, n, 1.0, EPS, DT, sim, e_before, e_after)
    sim.close()
    if rank != 0:
        return 0 if (checks is None or checks["ok"]) else 1

    peak_tf, sm_max = fp32_peak_tflops(info.sm_count)
    per_gpu_rate = value / world
    ach_tf = per_gpu_rate * FLOP_PER_INTERACTION / 1e12
    roof = {"bound": "fp32", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf,
            "traffic": traffic_bytes(n, world),
            "note": f"CUDA-core FP32 FMA roofline ({info.sm_count} SMs x 128 lanes x 2 x {sm_max:.0f} MHz from MEASURED_PEAKS.json "
                    "sm_max_mhz; the file holds no FP32 figure): the interaction is not a dense contraction, so neither "
                    "the HBM nor the tensor roofline bounds it. achieved = 20 flop x N*(N-1)/n_gpus interactions per "
                    "force-kernel launch / CUDA-event step time (force kernel >99.9% of the step, profiles/). traffic: "
                    "from the committed ncu --set full capture of this workload, not measured in this run.",
            # positions of all bodies once (16 B) + per own body: i-side sums (16 B) and the j-side sums, 3 x 5 64-bit
            # words of wide fixed point (120 B) written by atomics in the L2, then read and cleared by the finish kernel
            "algorithmic_bytes_per_launch": 16.0 * n + (16.0 + 3 * 120.0) * n / world,
            "flop_per_interaction_convention": FLOP_PER_INTERACTION,
            "flop_per_interaction_executed": executed_flops(info),
            "frac_executed": per_gpu_rate * executed_flops(info) / 1e12 / peak_tf}
    if clocks and clocks.get("sm_mhz"):
        roof["frac_at_measured_clock"] = ach_tf / (info.sm_count * 128 * 2 * clocks["sm_mhz"] * 1e6 / 1e12)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(world, n), **describe(info, world)),
            "pct_fp32_peak": 100.0 * ach_tf / peak_tf,
            "wall_ms_per_step_incl_l2_flush": t_wall / args.steps * 1e3,
            "roofline": roof, "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 28 * n, "d2h_bytes_per_step": 28 * n,
                    "note": "per step and rank: gs_upload_slice_f32 (pinned, reference 28-B records of the rank's own "
                            "bodies, mpi.c:184-187) + gs_step(1) + gs_download_slice_f32; bytes are totals over all ranks"},
            "gpu_launches": int(launches)}
    if checks is not None:
        line["checks"] = checks
    if world == 1 and not args.no_cpu_baseline:
        try:
            res = run_reference_mpi(steps=10, warmup=0)
            if res:
                line["cpu_baseline"] = {"value": res["value"], "unit": UNIT, "cores": res["cores"],
                                        "kind": res["kind"], "sample": res["sample"]}
        except Exception as ex:     # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                                    "sample": f"failed: {ex}"}
    if world == 1 and not args.no_cuda_baseline:
        try:
            line["cuda_baseline"] = run_reference_cuda()
        except Exception as ex:     # noqa: BLE001
            line["cuda_baseline"] = {"value": None, "unit": UNIT, "note": f"failed: {ex}"}
    print(json.dumps(line), flush=True)
    return 0 if (checks is None or checks["ok"]) else 1


def run_ref65536(env, args):
    """BASELINE.json configs[1]: N = 65,536 bodies in the reference's own distribution, the reference law (eps = 0:
    bare Newton, r < 1e-10 skipped), dt = 3600 s, 1000 steps, beside the reference's cuda.cu on the same inputs."""
    gs = env.gs
    n, steps = (args.n if args.n != N_BODIES else 65536), (args.steps if args.steps_given else 1000)
    b = gs.ic_reference_uniform(n, seed=4321)
    sim = env.sim(n, dt=3600.0)
    lo, hi = sim.my_range
    sim.upload_slice(b[lo:hi])
    sim.step(max(args.warmup, 3))
    sampler = ClockSampler(env.local_rank)
    if env.rank == 0:
        sampler.start()
    l0 = sim.kernel_launches
    ms, _ = env.time_steps(sim, 1, per_step_flush=False, batch=steps)
    launches = sim.kernel_launches - l0
    clocks = sampler.stop() if env.rank == 0 else None
    out = sim.download()
    info = sim.info
    sim.close()
    if env.rank != 0:
        return 0
    value = n * (n - 1.0) * steps / (ms * 1e-3)
    peak_tf, _ = fp32_peak_tflops(info.sm_count)
    line = {"metric": "body-body interactions/sec, N=65,536 reference-uniform bodies, reference force law (configs[1])",
            "value": value, "unit": UNIT, "n_gpus": env.world, "steps": steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": dict({"workload": f"reference_uniform_n{n}_eps0_euler", "n_bodies": n, "eps": 0.0, "dt": 3600.0,
                            "G": 6.6743e-11, "l2": "not flushed: the steps run back to back as the configuration says "
                            "(positions 1 MiB, L2-resident by nature)"}, **describe(info, env.world)),
            "pct_fp32_peak": 100.0 * value / env.world * FLOP_PER_INTERACTION / 1e12 / peak_tf,
            "all_finite": bool(np.isfinite(out).all()), "clocks": clocks, "gpu_launches": int(launches)}
    if env.world == 1 and not args.no_cuda_baseline:
        line["cuda_baseline"] = run_reference_cuda(n, CUDA_REF_STEPS, bodies=b)
        if line["cuda_baseline"].get("value"):
            line["speedup_vs_reference_cuda_cu"] = value / line["cuda_baseline"]["value"]
    print(json.dumps(line), flush=True)
    return 0


GALAXY_DRIFT_BOUND = 1e-6


def run_galaxies(env, args):
    """BASELINE.json configs[3]: two Plummer systems of N/2 bodies each on an offset approach orbit, N = 4,194,304,
    500 steps, kick-drift-kick leapfrog; total energy (gs_energy: fp64 accumulation) at the start and every 100
    steps; the run FAILS if max |E - E0| / |E0| exceeds GALAXY_DRIFT_BOUND."""
    gs = env.gs
    n, steps = (args.n if args.n != N_BODIES else 1 << 22), (args.steps if args.steps_given else 500)
    b = gs.ic_two_galaxy(n, seed=7)
    sim = env.sim(n, dt=DT, G=1.0, eps=EPS, integrator=gs.INTEGRATOR_LEAPFROG)
    lo, hi = sim.my_range
    sim.upload_slice(b[lo:hi])
    ke, pe, mom0 = sim.energy()
    e0 = ke + pe
    sampler = ClockSampler(env.local_rank)
    if env.rank == 0:
        sampler.start()
    l0 = sim.kernel_launches
    energies, ms_total, done = [{"step": 0, "E": e0, "KE": ke, "PE": pe}], 0.0, 0
    while done < steps:
        blk = min(100, steps - done)
        ms, _ = env.time_steps(sim, 1, per_step_flush=False, batch=blk)
        ms_total += ms
        done += blk
        ke, pe, mom = sim.energy()                  # closes the leapfrog half kick; not timed
        energies.append({"step": done, "E": ke + pe, "KE": ke, "PE": pe, "rel_drift": abs(ke + pe - e0) / abs(e0)})
    launches = sim.kernel_launches - l0
    clocks = sampler.stop() if env.rank == 0 else None
    part = sim.download_slice()
    finite = env.reduce([0.0 if np.isfinite(part).all() else 1.0])[0] == 0.0
    info = sim.info
    sim.close()
    drift = max(e["rel_drift"] for e in energies[1:])
    ok = bool(drift <= GALAXY_DRIFT_BOUND and finite)
    if env.rank == 0:
        value = n * (n - 1.0) * steps / (ms_total * 1e-3)
        peak_tf, _ = fp32_peak_tflops(info.sm_count)
        line = {"metric": "body-body interactions/sec, two-galaxy collision N=4M, leapfrog (configs[3])", "value": value,
                "unit": UNIT, "n_gpus": env.world, "steps": steps, "warmup": 0, "ms_per_step": ms_total / steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": dict({"workload": f"two_galaxy_n{n}_leapfrog", "n_bodies": n, "eps": EPS, "dt": DT, "G": 1.0,
                                "ic": "gs_ic_two_galaxy_f32(seed 7): 2 x N/2 Plummer, separation 8, impact 2, v_rel 0.5",
                                "l2": "not flushed: 500 consecutive steps; positions 64 MiB"}, **describe(info, env.world)),
                "pct_fp32_peak_per_gpu": 100.0 * value / env.world * FLOP_PER_INTERACTION / 1e12 / peak_tf,
                "energy": energies, "max_rel_energy_drift": drift, "energy_drift_bound": GALAXY_DRIFT_BOUND,
                "momentum_start": list(map(float, mom0)), "momentum_end": list(map(float, mom)),
                "all_finite": finite, "ok": ok, "clocks": clocks, "gpu_launches": int(launches)}
        print(json.dumps(line), flush=True)
    return 0 if ok else 1


def run_sweep(env, args):
    """BASELINE.json configs[4]: N = 16K ... 16M (factor 2) at this GPU count, Plummer spheres, the headline settings.
    One row per N; tools/merge_sweep.py combines the lines of P = 1, 2, 4, 8 into strong / weak efficiencies."""
    gs = env.gs
    rows = []
    n = 16384
    while n <= args.sweep_max:
        est = n * (n - 1.0) / (4.0e12 * env.world)                       # s per step, rough
        steps = int(min(50, max(1, round(1.5 / max(est, 1e-4)))))
        warm = 3 if est < 2.0 else 1
        b = gs.ic_plummer(n, seed=11)
        sim = env.sim(n, dt=DT, G=1.0, eps=EPS)
        lo, hi = sim.my_range
        sim.upload_slice(b[lo:hi])
        sim.step(warm)
        sampler = ClockSampler(env.local_rank)
        if env.rank == 0:
            sampler.start()
        ms, _ = env.time_steps(sim, steps, per_step_flush=(n * 16 <= (126 << 20)))
        clocks = sampler.stop() if env.rank == 0 else None
        info = sim.info
        part = sim.download_slice()
        finite = env.reduce([0.0 if np.isfinite(part).all() else 1.0])[0] == 0.0
        sim.close()
        rate = n * (n - 1.0) * steps / (ms * 1e-3)
        peak_tf, _ = fp32_peak_tflops(info.sm_count)
        rows.append({"n_bodies": n, "n_gpus": env.world, "steps": steps, "warmup": warm, "ms_per_step": ms / steps,
                     "value": rate, "pct_fp32_peak_per_gpu": 100.0 * rate / env.world * FLOP_PER_INTERACTION / 1e12 / peak_tf,
                     "frac_executed_per_gpu": rate / env.world * executed_flops(info) / 1e12 / peak_tf,
                     "variant": info.variant, "i_tile": info.i_tile, "symmetric": int(info.symmetric), "n_pad": info.n_pad,
                     "all_finite": finite, "sm_mhz": clocks.get("sm_mhz") if clocks else None,
                     "reasons": clocks.get("reasons") if clocks else None})
        if env.rank == 0:
            print(f"# sweep P={env.world} N={n}: {ms / steps:.3f} ms/step {rate:.4e} int/s", file=sys.stderr, flush=True)
        n *= 2
    if env.rank == 0:
        print(json.dumps({"metric": "body-body interactions/sec, N sweep (configs[4])", "unit": UNIT, "n_gpus": env.world,
                          "config": {"workload": "plummer_sweep_16k_16m_euler", "eps": EPS, "dt": DT, "G": 1.0,
                                     "l2": "flushed between timed steps while the positions fit the 126 MiB L2"},
                          "rows": rows}), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="plummer1m", choices=["plummer1m", "ref65536", "galaxies4m", "sweep"])
    ap.add_argument("--n", type=int, default=N_BODIES, help=argparse.SUPPRESS)      # exploration only
    ap.add_argument("--sweep-max", type=int, default=1 << 24)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cuda-baseline", action="store_true")
    ap.add_argument("--no-checks", action="store_true")
    args = ap.parse_args()
    args.steps_given = args.steps is not None
    if args.steps is None:
        args.steps = 5
    if args.impl == "reference":
        return main_reference(args, int(os.environ.get("RANK", "0")))
    args.warmup = max(args.warmup, 3)
    env = Env(args)
    rc = {"plummer1m": run_headline, "ref65536": run_ref65536, "galaxies4m": run_galaxies, "sweep": run_sweep}[args.workload](env, args)
    env.close()
    return rc


if __name__ == "__main__":
    sys.exit(main())
