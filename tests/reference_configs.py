"""BASELINE.json configs[0] and configs[1], ours next to the reference's own programs on the same inputs.

  configs[0]: N=4096 reference-uniform bodies, 100 steps, reference mpi.c with 4 ranks (fork shim)
  configs[1]: N=65,536 bodies, fp32, 1000 steps on 1 B200 vs reference cuda.cu (fewer steps: it is slow,
              and its output is non-finite -- timing comparator only)

Lives under tests/ because it drives the checkers in oracle/ (only tests, smoke() and bench.py's baseline legs may).
Run on the GPU box:  python tests/reference_configs.py > gpurun_out/reference_configs.json
"""
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
from __graft_entry__ import load_package  # noqa: E402
from conftest import Oracle, Reference    # noqa: E402  (checkers: this tool is measurement, not product)

gs = load_package()
out = {}

# ---- configs[0] --------------------------------------------------------------------------------
n, steps, dt = 4096, 100, 3600.0
b = gs.ic_reference_uniform(n, seed=1234)
ref = Reference()
with tempfile.TemporaryDirectory() as wd:
    ref_np4, ref_secs = ref.run_main(b.astype(np.float64), dt, steps, 4, wd)
with gs.GravSim(n, dt=dt) as sim:
    sim.upload(b); sim.step(2); sim.sync()
    sim.upload(b)
    t0 = time.perf_counter(); sim.step(steps); sim.sync(); ours_secs = time.perf_counter() - t0
    ours = sim.download()
sync = Oracle().run_sync(b.astype(np.float64), dt, steps)
scale = np.abs(sync[:, :3]).max()
out["config0_n4096_100steps"] = {
    "reference_mpi_c_np4_seconds": ref_secs, "reference_rate": n * (n - 1.0) * steps / ref_secs,
    "ours_seconds": ours_secs, "ours_rate": n * (n - 1.0) * steps / ours_secs,
    "max_pos_diff_ours_vs_synchronous_oracle_rel": float(np.abs(ours[:, :3] - sync[:, :3]).max() / scale),
    "max_pos_diff_mpi_np4_vs_synchronous_oracle_rel": float(np.abs(ref_np4[:, :3] - sync[:, :3]).max() / scale),
    "max_pos_diff_ours_vs_mpi_np4_rel": float(np.abs(ours[:, :3] - ref_np4[:, :3]).max() / scale),
    "note": "mpi.c as written updates in place, so it differs from the synchronous step by its own -np artefact"}

# ---- configs[1] --------------------------------------------------------------------------------
n, steps = 65536, 1000
b = gs.ic_reference_uniform(n, seed=4321)
with gs.GravSim(n, dt=dt) as sim:
    sim.upload(b); sim.step(3); sim.sync()
    sim.upload(b)
    t0 = time.perf_counter(); sim.step(steps); sim.sync(); ours_secs = time.perf_counter() - t0
    ours = sim.download()
    info = sim.info
res = {"ours_seconds_1000_steps": ours_secs, "ours_rate": n * (n - 1.0) * steps / ours_secs,
       "ours_all_finite": bool(np.isfinite(ours).all()), "variant": info.variant}
cuda_ref = os.path.join(REPO, "oracle", "_ref", "cuda_ref")
if os.path.exists(cuda_ref):
    ref_steps = 20
    with tempfile.TemporaryDirectory() as wd:
        sf = os.path.join(wd, "state.bin")
        b.tofile(sf)
        env = dict(os.environ, GS_REF_N=str(n), GS_REF_STEPS=str(ref_steps), GS_REF_STATE=sf)
        r = subprocess.run([cuda_ref], env=env, cwd=wd, capture_output=True, text=True, timeout=900)
    m = re.search(r"Simulation took ([0-9.eE+-]+) seconds", r.stdout)
    pos = re.findall(r"Particle \d+: \((.*)\)", r.stdout)
    finite = [all(np.isfinite(float(x)) for x in p.replace("-nan", "nan").split(",")) for p in pos]
    if m:
        secs = float(m.group(1))
        res.update({"reference_cuda_cu_seconds": secs, "reference_cuda_cu_steps": ref_steps,
                    "reference_cuda_cu_rate": n * (n - 1.0) * ref_steps / secs,   # N(N-1)/2 pair evaluations credited as 2
                    "reference_cuda_cu_printed_positions_finite": f"{sum(finite)}/{len(finite)}",
                    "speedup_vs_reference_cuda_cu": (n * (n - 1.0) * steps / ours_secs) / (n * (n - 1.0) * ref_steps / secs)})
    else:
        res["reference_cuda_cu"] = "failed: " + (r.stderr or r.stdout)[-300:]
out["config1_n65536"] = res
print(json.dumps(out, indent=1))
