set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python - <<'PY'
import sys, time
sys.path.insert(0, '.')
from __graft_entry__ import load_package
gs = load_package()
for n, eps in ((1048576, 0.01), (1048576, 0.0), (65536, 0.01)):
    b = gs.ic_plummer(n, seed=1)
    with gs.GravSim(n, dt=1e-3, G=1.0, eps=eps) as sim:
        sim.upload(b); sim.energy()
        t0 = time.perf_counter(); e = sim.energy(); de = time.perf_counter() - t0
    print(f"N={n} eps={eps}: gs_energy {de*1e3:.2f} ms  PE={e[1]:.9f}", flush=True)
PY
