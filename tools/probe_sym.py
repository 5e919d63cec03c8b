"""Exploration tool (not the bench): rate of gs_step for a list of (variant, eps, flags, env) set-ups of the
pair-symmetric pass at one N.  usage: python tools/probe_sym.py N 'variant:eps:flags[:ENV=VAL,...]' ..."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

gs = load_package()
PEAK = 148 * 128 * 2 * 1.965e9
KNOBS = ("GRAVSIM_SYM_NPAIR", "GRAVSIM_SYM_SHFL", "GRAVSIM_CUTMODE", "GRAVSIM_GRID")


def probe(n, variant, eps, flags, env, steps=None):
    steps = steps or max(4, int(2e12 / (float(n) * n)))
    for k in KNOBS:
        os.environ.pop(k, None)
    os.environ.update(env)
    b = gs.ic_plummer(n, seed=1) if eps > 0 else gs.ic_reference_uniform(n, seed=1)
    with gs.GravSim(n, dt=1e-3 if eps > 0 else 3600.0, G=1.0 if eps > 0 else gs.G_SI, eps=eps, variant=variant,
                    flags=flags) as sim:
        sim.upload(b)
        sim.step(2)
        sim.sync()
        t0 = time.perf_counter()
        sim.step(steps)
        sim.sync()
        dt = (time.perf_counter() - t0) / steps
        info = sim.info
    rate = n * (n - 1.0) / dt
    print(f"N={n} v={info.variant:2d} thr={info.threads} itile={info.i_tile} sym={info.symmetric} uni={info.uniform_mass} "
          f"eps={eps} flags={flags} env={env} regs={info.regs_per_thread} {dt * 1e3:9.3f} ms/step {rate:.4e} int/s "
          f"{rate * 20 / PEAK * 100:6.1f}%", flush=True)


if __name__ == "__main__":
    n = int(sys.argv[1])
    for spec in sys.argv[2:]:
        f = spec.split(":")
        env = dict(kv.split("=") for kv in f[3].split(",")) if len(f) > 3 and f[3] else {}
        try:
            probe(n, int(f[0]), float(f[1]), int(f[2]), env)
        except Exception as ex:      # noqa: BLE001
            print(f"spec {spec}: FAILED {ex}", flush=True)
