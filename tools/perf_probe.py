"""Exploration tool (not the bench): wall-clock rate of gs_step per kernel variant / mode / N."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

gs = load_package()
PEAK = 148 * 128 * 2 * 1.965e9


def probe(n, variant, eps, steps=4, grid=None, flags=0):
    b = gs.ic_plummer(n, seed=1) if eps > 0 else gs.ic_reference_uniform(n, seed=1)
    if grid:
        os.environ["GRAVSIM_GRID"] = str(grid)
    else:
        os.environ.pop("GRAVSIM_GRID", None)
    with gs.GravSim(n, dt=1e-3 if eps > 0 else 3600.0, G=1.0 if eps > 0 else gs.G_SI, eps=eps, variant=variant, flags=flags) as sim:
        sim.upload(b)
        sim.step(2)
        sim.sync()
        t0 = time.perf_counter()
        sim.step(steps)
        sim.sync()
        dt = (time.perf_counter() - t0) / steps
        info = sim.info
    rate = n * (n - 1.0) / dt
    print(f"N={n:8d} variant={variant:2d} thr={info.threads} itile={info.i_tile} mass={'general' if (flags & 4 or eps == 0) else 'uniform'}{' SYM' if info.symmetric else ''} mode={'soft' if eps > 0 else 'cut '} grid={info.grid} regs={info.regs_per_thread} "
          f"{dt * 1e3:9.3f} ms/step  {rate:.3e} int/s  {rate * 20 / PEAK * 100:5.1f}% of 74.5 TF", flush=True)


if __name__ == "__main__":
    ns = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [262144]
    variants = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 2, 3, 4, 5, 6]
    for n in ns:
        for v in variants:
            probe(n, v, 0.01, flags=4 | 128)   # row gather, general masses, softened
            probe(n, v, 0.01, flags=128)       # row gather, equal masses (Plummer): hoisted-mass kernel
            if len(sys.argv) > 3 and 'cut' in sys.argv[3]:
                probe(n, v, 0.0, flags=128)  # row gather, cutoff mode (reference-uniform SI bodies)
            if len(sys.argv) > 3 and 'sym' in sys.argv[3]:
                probe(n, v, 0.01, flags=4)   # pair-symmetric pass, general masses
                probe(n, v, 0.01, flags=0)       # pair-symmetric pass, equal masses
                probe(n, v, 0.0, flags=0)        # pair-symmetric pass, cutoff mode
