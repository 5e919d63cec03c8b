"""Generates tests/golden/*.npz from the REFERENCE ITSELF, compiled here from /root/reference by
oracle/Makefile (oracle/_ref/).  Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference ships no golden vectors of its own; these fixtures pin the
oracle restatement (tests/test_oracle.py, runs everywhere) and the CUDA path (tests -m gpu) to the
reference's arithmetic without needing /root/reference at test time.
"""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conftest import GOLDEN_DIR, Reference, load_package  # noqa: E402


def main():
    ref = Reference()
    assert ref.available, "oracle/_ref is missing: run `make -C oracle ref` where /root/reference exists"
    gs = load_package()
    with tempfile.TemporaryDirectory() as wd:
        # 1. Sun/Earth/Mars, the constants of cuda.cu:82-93 == mpi.c:76-94; 500 steps of 3600 s
        #    (cuda.cu:155,123; mpi.c:147-148), the reference's own main() with 1 rank (as written)
        #    and with 3 ranks (one body per rank == synchronous step).
        solar = gs.ic_reference_uniform(3, seed=0, dtype=np.float64)
        s1, _ = ref.run_main(solar, 3600.0, 500, 1, wd)
        s3, _ = ref.run_main(solar, 3600.0, 500, 3, wd)
        np.savez(os.path.join(GOLDEN_DIR, "solar3_500steps.npz"), initial=solar, final_np1=s1, final_np3=s3)

        # 2. frozen-snapshot accelerations from the reference's calculate_force (mpi.c:59-73,197-206)
        for n, seed in ((8, 11), (64, 12), (1000, 13)):
            p = gs.ic_reference_uniform(n, seed=seed, dtype=np.float64)
            np.savez(os.path.join(GOLDEN_DIR, f"accel_uniform_n{n}.npz"), bodies=p, accel=ref.accel(p), seed=seed)

        # 3. trajectories from the reference's own main(): N=256, 20 steps, np=1 and np=4 (as
        #    written, partition-dependent) and the synchronous step built from its functions.
        p = gs.ic_reference_uniform(256, seed=21, dtype=np.float64)
        t1, _ = ref.run_main(p, 3600.0, 20, 1, wd)
        t4, _ = ref.run_main(p, 3600.0, 20, 4, wd)
        ts = ref.step_sync(p, 3600.0, 20)
        np.savez(os.path.join(GOLDEN_DIR, "traj_uniform_n256_20steps.npz"), initial=p, final_np1=t1, final_np4=t4,
                 final_sync=ts)

        # 4. two coincident bodies + one distant: cutoff semantics (mpi.c:64)
        c = np.array([[1e9, 2e9, 3e9, 0, 0, 0, 1e24], [1e9, 2e9, 3e9, 0, 0, 0, 2e24], [5e10, 0, 0, 0, 0, 0, 3e24]])
        np.savez(os.path.join(GOLDEN_DIR, "coincident3.npz"), bodies=c, accel=ref.accel(c))
    print("golden fixtures written to", GOLDEN_DIR)


if __name__ == "__main__":
    main()
