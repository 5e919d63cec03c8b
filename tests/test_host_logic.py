"""CPU: the C-ABI library loads, exports what include/gravsim.h declares, and its host-side logic
(work partition, slices, initial conditions, state files, text dumps) is right.  No GPU compute."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import PKG_DIR, REPO


def test_library_exports_every_declared_symbol(gs):
    hdr = open(os.path.join(REPO, "include", "gravsim.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(gs.EXPORTED_SYMBOLS), declared ^ set(gs.EXPORTED_SYMBOLS)
    L = gs.lib()
    for s in declared:
        assert getattr(L, s) is not None
    assert b"sm_100a" in L.gs_version()
    assert L.gs_strerror(-6).decode().startswith("no CUDA device")


def test_no_cpu_fallback(gs):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(gs.GravSimError) as e:
        gs.GravSim(1024)
    assert e.value.code == -6          # GS_ENOGPU: the product never computes on the CPU


def test_product_does_not_reference_the_oracle():
    for root, _, files in os.walk(PKG_DIR):
        if os.sep + "build" in root:
            continue
        for f in files:
            if f.endswith((".cu", ".c", ".h", ".py", "Makefile")):
                txt = open(os.path.join(root, f), errors="ignore").read()
                assert "libgsoracle" not in txt and "oracle/" not in txt and "gso_" not in txt, f


@pytest.mark.parametrize("ni,nj,grid", [(1, 1, 1), (1, 3, 8), (7, 5, 3), (16, 16, 296), (1024, 2048, 296),
                                        (128, 256, 148), (3, 1000, 7), (50, 1, 296), (256, 1792, 296)])
def test_stream_k_partition_covers_every_unit_once(gs, ni, nj, grid):
    U = ni * nj
    ranges = [gs.plan_cta_range(ni, nj, grid, c) for c in range(grid)]
    assert ranges[0][0] == 0 and ranges[-1][1] == U
    for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0 and a0 <= a1
    sizes = [b - a for a, b in ranges]
    assert max(sizes) - min(sizes) <= 1                      # balanced to one unit
    # owners/split flags agree with the ranges; each CTA has at most two partial tiles
    partial_tiles = {c: set() for c in range(grid)}
    for it in range(ni):
        first, last, split = gs.plan_tile_owners(ni, nj, grid, it)
        lo, hi = it * nj, (it + 1) * nj
        touching = [c for c, (a, b) in enumerate(ranges) if a < b and a < hi and b > lo]
        assert touching[0] == first and touching[-1] == last
        assert split == (len(touching) > 1)
        if split:
            for c in touching:
                partial_tiles[c].add(it)
    for c, tiles in partial_tiles.items():
        assert len(tiles) <= 2
        if len(tiles) == 2:     # slot rule: first tile of the range -> slot 0, the other -> slot 1
            assert min(tiles) == ranges[c][0] // nj


@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("variant", [3, 6])
def test_symmetric_plan_covers_every_pair_once(gs, world, variant):
    # pair-symmetric pass: over all ranks every unordered pair of bodies must be evaluated exactly once.
    # At tile level: count[(i-tile row as j-tile range, j-tile)] -- plain units cover a row's own j-tiles in
    # one direction (both orders inside the row's bodies), symmetric units cover (row, other j-tile) once.
    i_tile, j_tile = gs.plan_tiles(variant)
    diag = i_tile // j_tile
    n = 4 * i_tile * world - 5                       # 4 rows per rank, a little padding
    n_pad = gs.plan_slice(n, world, 0, variant)[0]
    njt = n_pad // j_tile
    seen = np.zeros((njt, njt), dtype=np.int32)      # [j-tile a][j-tile b], a != b: symmetric coverage
    for rank in range(world):
        rows = gs.plan_sym_row(n, world, rank, variant, 0)[0]
        assert rows == n_pad // world // i_tile
        for row in range(rows):
            _, units, plain = gs.plan_sym_row(n, world, rank, variant, row)
            assert plain == diag
            own = [(rank * rows + row) * diag + d for d in range(diag)]      # the row's own j-tiles
            for k in range(units):
                jt = gs.plan_sym_unit(n, world, rank, variant, row, k)
                if k < plain:
                    assert jt == own[k]                                         # one-directional, all ordered pairs
                else:
                    assert jt not in own
                    for a in own:
                        seen[min(a, jt), max(a, jt)] += 1
    iu = np.triu_indices(njt, 1)
    same_row = (iu[0] // diag) == (iu[1] // diag)       # pairs of j-tiles inside one row are done by its plain units
    assert (seen[iu][~same_row] == 1).all() and (seen[iu][same_row] == 0).all()
    # and the CTA ranges of one rank tile its units exactly once
    grid = 37
    total = sum(gs.plan_sym_row(n, world, 0, variant, r)[1] for r in range(rows))
    pos, count = (0, 0), 0
    for cta in range(grid):
        row0, k0, cnt = gs.plan_sym_cta(n, world, 0, variant, grid, cta)
        if cnt:
            assert (row0, k0) == pos
            for _ in range(cnt):
                r, k = pos
                k += 1
                if k >= gs.plan_sym_row(n, world, 0, variant, r)[1]:
                    r, k = r + 1, 0
                pos = (r, k)
        count += cnt
    assert count == total


def test_slices_tile_aligned_and_cover(gs):
    for variant in (1, 3, 5, 6, 7, 10, 11, 12, 13, 16):
        it, jt = gs.plan_tiles(variant)
        for n, w in ((3, 1), (8, 4), (4096, 4), (65536, 1), (1048576, 8), (1000003, 2)):
            prev_end = 0
            for r in range(w):
                n_pad, b, e = gs.plan_slice(n, w, r, variant)
                assert b == prev_end and (e - b) % it == 0 and (e - b) % jt == 0 and n_pad >= n
                prev_end = e
            assert prev_end == n_pad
    assert gs.plan_slice(1048576, 8, 0, 1)[0] == 1048576      # the headline config needs no padding


def test_initial_conditions(gs):
    a = gs.ic_reference_uniform(1000, seed=7)
    b = gs.ic_reference_uniform(1000, seed=7, dtype=np.float64)
    assert np.array_equal(a.astype(np.float64), b)             # fp32-representable: same bodies both ways
    assert np.array_equal(a[:500], gs.ic_reference_uniform(500, seed=7))   # per-body streams
    assert not np.array_equal(a[3:], gs.ic_reference_uniform(1000, seed=8)[3:])
    # Sun / Earth / Mars: cuda.cu:82-93, mpi.c:76-94
    assert a[0, 6] == np.float32(1.989e30) and a[1, 0] == np.float32(1.496e11) and a[2, 4] == np.float32(24.077e3)
    assert np.abs(a[3:, :3]).max() <= 3e11 and np.abs(a[3:, 3:6]).max() <= 3e4
    assert a[3:, 6].min() >= 1e23 and a[3:, 6].max() <= 1e25
    assert abs(a[3:, 0].mean()) < 3e10 and abs(a[3:, 6].mean() / 5.05e24 - 1) < 0.05
    p = gs.ic_plummer(20000, seed=1)
    r = np.linalg.norm(p[:, :3], axis=1)
    assert abs(np.median(r) / 1.305 - 1) < 0.05 and r.max() <= 10.5      # half-mass radius of a=1 Plummer
    ke = (0.5 * p[:, 6] * (p[:, 3:6] ** 2).sum(1)).sum()
    assert abs(ke / (3 * np.pi / 64) - 1) < 0.05 and abs(p[:, 6].sum() - 1) < 1e-4
    g = gs.ic_two_galaxy(10000, seed=1, sep=8.0, impact=2.0, vrel=0.5)
    assert abs(g[:5000, 0].mean() + 4) < 0.05 and abs(g[5000:, 0].mean() - 4) < 0.05
    assert abs(g[:5000, 3].mean() - 0.25) < 0.01 and np.abs((g[:, 6:7] * g[:, 3:6]).sum(0)).max() < 1e-3


def test_state_file_round_trip_and_dumps(gs, tmp_path):
    a = gs.ic_reference_uniform(37, seed=3)
    path = str(tmp_path / "s.gs")
    gs.state_write(path, a, step=5, time=18000.0, dt=3600.0)
    b, h = gs.state_read(path)
    assert np.array_equal(a, b) and h.step == 5 and h.dt == 3600.0 and h.record_bytes == 28
    d = a.astype(np.float64)
    gs.state_write(path, d)
    b, h = gs.state_read(path)
    assert np.array_equal(d, b) and h.record_bytes == 56
    with open(path, "r+b") as f:
        f.write(b"XXXX")
    with pytest.raises(gs.GravSimError):
        gs.state_read(path)
    out = str(tmp_path / "dump.txt")
    gs.dump_positions_cuda_format(out, a, 10)
    lines = open(out).read().splitlines()
    assert lines[1] == "Final positions of first 10 out of 37 particles:"                 # cuda.cu:100
    assert lines[3] == "Particle 1: (%.14f, %.14f, %.14f)" % tuple(a[1, :3])               # cuda.cu:103-108
    gs.dump_positions_mpi_format(out, d)
    lines = open(out).read().splitlines()
    assert lines[1] == "Final positions:" and lines[3] == "Particle 1: (%e, %e, %e)" % tuple(d[1, :3])   # mpi.c:249-256


def test_cli_driver_built_and_fails_loudly_without_gpu(gs):
    exe = os.path.join(PKG_DIR, "bin", "gravsim")
    assert os.path.exists(exe)
    assert subprocess.run([exe], capture_output=True).returncode == 2
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([exe, "64", "1", "3600", "--quiet"], capture_output=True, text=True)
        assert r.returncode == 1 and "no CUDA device" in r.stderr


def test_bench_reference_arm_line_and_its_bodies(gs):
    """bench.py --impl reference: honest config (what it ran), zero GPU bytes, and bodies generated without the
    product library in the reference's own distribution (same Sun/Earth/Mars records as the library's generator)."""
    import importlib.util
    import json
    import sys
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(REPO, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    b = bench.reference_uniform_numpy(1000, seed=3)
    assert np.array_equal(b[:3].astype(np.float32), gs.ic_reference_uniform(3, seed=1))
    assert np.abs(b[3:, :3]).max() <= 3e11 and np.abs(b[3:, 3:6]).max() <= 3e4
    assert b[3:, 6].min() >= 1e23 and b[3:, 6].max() <= 1e25 and np.array_equal(b, b.astype(np.float32).astype(np.float64))
    if not os.path.exists(os.path.join(REPO, "oracle", "_ref", "mpi_ref")):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    r = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["config"]["workload"] == "reference_uniform_n32768_eps0_sample"
    assert line["config"]["extrapolated"] is True and line["value"] > 0 and line["unit"] == "interactions/s"
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "reference"


def test_merge_sweep_efficiencies(tmp_path):
    import json
    import sys
    rows = lambda p, scale: [{"n_bodies": n, "n_gpus": p, "ms_per_step": 1e-9 * n * n / (p * scale), "value": 1e12 * p * scale,
                              "pct_fp32_peak_per_gpu": 100.0 * scale} for n in (16384, 32768)]
    files = []
    for p, scale in ((1, 1.0), (4, 0.9)):
        f = tmp_path / f"s{p}.json"
        f.write_text(json.dumps({"metric": "m", "unit": "u", "config": {}, "rows": rows(p, scale)}) + "\n")
        files.append(str(f))
    out = tmp_path / "merged.json"
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", "merge_sweep.py"), str(out)] + files, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    m = {(x["n_bodies"], x["n_gpus"]): x for x in json.load(open(out))["rows"]}
    assert abs(m[(32768, 4)]["strong_efficiency"] - 0.9) < 1e-9
    assert abs(m[(32768, 4)]["weak_efficiency_vs_half_n_quarter_p"] - 0.9) < 1e-9
    assert m[(16384, 1)]["strong_efficiency"] == 1.0 and m[(16384, 4)]["weak_efficiency_vs_half_n_quarter_p"] is None
