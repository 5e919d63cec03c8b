"""Merges the `bench.py --workload sweep` lines of P = 1, 2, 4, 8 (one file each) into profiles/r2_sweep.json:
per (N, P) interactions/s, both roofline fractions, and
  strong efficiency  T_1(N) / (P * T_P(N))                       (same N, P GPUs)
  weak efficiency    rate(2N, 4P') / (4 * rate(N, P'))            (N^2 / P fixed: N doubles when P quadruples)
usage: python tools/merge_sweep.py out.json sweep_p1.json sweep_p2.json ..."""
import json
import sys

rows = {}
meta = None
for path in sys.argv[2:]:
    line = [l for l in open(path) if l.startswith("{")][-1]
    d = json.loads(line)
    meta = meta or {k: d[k] for k in ("metric", "unit", "config")}
    for r in d["rows"]:
        rows[(r["n_bodies"], r["n_gpus"])] = r
for (n, p), r in rows.items():
    one = rows.get((n, 1))
    r["strong_efficiency"] = one["ms_per_step"] / (p * r["ms_per_step"]) if one else None
    base = rows.get((n // 2, p // 4)) if p % 4 == 0 else None
    r["weak_efficiency_vs_half_n_quarter_p"] = r["value"] / (4.0 * base["value"]) if base else None
out = dict(meta or {}, build="one build: the shipped libgravsim.so of this commit",
           rows=[rows[k] for k in sorted(rows)])
json.dump(out, open(sys.argv[1], "w"), indent=1)
print("N".rjust(9), *[f"P={p}: int/s   %peak  strong  weak".rjust(40) for p in (1, 2, 4, 8)])
for n in sorted({k[0] for k in rows}):
    cells = []
    for p in (1, 2, 4, 8):
        r = rows.get((n, p))
        if not r:
            cells.append(" " * 40)
            continue
        se = f"{r['strong_efficiency']:.3f}" if r["strong_efficiency"] else "  -  "
        we = f"{r['weak_efficiency_vs_half_n_quarter_p']:.3f}" if r["weak_efficiency_vs_half_n_quarter_p"] else "  -  "
        cells.append(f"{r['value']:.3e} {r['pct_fp32_peak_per_gpu']:6.1f}% {se} {we}".rjust(40))
    print(str(n).rjust(9), *cells)
