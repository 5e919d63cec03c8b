"""Turns gpurun_out/{launches.csv,prof.ncu-rep} into the small tracked summaries under profiles/."""
import csv
import json
import os
import subprocess
import sys
from collections import defaultdict

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
out_dir = os.path.join(REPO, "profiles")
src = os.path.join(REPO, "gpurun_out")

KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]


def launches():
    path = os.path.join(src, "launches.csv")
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    d = defaultdict(list)
    for r in rows[1:]:
        d[r[ki].split("(")[0]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in d.values())
    with open(os.path.join(out_dir, f"{tag}_launches_summary.csv"), "w") as f:
        f.write("kernel,launches,total_ms,share_pct,avg_ms\n")
        for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"\"{k}\",{len(v)},{sum(v) / 1e6:.4f},{100 * sum(v) / tot:.3f},{sum(v) / len(v) / 1e6:.4f}\n")
    # the raw per-launch list as well (small)
    with open(os.path.join(out_dir, f"{tag}_launches.csv"), "w") as f:
        f.write("id,kernel,grid,block,duration_ns\n")
        bi = hdr.index("Block Size")
        for r in rows[1:]:
            f.write(f"{r[0]},\"{r[ki].split('(')[0]}\",\"{r[gi]}\",\"{r[bi]}\",{r[vi]}\n")


def full():
    rep = os.path.join(src, "prof.ncu-rep")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    summary = []
    for rec in rows[2:]:
        m = {h: (u, v) for h, u, v in zip(hdr, units, rec)}
        item = {"kernel": m.get("Kernel Name", ("", ""))[1]}
        for k in KEYS:
            if k in m:
                item[k] = {"unit": m[k][0], "value": m[k][1]}
        summary.append(item)
    json.dump(summary, open(os.path.join(out_dir, f"{tag}_force_kernel_ncu_full.json"), "w"), indent=1)
    one = summary[0]
    tb = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
    rd = float(one["dram__bytes_read.sum"]["value"].replace(",", "")) * tb[one["dram__bytes_read.sum"]["unit"]]
    wr = float(one["dram__bytes_write.sum"]["value"].replace(",", "")) * tb[one["dram__bytes_write.sum"]["unit"]]
    json.dump({"kernel": one["kernel"], "workload": "plummer_n1048576", "dram_traffic_bytes_per_launch": rd + wr,
               "source": f"profiles/{tag}_force_kernel_ncu_full.json (ncu --set full, 1 launch)"},
              open(os.path.join(out_dir, "force_kernel_traffic.json"), "w"), indent=1)
    # per-instruction hot spots
    srcp = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    open(os.path.join(out_dir, f"{tag}_force_kernel_source_page.csv"), "w").write(srcp)


if __name__ == "__main__":
    os.makedirs(out_dir, exist_ok=True)
    launches()
    full()
    print("profiles written")
