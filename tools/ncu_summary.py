#!/usr/bin/env python3
"""Summarise the ncu artefacts of tools/profile.sh into small text files for profiles/.

    python tools/ncu_summary.py <tag>      (reads gpurun_out/<tag>_launches.csv and
                                            gpurun_out/<tag>_k_query.ncu-rep, needs `ncu` on PATH)
Writes profiles/<tag>_launches.txt (per-kernel share of the profiled command),
profiles/<tag>_k_query_metrics.txt (headline counters of the hot kernel),
profiles/<tag>_k_query_hot_lines.txt (per-source-line samples / instructions / active lanes) and
updates profiles/traffic.json (DRAM bytes per launch of k_query, read by bench.py)."""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_static", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct", "smsp__inst_executed.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max"]


def to_ns(v, unit):
    v = float(v.replace(",", ""))
    return v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)


def launches(tag):
    path = os.path.join(ROOT, "gpurun_out", f"{tag}_launches.csv")
    if not os.path.exists(path):
        return
    lines = [l for l in open(path) if l.startswith('"')]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(io.StringIO("".join(lines))):
        a = agg[row["Kernel Name"][:90]]
        a[0] += 1
        a[1] += to_ns(row["Metric Value"], row["Metric Unit"])
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(ROOT, "profiles", f"{tag}_launches.txt"), "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none (tools/profile.sh {tag}); cold-cache, serialised: compare shares\n")
        f.write(f"# total {tot / 1e6:.3f} ms over {sum(v[0] for v in agg.values())} launches\n")
        for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write(f"{v[0]:5d} launches {v[1] / 1e6:10.3f} ms {100 * v[1] / tot:5.1f} %  {k}\n")


def metrics(tag, n_queries):
    rep = os.path.join(ROOT, "gpurun_out", f"{tag}_k_query.ncu-rep")
    if not os.path.exists(rep):
        return
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    traffic = []
    with open(os.path.join(ROOT, "profiles", f"{tag}_k_query_metrics.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none -k regex:k_query (tools/profile.sh {tag}); values per launch\n")
        for row in rows[2:]:
            f.write(row[hdr.index("Kernel Name")] + "\n")
            d = {}
            for k in KEYS:
                if k in hdr:
                    d[k] = row[hdr.index(k)]
                    f.write(f"  {k:70s} {row[hdr.index(k)]} {units[hdr.index(k)]}\n")
            def gb(k):
                i = hdr.index(k)
                return float(row[i].replace(",", "")) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[units[i]]
            t = gb("dram__bytes_read.sum") + gb("dram__bytes_write.sum")
            traffic.append(t)
            extra = {k: float(d[k].replace(",", "")) for k in ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
                                                                 "l1tex__t_sector_hit_rate.pct", "smsp__thread_inst_executed_per_inst_executed.ratio",
                                                                 "sm__warps_active.avg.pct_of_peak_sustained_active",
                                                                 "dram__throughput.avg.pct_of_peak_sustained_elapsed") if k in d}
            if n_queries:
                f.write(f"  -> DRAM traffic {t / 1e9:.2f} GB per launch = {t / n_queries:.0f} B/query at {n_queries} queries per launch\n")
    if traffic:
        json.dump({"k_query_dram_bytes_per_launch": sum(traffic) / len(traffic), "queries_per_launch": n_queries, "source": f"{tag}_k_query.ncu-rep",
                   "counters": extra},
                  open(os.path.join(ROOT, "profiles", "traffic.json"), "w"))
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    hdr, out, seen = None, [], 0
    for line in src.splitlines():
        # not csv.reader: source lines with inline asm carry unescaped quotes; split on the
        # separator and re-join whatever the source column was broken into
        if not line.startswith('"'):
            continue
        r = line.strip()[1:-1].split('","')
        if r[0] == "Line No":
            hdr = r
            seen += 1
            continue
        if hdr is None or seen > 1 or not r[0].isdigit():
            continue
        extra = len(r) - len(hdr)
        if extra > 0:
            r = [r[0], '","'.join(r[1:2 + extra])] + r[2 + extra:]
        out.append((int(r[0]), r[1], int(r[6]), int(r[7]), int(r[8]), r[hdr.index("stall_long_sb")]))
    ts, ti = sum(o[2] for o in out) or 1, sum(o[3] for o in out) or 1
    with open(os.path.join(ROOT, "profiles", f"{tag}_k_query_hot_lines.txt"), "w") as f:
        f.write(f"# first k_query launch of {tag}_k_query.ncu-rep; samples {ts}, warp instructions {ti}\n")
        f.write("# line  samples%  instr%  active lanes/instr  long-scoreboard samples | source\n")
        for o in sorted(out, key=lambda x: -x[2])[:40]:
            f.write(f"{o[0]:5d} {100 * o[2] / ts:6.1f} {100 * o[3] / ti:6.1f} {o[4] / max(o[3], 1):6.1f} {o[5]:>8} | {o[1].strip()[:110]}\n")


if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    nq = int(float(sys.argv[2])) if len(sys.argv) > 2 else 1 << 26
    launches(tag)
    metrics(tag, nq)
