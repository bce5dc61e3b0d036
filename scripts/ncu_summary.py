#!/usr/bin/env python
"""Markdown table + roofline_traffic.json from `ncu --set full` reports (read with `ncu -i … --page raw --csv`).

usage: ncu_summary.py <tag> <gemm.ncu-rep> [<attention.ncu-rep>] [--workload config2]
       -> stdout = markdown; updates key gemm_bytes_per_launch_<workload> of profiles/roofline_traffic.json
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = [("time us", "gpu__time_duration.sum", 1.0), ("tensor pipe active % (elapsed)", "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed", 1.0),
        ("tensor active % (while SM active)", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
        ("issue active %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
        ("DRAM read MB", "dram__bytes_read.sum", None), ("DRAM write MB", "dram__bytes_write.sum", None),
        ("L2 hit %", "lts__t_sector_hit_rate.pct", 1.0), ("L2->SM % of peak", "lts__t_sectors_srcunit_tex.avg.pct_of_peak_sustained_elapsed", 1.0)]
UNIT = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "ns": 1e-3, "us": 1.0, "ms": 1e3}


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    hdr, units = r[0], r[1]
    return hdr, units, r[2:]


def val(hdr, units, row, name):
    if name not in hdr:
        return None
    i = hdr.index(name)
    try:
        v = float(row[i].replace(",", ""))
    except ValueError:
        return None
    return v * UNIT.get(units[i], 1.0)


def short(n):
    n = re.sub(r"void |ace::|<unnamed>::|\(anonymous namespace\)::", "", n)
    return re.sub(r"\(CUtensorMap.*", "", n)


def table(rep):
    hdr, units, rows = rows_of(rep)
    lines = ["| # | kernel | grid | " + " | ".join(c[0] for c in COLS) + " |", "|---|---|---|" + "---|" * len(COLS)]
    traffic = []
    for k, row in enumerate(rows):
        name = short(row[hdr.index("Kernel Name")])
        grid = row[hdr.index("launch__grid_size")] if "launch__grid_size" in hdr else "?"
        vals = [val(hdr, units, row, c[1]) for c in COLS]
        lines.append(f"| {k} | `{name}` | {grid} | " + " | ".join("-" if v is None else f"{v:.1f}" for v in vals) + " |")
        rd, wr = val(hdr, units, row, "dram__bytes_read.sum"), val(hdr, units, row, "dram__bytes_write.sum")
        if rd is not None and wr is not None:
            traffic.append((rd + wr) * 1e6)
    return "\n".join(lines), traffic


def main():
    workload = "config2"
    if "--workload" in sys.argv:
        i = sys.argv.index("--workload")
        workload = sys.argv[i + 1]
        del sys.argv[i:i + 2]
    tag, gemm_rep = sys.argv[1], sys.argv[2]
    t, traffic = table(gemm_rep)
    print(f"## tap-GEMM launches ({os.path.basename(gemm_rep)})\n\n{t}\n")
    if traffic:
        mean = sum(traffic) / len(traffic)
        print(f"Mean DRAM traffic per launch: {mean / 1e6:.1f} MB over {len(traffic)} launches.\n")
        path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        try:
            doc = json.load(open(path))
        except Exception:
            doc = {}
        doc[f"gemm_bytes_per_launch_{workload}"] = mean
        doc[f"source_{workload}"] = (f"profiles/{tag}.md: mean of dram__bytes_read.sum + dram__bytes_write.sum over {len(traffic)} consecutive gemm_tc* launches of "
                                     f"`bench.py --workload {workload} --steps 2 --warmup 3` (ncu --set full)")
        json.dump(doc, open(path, "w"), indent=1)
    if len(sys.argv) > 3:
        t, _ = table(sys.argv[3])
        print(f"## attention launches ({os.path.basename(sys.argv[3])})\n\n{t}\n")


if __name__ == "__main__":
    main()
