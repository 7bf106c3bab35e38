#!/usr/bin/env python
"""Regenerates profiles/traffic_bytes_per_launch.json from an `ncu --set full` report (read on the CPU box).

    python tools/ncu_traffic.py gpurun_out/prof_spmv.ncu-rep [--local gpurun_out/prof_grid_stream.ncu-rep] [--raw-out profiles/rNN_x_raw.csv]

The file carries the kernel NAME the counters belong to and the git HEAD it was captured at; bench.py attaches
`roofline.traffic` only when the kernel that ran in the bench has the same name (a stale capture of another
schedule is never reported as this one's traffic)."""
import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    return out, hdr, rows[2:]  # row 1 = units


def summarise(rep):
    text, hdr, rows = raw_rows(rep)
    col = {n: i for i, n in enumerate(hdr)}
    r = rows[-1]  # the last captured launch (warm)
    f = lambda name: float(r[col[name]].replace(",", ""))
    unit_row = list(csv.reader(io.StringIO(text)))[1]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rd = f("dram__bytes_read.sum") * scale.get(unit_row[col["dram__bytes_read.sum"]], 1)
    wr = f("dram__bytes_write.sum") * scale.get(unit_row[col["dram__bytes_write.sum"]], 1)
    tu = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
    t = f("gpu__time_duration.sum") * tu.get(unit_row[col["gpu__time_duration.sum"]], 1.0)
    return text, {"kernel": r[col["Kernel Name"]], "dram_bytes_read": int(rd), "dram_bytes_write": int(wr),
                  "dram_bytes_per_launch": int(rd + wr), "gpu_time_us": t, "launches_captured": len(rows)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--local", default=None, help="report of the local-gathers (grid sweep) capture")
    ap.add_argument("--raw-out", default=None)
    ap.add_argument("--local-raw-out", default=None)
    a = ap.parse_args()
    head = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    text, top = summarise(a.report)
    if a.raw_out:
        open(a.raw_out, "w").write(text)
    out = dict(top, workload="config 2 (16M rows x 16 nnz/row, f32 plus-times)", algorithmic_bytes_per_launch=2348810244,
               git_head=head, source=f"{a.raw_out or a.report} (ncu --set full --clock-control none, tools/gpu_ncu.sh)")
    if a.local:
        ltext, loc = summarise(a.local)
        if a.local_raw_out:
            open(a.local_raw_out, "w").write(ltext)
        out["local_gathers"] = dict(loc, workload="config 4 operator (8192x8192 grid), one dense min-plus sweep",
                                    algorithmic_bytes_per_launch=2952527876, git_head=head,
                                    source=f"{a.local_raw_out or a.local} (ncu --set full --clock-control none)")
    else:
        try:
            out["local_gathers"] = json.load(open(os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json")))["local_gathers"]
        except Exception:
            pass
    json.dump(out, open(os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    sys.exit(main())
