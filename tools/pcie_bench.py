#!/usr/bin/env python
"""Host<->device copy ceiling of this box, for the `e2e` line of bench.py (one step = 64 MB up + 64 MB down):
pinned 64 MB buffers, H2D alone, D2H alone, both directions at once on two streams, and both at once while a
config-2-sized SpMV keeps the device memory busy.  Prints one JSON line.

    python tools/pcie_bench.py [--mb 64] [--reps 20]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=64)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import torch
    n = a.mb << 20
    dev = torch.device("cuda:0")
    h_up = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_dn = torch.empty(n, dtype=torch.uint8).pin_memory()
    d_up = torch.empty(n, dtype=torch.uint8, device=dev)
    d_dn = torch.empty(n, dtype=torch.uint8, device=dev)
    s_up, s_dn, s_k = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    big_a = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    big_b = torch.empty(1 << 30, dtype=torch.uint8, device=dev)

    def timed(up, dn, busy):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s_up.wait_event(e0), s_dn.wait_event(e0), s_k.wait_event(e0)
        for _ in range(a.reps):
            if up:
                with torch.cuda.stream(s_up):
                    d_up.copy_(h_up, non_blocking=True)
            if dn:
                with torch.cuda.stream(s_dn):
                    h_dn.copy_(d_dn, non_blocking=True)
            if busy:
                with torch.cuda.stream(s_k):
                    big_b.copy_(big_a)  # 2 GB of device traffic per step, like one config-2 SpMV
        for s in (s_up, s_dn, s_k):
            ev = torch.cuda.Event()
            ev.record(s)
            torch.cuda.current_stream().wait_event(ev)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        return {"ms_per_step": ms, "gbs_per_direction": n / ms / 1e6}

    for _ in range(2):
        timed(True, True, False)
    out = {"mb": a.mb, "h2d_alone": timed(True, False, False), "d2h_alone": timed(False, True, False),
           "both": timed(True, True, False), "both_with_device_traffic": timed(True, True, True)}
    try:
        q = subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current",
                            "--format=csv,noheader", "-i", "0"], capture_output=True, text=True, timeout=20).stdout.strip()
        out["pcie_link"] = q
    except Exception as e:  # noqa: BLE001
        out["pcie_link"] = f"unavailable: {e}"
    print(json.dumps(out))


if __name__ == "__main__":
    main()
