#!/usr/bin/env python
"""Summarise `ncu --set full` reports (run here, on the CPU box: `ncu -i <rep> --page raw --csv`) into a small JSON that
is committed under profiles/ and read by bench.py for roofline.traffic.

    python tools/ncu_summary.py gpurun_out/prof_*.ncu-rep -o profiles/ncu_summary_latest.json
"""
import argparse
import csv
import io
import json
import subprocess
import sys

KEEP = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_active_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "launch__registers_per_thread": "registers",
    "smsp__issue_active.avg.per_cycle_active": "issue_active_per_cycle",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0, "second": 1.0,
        "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9}


def parse(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    res = []
    for r in rows[2:]:
        d = {"kernel": r[ki].split("(")[0].replace("void ", "").strip()}
        for i, h in enumerate(hdr):
            if h in KEEP and r[i] not in ("", "n/a"):
                try:
                    v = float(r[i].replace(",", ""))
                except ValueError:
                    continue
                if KEEP[h] in ("duration", "dram_read", "dram_write"):
                    v *= UNIT.get(units[i], 1.0)
                d[KEEP[h]] = v
        res.append(d)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("reports", nargs="+")
    ap.add_argument("-o", "--out", required=True)
    ap.add_argument("--note", default="")
    ap.add_argument("--capture", default="", help="what was captured (command line / round tag); stored for bench.py's traffic_source")
    a = ap.parse_args()
    launches = []
    for rep in a.reports:
        launches += parse(rep)
    agg = {}
    for d in launches:
        k = agg.setdefault(d["kernel"], {"launches": 0})
        k["launches"] += 1
        for key, v in d.items():
            if key == "kernel":
                continue
            k[key] = k.get(key, 0.0) + v
    for k in agg.values():
        n = k["launches"]
        for key in list(k):
            if key != "launches":
                k[key] = k[key] / n
        if "dram_read" in k and "dram_write" in k:
            k["dram_bytes_per_launch"] = k["dram_read"] + k["dram_write"]
            if k.get("duration"):
                k["dram_gbs_under_ncu"] = k["dram_bytes_per_launch"] / k["duration"] / 1e9
    import os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from solvingpapers_b200 import build as spbuild
    # lib_stamp: hash of the library sources at summary time.  bench.py prints roofline.traffic only while the loaded
    # library still has this stamp - run this script right after the capture, before touching csrc/.
    json.dump({"note": a.note or "per-launch averages from ncu --set full (cold-cache, serialised launches)", "lib_stamp": spbuild._stamp(),
               "capture": a.capture, "kernels": agg},
              open(a.out, "w"), indent=1, sort_keys=True)
    for name, k in sorted(agg.items(), key=lambda kv: -kv[1].get("duration", 0) * kv[1]["launches"]):
        print(f"{name[:60]:60s} n={k['launches']:2d} {k.get('duration', 0)*1e6:9.1f} us  dram {k.get('dram_bytes_per_launch', 0)/1e6:8.1f} MB  "
              f"tensor {k.get('tensor_pipe_active_pct', 0):5.1f}%  regs {int(k.get('registers', 0))}")


if __name__ == "__main__":
    main()
