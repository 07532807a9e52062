#!/usr/bin/env python
"""Prints the BASELINE.md stage table (markdown) from a bench.py JSON line:  python tools/results_table.py profiles/rXX_bench.json"""
import json
import sys

d = json.load(open(sys.argv[1]))
print("value %.0f pictures/s (%.2f ms per step), e2e %.0f pictures/s, launches %d, clocks %s" %
      (d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], d.get("clocks")))
cb = d.get("cpu_baseline")
if cb:
    print("cpu arm %.0f pictures/s on %d cores (%s)" % (cb["value"], cb["cores"], cb["kind"]))
print("| stage | launches / segment | ms / segment | share | avg launch | GB/s | of measured peak |")
print("|---|---|---|---|---|---|---|")
for k, v in d["stages"].items():
    g = v.get("achieved_gbs")
    print("| %s | %d | %.3f | %.1f %% | %.1f us | %s | %s |" % (
        k, round(v["launches_per_segment"]), v["ms_per_segment"], 100 * v["share_of_kernel_time"], 1e3 * v["avg_launch_ms"],
        "%.0f" % g if g else "–", "%.2f %%" % (100 * v["frac_of_peak"]) if g else "–"))
