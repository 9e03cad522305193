#!/usr/bin/env python
"""tools/ncu_table.py PLAN.json NCU.csv > profiles/rN_ncu_per_kernel_dram.md

Joins the plan printed by tools/ncu_each_kernel.py (families in launch order, kernels per call, algorithmic bytes) with
the CSV of `ncu --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,
dram__throughput.avg.pct_of_peak_sustained_elapsed` over the same command: per family, the LAST call's kernels."""
import csv
import json
import sys
from collections import OrderedDict


def main():
    plan = json.loads([l for l in open(sys.argv[1]) if l.startswith("[")][-1])
    rows = OrderedDict()
    with open(sys.argv[2], newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if not any(k in r["Kernel Name"] for k in ("mt_table_kernel", "norm_finalize", "norm_blend", "seg_sq_finalize")):
            continue
        d = rows.setdefault(r["ID"], {"name": r["Kernel Name"]})
        val = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        m = r["Metric Name"]
        if m.startswith("dram__bytes"):
            val *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        if m == "gpu__time_duration.sum":
            val *= {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)      # -> microseconds
        d[m] = val
    kernels = list(rows.values())
    print("| family | kernels | algorithmic MB | DRAM MB (read+write) | DRAM/alg | time us | GB/s (alg) | dram throughput % of peak |")
    print("|---|---|---|---|---|---|---|---|")
    i = 0
    for fam in plan:
        k = fam["kernels_per_call"]
        n = k * fam["launches"]
        last = kernels[i + n - k:i + n]
        i += n
        if len(last) < k or k == 0:
            print(f"| {fam['family']} | {k} | (kernels missing from the capture) |")
            continue
        dram = sum(x.get("dram__bytes_read.sum", 0) + x.get("dram__bytes_write.sum", 0) for x in last)
        t = sum(x.get("gpu__time_duration.sum", 0) for x in last)
        pct = sum(x.get("dram__throughput.avg.pct_of_peak_sustained_elapsed", 0) * x.get("gpu__time_duration.sum", 0)
                  for x in last) / max(t, 1e-9)
        alg = fam["algorithmic_bytes_per_launch"]
        print(f"| {fam['family']} | {k} | {alg / 1e6:.0f} | {dram / 1e6:.0f} | {dram / alg:.3f} | {t:.1f} | "
              f"{alg / t / 1e3:.0f} | {pct:.1f} |")


if __name__ == "__main__":
    main()
