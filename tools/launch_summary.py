"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total,
time per pipeline run and share.  usage: python tools/launch_summary.py launches.csv [bench.json]"""
import collections
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[hdr + 2:]:
    if len(r) > vi:
        k = r[ki].split("(")[0].replace("void ", "")
        tot[k] += float(r[vi].replace(",", ""))
        cnt[k] += 1
runs = max(cnt.get("rfq::csv2_scan_marks_kernel", cnt.get("rfq::csv2_scan_kernel", cnt.get("rfq::csv_tilefn_kernel", 1))), 1)
T = sum(v for k, v in tot.items() if k.startswith("rfq::"))
print("%d launches, %d pipeline runs (warm-up, timed, parity gate, e2e)" % (len(rows) - hdr - 2, runs))
print("per-launch times are cold-cache and serialised under ncu: compare SHARES with bench.py's "
      "kernel_ms_per_step, not absolutes\n")
print("%-46s %5s %10s %9s %7s" % ("kernel", "n", "total us", "us/run", "share"))
for k, v in tot.most_common():
    if k.startswith("rfq::"):
        print("%-46s %5d %10.1f %9.1f %6.1f%%" % (k[:46], cnt[k], v / 1000, v / 1000 / runs, 100 * v / T))
if len(sys.argv) > 2:
    j = json.load(open(sys.argv[2]))
    km = {k: v for k, v in j["kernel_ms_per_step"].items() if k.endswith("_kernel")}
    print("\nbench.py (CUDA events, no profiler), %.3f ms per step:" % j["ms_per_step"],
          ", ".join("%s %.1f %%" % (k.replace("csv_", "").replace("_kernel", ""), 100 * v / j["ms_per_step"])
                    for k, v in sorted(km.items(), key=lambda x: -x[1])[:6]))
