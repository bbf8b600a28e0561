"""Source-level hot spots of one kernel in an .ncu-rep captured with --import-source on:
share of executed warp instructions, share of stall samples, active lanes per instruction.
usage: python tools/ncu_hotspots.py report.ncu-rep kernel_regex [top_n]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--kernel-name", "regex:" + kern, "--page", "source",
                      "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur, hdr, agg = None, None, {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if r[0] == "Function Name":
        continue
    if hdr and r[0].isdigit() and r[2] == "-":
        try:
            inst = int(r[hdr.index("Instructions Executed")])
            smp = int(r[hdr.index("# Samples")])
            ti = int(r[hdr.index("Thread Instructions Executed")])
        except Exception:
            continue
        a = agg.setdefault((cur, int(r[0]), r[1][:80]), [0, 0, 0])
        a[0] += inst; a[1] += smp; a[2] += ti
ti = sum(a[0] for a in agg.values()) or 1
ts = sum(a[1] for a in agg.values()) or 1
tt = sum(a[2] for a in agg.values())
print("tot warp inst", ti, "avg active %.1f" % (tt / ti), "samples", ts)
for k, a in sorted(agg.items(), key=lambda x: -x[1][1 if "--by-samples" in sys.argv else 0])[:int(sys.argv[3]) if len(sys.argv) > 3 else 30]:
    print("%5.1f%% inst %5.1f%% smp act %4.1f %-16s:%4d  %s" % (100 * a[0] / ti, 100 * a[1] / ts, a[2] / max(a[0], 1), k[0], k[1], k[2]))
