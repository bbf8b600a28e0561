"""Per-kernel metric table + DRAM traffic json from an `ncu --set full` report.
usage: python tools/ncu_summary.py report.ncu-rep out_summary.txt out_traffic.json"""
import csv
import io
import json
import subprocess
import sys

rep, out_txt, out_json = sys.argv[1:4]
mets = ("gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,"
        "dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,"
        "sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,"
        "smsp__thread_inst_executed_per_inst_executed.ratio,launch__registers_per_thread,"
        "launch__occupancy_limit_registers,launch__occupancy_limit_shared_mem,l1tex__t_sector_hit_rate.pct,"
        "lts__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", mets],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, u = rows[0], rows[1]
cols = [i for i, a in enumerate(h) if a in mets.split(",")]
lines = ["ncu --set full --import-source on --clock-control none, one launch of each hot kernel of the CSV row pipeline (round 2)",
         "workload: bench.py C2 (1 M rows x 8 cols, 86.8 MB in -> 135.1 MB out), B200; report %s (not tracked)" % rep,
         "times under ncu are cold-cache / serialised; bench.py's CUDA-event times are the measured ones", ""]
traffic = {"source": "%s (ncu --set full, C2 1M rows, round-2 kernels): dram__bytes_read.sum + "
                     "dram__bytes_write.sum per launch" % out_txt}
mult = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}
seen = set()
for r in rows[2:]:
    name = r[h.index("Kernel Name")].split("(")[0]
    if name in seen:
        continue
    seen.add(name)
    lines.append(name)
    d = {}
    for i in cols:
        lines.append("    %-62s %14s %s" % (h[i], r[i], u[i]))
        d[h[i]] = (float(r[i].replace(",", "")), u[i])
    traffic[name] = int(d["dram__bytes_read.sum"][0] * mult[d["dram__bytes_read.sum"][1]] +
                        d["dram__bytes_write.sum"][0] * mult[d["dram__bytes_write.sum"][1]])
    lines.append("")
open(out_txt, "w").write("\n".join(lines))
json.dump(traffic, open(out_json, "w"), indent=1)
print(json.dumps(traffic, indent=1))
