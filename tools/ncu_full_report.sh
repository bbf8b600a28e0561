#!/bin/bash
# profiles/<name>_ncu_full_summary.txt + traffic json from one `ncu --set full --import-source on` report of
# bench.py's C2 step: per-kernel metrics, DRAM traffic, source hot spots, stall reasons.
# usage: tools/ncu_full_report.sh gpurun_out/prof.ncu-rep profiles/r02_ncu_full_summary.txt profiles/traffic_r02.json
set -e
rep=$1; out=$2; traffic=$3
here=$(dirname "$0")
python "$here/ncu_summary.py" "$rep" "$out" "$traffic" > /dev/null
for k in csv2_scan_marks_kernel csv2_scan_write_kernel csv2_type_kernel csv2_emit_kernel; do
  echo "== $k: source hot spots (share of executed warp instructions, share of stall samples, active lanes)" >> "$out"
  python "$here/ncu_hotspots.py" "$rep" "$k" 14 >> "$out"
done
echo "== stall reasons (cycles a warp waits per issued instruction, top 8 per kernel)" >> "$out"
ncu -i "$rep" --page raw --csv 2>/dev/null | python -c "
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[0]; ki = hdr.index('Kernel Name')
cols = [i for i, h in enumerate(hdr) if 'issue_stalled' in h and h.endswith('.ratio') and 'not_issued' not in h and 'per_warp_active' not in h]
for r in rows[2:]:
    vals = sorted(((float(r[i].replace(',', '')) if r[i] else 0., hdr[i].split('issue_stalled_')[1].split('_per')[0]) for i in cols), reverse=True)
    print(r[ki].split('(')[0], ' '.join('%s %.2f' % (n, v) for v, n in vals[:8]))
" >> "$out"
