"""Transpose `ncu --page raw --csv` of one kernel launch into metric,unit,value rows (what profiles/r1_*_ncu_full_summary.csv hold):
    python profiles/ncu_summary.py gpurun_out/<tag>/solve_full_raw.csv [earlier_summary.csv] > profiles/r1_<tag>_ncu_full_summary.csv"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if r]
hdr, units, vals = rows[0], rows[1], rows[-1]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit", "value"])
skip = {"ID", "Process ID", "Process Name", "Host Name", "Context", "Stream", "Device", "CC", "Section Name", "Metric Name"}
pick = None
if len(sys.argv) > 2:           # optional: keep only the metrics listed in an earlier summary
    pick = {r[0] for r in csv.reader(open(sys.argv[2]))}
for h, u, v in zip(hdr, units, vals):
    if h in skip or (pick is not None and h not in pick):
        continue
    w.writerow([h, u, v])
