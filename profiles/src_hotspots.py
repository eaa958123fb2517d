"""Aggregate an `ncu --page source --print-source cuda,sass --csv` dump by source file:line.
usage: python profiles/src_hotspots.py dump.csv [top]"""
import csv
import sys

def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file, hdr, out = None, None, []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and r and r[0].strip().isdigit():
        d = dict(zip(hdr, r))
        out.append((cur_file, int(r[0]), r[1].strip()[:90], num(d["# Samples"]), num(d["Instructions Executed"])))
tot_s = sum(o[3] for o in out) or 1
tot_i = sum(o[4] for o in out) or 1
print(f"total samples {tot_s}  total warp-instructions {tot_i}")
for o in sorted(out, key=lambda o: -o[3])[:top]:
    print(f"{100*o[3]/tot_s:5.1f}% smp {100*o[4]/tot_i:5.1f}% inst  {o[0]}:{o[1]}  {o[2]}")

# optional: aggregate by named line ranges of srbd_solve_warp.cuh given as name:lo-hi,...
if len(sys.argv) > 3:
    print()
    for spec in sys.argv[3].split(","):
        name, rng = spec.split(":")
        lo, hi = [int(x) for x in rng.split("-")]
        sel = [o for o in out if o[0] == "srbd_solve_warp.cuh" and lo <= o[1] <= hi]
        print(f"{name:14s} {100*sum(o[3] for o in sel)/tot_s:5.1f}% smp {100*sum(o[4] for o in sel)/tot_i:5.1f}% inst")
    oth = [o for o in out if o[0] != "srbd_solve_warp.cuh"]
    print(f"{'other files':14s} {100*sum(o[3] for o in oth)/tot_s:5.1f}% smp {100*sum(o[4] for o in oth)/tot_i:5.1f}% inst")
