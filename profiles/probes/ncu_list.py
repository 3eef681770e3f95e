"""Print (id, kernel, grid, duration ns) from an `ncu --metrics gpu__time_duration.sum --csv` log; optional id range."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
cols = rows[hdr]
ki, vi, gi = cols.index("Kernel Name"), cols.index("Metric Value"), cols.index("Grid Size")
lo, hi = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (0, 10 ** 9)
for r in rows[hdr + 1:]:
    if len(r) > vi and lo <= int(r[0]) <= hi:
        print(r[0], r[ki][:70].replace("<unnamed>::", ""), r[gi], r[vi])
