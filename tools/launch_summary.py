"""Prints the kernels of ONE step (between two dgs_header_init_kernel launches) from an ncu launch-list CSV."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
seq = [(r[ki].split("(")[0][-60:], float(r[vi].replace(",", ""))) for r in rows[hi + 2:] if len(r) > vi]
idx = [i for i, (k, v) in enumerate(seq) if "header_init" in k]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 1
a, b = idx[which], idx[which + 1]
tot = 0.0
for k, v in seq[a:b]:
    print(f"{v / 1000:9.1f} us  {k}")
    tot += v
print(f"total {tot / 1000:.1f} us in {b - a} launches")
