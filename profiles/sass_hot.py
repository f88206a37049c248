"""Per-SASS-instruction executed counts from `ncu --page source --csv` (gzip), normalised per warp-step.
usage: python profiles/sass_hot.py file.source.csv.gz <warp_steps> [min_per_step]"""
import csv, gzip, sys
rows = list(csv.reader(gzip.open(sys.argv[1], "rt")))
ws = float(sys.argv[2]); thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ci, cs, ct, cst = h.index("Instructions Executed"), h.index("Source"), h.index("Avg. Threads Executed"), h.index("# Samples")
tot = 0.0
for n, r in enumerate(rows[hi + 1:]):
    if len(r) <= ci: continue
    x = float(r[ci]) / ws
    tot += x
    if x >= thr:
        print("%4d %7.3f %5s %6s  %s" % (n, x, r[ct][:5], r[cst], r[cs].strip()))
print("total per warp-step: %.1f" % tot)
