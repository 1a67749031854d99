#!/usr/bin/env python
"""Print the LAST `n` launches of an ncu --csv launch list (kernel, grid, us)."""
import csv, sys
path = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 20
skip = sys.argv[3] if len(sys.argv) > 3 else None
rows = [r for r in csv.reader(open(path)) if len(r) > 10]
h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value'); gi = h.index('Grid Size')
rows = [r for r in rows[1:] if not (skip and skip in r[ki])]
for r in rows[-n:]:
    print(f"{r[ki][:60]:60s} {r[gi]:14s} {float(r[vi].replace(',', '')) / 1e3:9.2f} us")
