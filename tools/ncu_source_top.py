#!/usr/bin/env python
"""Top stalled SASS instructions of a kernel from `ncu -i rep --page source --csv`."""
import csv
import subprocess
import sys


def main(path, top=25, kernel_idx=0):
    out = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
    lines = out.splitlines()
    # split per kernel: each block starts with "Kernel Name",...
    blocks, cur = [], []
    for ln in lines:
        if ln.startswith('"Kernel Name"'):
            if cur:
                blocks.append(cur)
            cur = [ln]
        elif cur:
            cur.append(ln)
    if cur:
        blocks.append(cur)
    for b in blocks:
        print('=' * 20, b[0][:150])
        rows = list(csv.reader(b[1:]))
        h = rows[0]
        si = h.index('# Samples')
        src = h.index('Source')
        stall_cols = [i for i, k in enumerate(h) if k.startswith('stall_') and 'Not Issued' not in k]
        data = []
        total = 0
        for r in rows[1:]:
            try:
                n = int(r[si])
            except Exception:
                continue
            total += n
            data.append((n, r))
        print('total samples', total)
        tot_st = {h[i]: 0 for i in stall_cols}
        for n, r in data:
            for i in stall_cols:
                try:
                    tot_st[h[i]] += int(r[i])
                except Exception:
                    pass
        print('stall totals:', {k: v for k, v in sorted(tot_st.items(), key=lambda kv: -kv[1]) if v})
        order = sorted(range(len(data)), key=lambda i: -data[i][0])[:top]
        for i in sorted(order):
            n, r = data[i]
            st = {h[c][6:]: int(r[c]) for c in stall_cols if r[c] not in ('', '0')}
            st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
            print(f'{i:5d} {n:6d} {r[src][:90]:90s} {st}')


if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
