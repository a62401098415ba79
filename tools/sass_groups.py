#!/usr/bin/env python
"""Group the per-line instruction counts of tools/sass_by_line.py by line ranges of forward_fx.cuh.
Usage: tools/sass_by_line.py ... 500 | tools/sass_groups.py name:lo-hi ..."""
import re, sys
ranges = []
for a in sys.argv[1:]:
    n, r = a.split(':'); lo, hi = r.split('-'); ranges.append((n, int(lo), int(hi)))
g, st = {}, {}
for l in sys.stdin:
    m = re.match(r'(\S+)\s*:\s*(\d+)\s+([\d.]+)M\s+([\d.]+)%\s+stall\s+([\d.]+)%', l)
    if not m:
        continue
    f, ln, n, s = m.group(1), int(m.group(2)), float(m.group(3)), float(m.group(5))
    k = 'other:' + f
    if f == 'forward_fx.cuh':
        k = 'fx:unassigned'
        for name, lo, hi in ranges:
            if lo <= ln <= hi:
                k = name
    g[k] = g.get(k, 0) + n; st[k] = st.get(k, 0) + s
for k in sorted(g, key=lambda k: -g[k]):
    print(f'{k:40s} {g[k]:8.1f}M  stall {st[k]:5.1f}%')
print('total', round(sum(g.values()), 1))
