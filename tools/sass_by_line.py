#!/usr/bin/env python
"""Join ncu's per-SASS-instruction counters with nvdisasm line info: executed warp instructions and stall
samples per CUDA source line.  Usage: tools/sass_by_line.py report.ncu-rep cubin kernel-regex mangled-substring"""
import csv, io, re, subprocess, sys
from collections import defaultdict

rep, cubin, kre, mangled = sys.argv[1:5]
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[2:]:
    if len(r) < len(hdr) or r[0] in ('Kernel Name', 'Address'):
        break
    data.append((r[ix['Source']].strip(), int(r[ix['Instructions Executed']]), int(r[ix['# Samples']])))
dis = subprocess.run(['nvdisasm', '--print-line-info', '-c', cubin], capture_output=True, text=True).stdout
# find the function section
lines = dis.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('.text.') and mangled in l)
cur = ('?', 0)
insts = []
for l in lines[start + 1:]:
    if l.startswith('.text.') or l.startswith('.section'):
        if insts:
            break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        inl = ' (inlined)' if 'inlined' in l else ''
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*?);', l)
    if m:
        insts.append((cur, m.group(2).strip()))
print('ncu instructions', len(data), 'nvdisasm instructions', len(insts))
by = defaultdict(lambda: [0, 0])
for (srcl, n, s), (loc, txt) in zip(data, insts):
    by[loc][0] += n
    by[loc][1] += s
ti = sum(v[0] for v in by.values()); ts = sum(v[1] for v in by.values())
cache = {}
def text(loc):
    f, ln = loc
    for base in ('/root/repo/syn3r_b200/csrc/', ):
        try:
            if f not in cache:
                cache[f] = open(base + f).read().splitlines()
            return cache[f][ln - 1].strip()[:90]
        except Exception:
            pass
    return ''
for loc, (n, s) in sorted(by.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[5]) if len(sys.argv) > 5 else 60]:
    print(f'{loc[0]:22s}:{loc[1]:4d} {n / 1e6:8.2f}M {n / ti * 100:5.1f}%  stall {s / ts * 100:5.1f}%  | {text(loc)}')
