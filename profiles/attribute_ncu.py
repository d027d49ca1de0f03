#!/usr/bin/env python
"""Attributes executed instructions of one kernel to source functions.
usage: attribute_ncu.py <ncu source-page csv> <nvdisasm -g -c listing> <mangled kernel name prefix>
The csv comes from `ncu -i X.ncu-rep --page source --csv --launch-count 1`, the listing from
`cuobjdump -xelf all librrt_b200.so; nvdisasm -g -c rrt_api.sm_100a.cubin`."""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1])))
txt = open(sys.argv[2]).read().split('\n')
name = sys.argv[3]
h = rows[1]
start = [i for i, l in enumerate(txt) if l.startswith('\t.section\t.text.' + name)][0]
cur = None; m = {}
for l in txt[start + 1:]:
    if (l.startswith('\t.section') or l.startswith('//-----')) and m:
        break
    mm = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if mm:
        cur = (mm.group(1).split('/')[-1], int(mm.group(2))); continue
    mm = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(\S.*?);', l)
    if mm:
        m[int(mm.group(1), 16)] = cur
iA = h.index('Address'); iI = h.index('Instructions Executed'); iT = h.index('Thread Instructions Executed'); iN = h.index('# Samples')
SRC = {}
def region(k):
    if k is None:
        return 'unknown'
    f, l = k
    if f in ('rrt_trace.cuh', 'rrt_wavetrace.cuh'):
        if f not in SRC:
            SRC[f] = open('/root/repo/relativistic-ray-tracer_b200/csrc/' + f).read().split('\n')
        for i in range(min(l, len(SRC[f])) - 1, -1, -1):
            mm = re.match(r'__device__ __forceinline__ \S+ (\w+)\(', SRC[f][i])
            if mm:
                return f[4:-4] + '.' + mm.group(1)
        return f
    if f == 'rrt_math.cuh':
        return 'math (vector helpers)'
    return f
base = None
reg = collections.defaultdict(lambda: [0, 0, 0])
for r in rows[2:]:
    try:
        a = int(r[iA], 16); n = int(r[iI]); t = int(r[iT]); s = int(r[iN])
    except (ValueError, IndexError):
        continue
    if base is None:
        base = a
    g = reg[region(m.get(a - base))]; g[0] += n; g[1] += t; g[2] += s
tot = sum(v[0] for v in reg.values()); ts = sum(v[2] for v in reg.values()); tt = sum(v[1] for v in reg.values())
print("kernel %s: %d warp instructions, %.1f active threads per instruction" % (rows[0][1][:60], tot, tt / tot))
for k, v in sorted(reg.items(), key=lambda x: -x[1][0])[:16]:
    print("%-34s inst %5.1f%%  threads/inst %5.1f  stall samples %5.1f%%" % (k, 100 * v[0] / tot, v[1] / max(v[0], 1), 100 * v[2] / max(ts, 1)))
