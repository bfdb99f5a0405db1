"""Segments of a kernel's SASS by execution count (ncu source page CSV): where the non-arithmetic instructions are.
usage: python scripts/sass_segments.py gpurun_out/one_TAG_sass.csv [KERNEL_INDEX] [MIN_SHARE_PCT]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
min_share = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hdr = rows[starts[which]]
idx = {h: i for i, h in enumerate(hdr)}
end = starts[which + 1] if which + 1 < len(starts) else len(rows)
data = [r for r in rows[starts[which] + 1:end] if len(r) == len(hdr)]
ie, sc = idx['Instructions Executed'], idx['Warp Stall Sampling (All Samples)']
ti = sum(int(r[ie]) for r in data)
ts = sum(int(r[sc]) for r in data)
prev, seg = None, []
for k, r in enumerate(data):
    n = int(r[ie])
    if prev is None or abs(n - prev) > 0.02 * max(n, prev, 1):
        seg.append([k, k, 0, collections.Counter(), 0])
    s = seg[-1]
    s[1] = k
    s[2] += n
    s[4] += int(r[sc])
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)', r[1])
    s[3][m.group(2)] += 1
    prev = n
for a, b, n, c, smp in seg:
    share = 100.0 * n / ti
    if share >= min_share:
        fp = sum(v for k, v in c.items() if k in ('FADD', 'FFMA', 'FMUL', 'MUFU', 'FMNMX'))
        print("%5d-%5d  exec/instr %10d  len %4d  instr share %5.1f%%  samples %5.1f%%  fp %3d%%  %s" %
              (a, b, n // (b - a + 1), b - a + 1, share, 100.0 * smp / ts, 100 * fp // (b - a + 1), c.most_common(7)))
