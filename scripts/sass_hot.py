"""Hot spots of a kernel from `ncu --page source --csv --print-source sass` (scripts/ncu_one.sh): opcode histogram with modifiers,
stall totals, the most sampled instructions with context.  usage: python scripts/sass_hot.py gpurun_out/one_TAG_sass.csv [KERNEL_INDEX] [TOP]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ntop = int(sys.argv[3]) if len(sys.argv) > 3 else 14
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
start = starts[which]
end = starts[which + 1] if which + 1 < len(starts) else len(rows)
hdr = rows[start]
idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[start + 1:end] if len(r) == len(hdr)]
c, ie = idx['Warp Stall Sampling (All Samples)'], idx['Instructions Executed']
tot = sum(int(r[c]) for r in data)
ti = sum(int(r[ie]) for r in data)
print("samples", tot, "warp instructions", ti, "static", len(data))
print("stalls:", ", ".join("%s %.1f%%" % (k[6:], 100.0 * sum(int(r[idx[k]]) for r in data) / tot) for k in hdr if k.startswith('stall_') and 'Not Issued' not in k
                           and sum(int(r[idx[k]]) for r in data) > 0.01 * tot))
ops = collections.Counter()
for r in data:
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_]+(\.[A-Z0-9_]+)*)', r[1])
    ops[m.group(2) if m else '?'] += int(r[ie])
print("opcodes:", "  ".join("%s %.1f%%" % (k, 100.0 * v / ti) for k, v in ops.most_common(40)))
top = sorted(enumerate(data), key=lambda kr: -int(kr[1][c]))[:ntop]
for k, r in top:
    st = max((int(r[idx[h]]), h[6:]) for h in hdr if h.startswith('stall_') and 'Not Issued' not in h)
    print("%5d  %5.1f%%  %-10s %s" % (k, 100.0 * int(r[c]) / tot, st[1], r[1][:90]))
