"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
usage: python scripts/launch_summary.py launches.csv "header line" > profiles/xxx_summary.txt"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
tot = collections.defaultdict(float)
cnt = collections.Counter()
for r in rows[1:]:
    if r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    us = v / 1e3 if unit in ("ns", "nsecond") else (v * 1e3 if unit in ("ms", "msecond") else v)
    name = r[ix["Kernel Name"]]
    tot[name] += us
    cnt[name] += 1
total = sum(tot.values())
for line in sys.argv[2:]:
    print("# " + line)
print("# %d launches, %.1f us of GPU time (serialised by the profiler, cold caches)" % (sum(cnt.values()), total))
group = collections.defaultdict(float)
for name, us in tot.items():
    m = re.search(r"(fwd_xy|z_pass|z_warp|z_part|z_tma|inv_xy|inv_ring|sep_xy|sep_z)_kernel", name)
    fam = {"z_warp": "z_pass", "z_part": "z_pass", "z_tma": "z_pass", "inv_ring": "inv_xy", "sep_xy": "direct_separable", "sep_z": "direct_separable"}
    group[fam.get(m.group(1), m.group(1)) if m else "other"] += us
print("# by kernel family: " + ", ".join("%s %.3f" % (k, v / total) for k, v in sorted(group.items(), key=lambda kv: -kv[1])))
for name, us in sorted(tot.items(), key=lambda kv: -kv[1])[:40]:
    print("%-70.70s   n=%5d  total_us=%10.1f  avg_us=%8.1f  share=%.3f" % (name, cnt[name], us, us / cnt[name], us / total))
