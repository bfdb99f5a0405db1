"""DRAM traffic of the convolution kernels against the byte model, from an `ncu --set full` raw page.
usage: python scripts/traffic_summary.py gpurun_out/prof_TAG_raw.csv gpurun_out/prof_TAG.log VIEWS_CAPTURED/VIEWS > profiles/rNN_TAG_dram_traffic.json
The log holds prof_conv.py's 'model' lines (algorithmic bytes per kernel family and iteration)."""
import collections
import csv
import json
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
frac = eval(sys.argv[3]) if len(sys.argv) > 3 else 1.0
model = {}
for line in open(sys.argv[2]):
    m = re.match(r"model (\w+): (\d+) launches, ([\d.]+) ms, ([\d.]+) GB", line)
    if m:
        model[m.group(1)] = dict(launches_per_iteration=int(m.group(2)), event_ms_per_iteration=float(m.group(3)), alg_gb_per_iteration=float(m.group(4)))


def num(d, key):
    v = float(d[ix[key]].replace(",", ""))
    u = units[ix[key]]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "nsecond": 1e-3, "usecond": 1, "msecond": 1e3}.get(u, 1)
    return v * scale


fam = collections.defaultdict(lambda: dict(launches=0, dram_read_mb=0.0, dram_write_mb=0.0, us=0.0))
detail = []
for d in data:
    name = d[ix["Kernel Name"]]
    m = re.search(r"(fwd_xy|z_pass|z_warp|z_part|z_tma|inv_xy|inv_ring)_kernel", name)
    if not m:
        continue
    family = {"z_warp": "z_pass", "z_part": "z_pass", "z_tma": "z_pass", "inv_ring": "inv_xy"}.get(m.group(1), m.group(1))  # kernel generations of one pass
    f = fam[family]
    rd, wr, us = num(d, "dram__bytes_read.sum") / 1e6, num(d, "dram__bytes_write.sum") / 1e6, num(d, "gpu__time_duration.sum")
    f["launches"] += 1
    f["dram_read_mb"] += rd
    f["dram_write_mb"] += wr
    f["us"] += us
    detail.append(dict(kernel=name[:60], grid=d[ix["Grid Size"]], dram_read_mb=round(rd, 1), dram_write_mb=round(wr, 1), us=round(us, 1),
                       dram_pct_of_peak=float(d[ix["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]]),
                       issue_active_pct=float(d[ix["sm__issue_active.avg.pct_of_peak_sustained_elapsed"]])))
out = dict(source="ncu --set full --clock-control none, %d launches of scripts/prof_conv.py 1024 1024 512 31 1 (one B200)" % len(detail),
           note="model = the per-kernel algorithmic byte model of DESIGN.md section 4.2 (what roofline.achieved divides by), "
                "scaled to the captured share (%s) of one iteration; durations are ncu's (serialised, cold caches)" % sys.argv[3],
           kernels={}, launches=detail)
for k, f in fam.items():
    mb = model[k]["alg_gb_per_iteration"] * 1e3 * frac
    out["kernels"][k] = dict(launches=f["launches"], dram_read_mb=f["dram_read_mb"], dram_write_mb=f["dram_write_mb"], model_mb=mb,
                             traffic_over_model=(f["dram_read_mb"] + f["dram_write_mb"]) / mb, us=f["us"],
                             dram_gbs=(f["dram_read_mb"] + f["dram_write_mb"]) / f["us"] * 1e3 / 1e3)
print(json.dumps(out, indent=1))
