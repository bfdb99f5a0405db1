"""Summarise an .ncu-rep (raw page + SASS instruction mix) -- run in the build container (no GPU needed).
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep"""
import collections
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed']
stalls = [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')]
idx = {h: i for i, h in enumerate(hdr)}
for d in data:
    print('-' * 100)
    for w in want:
        if w in idx:
            print("%-66s %s %s" % (w, d[idx[w]], units[idx[w]]))
    st = sorted(((float(d[idx[h]]), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]) for h in stalls), reverse=True)
    print("stalls per issue: " + ", ".join("%s %.2f" % (n, v) for v, n in st[:8]))
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kern, seen, mix = None, set(), {}
for r in csv.reader(io.StringIO(sass)):
    if r and r[0] == 'Kernel Name':
        kern = r[1] if r[1] not in seen else None
        if kern:
            seen.add(kern)
            mix[kern] = collections.Counter()
        continue
    if kern is None or not r or r[0] == 'Address':
        continue
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)(\.[A-Z0-9_.]+)?', r[1])
    op = m.group(2) if m else '?'
    suf = (m.group(3) or '') if m else ''
    if op in ('LDS', 'STS', 'LDG', 'STG', 'LDL', 'STL'):
        op += '.128' if '.128' in suf else '.64' if '.64' in suf else ''
    mix[kern][op] += int(r[5])
for k, c in mix.items():
    tot = sum(c.values())
    print('=' * 100)
    print(k[:90], "warp instructions:", tot)
    print("  " + "  ".join("%s %.1f%%" % (o, 100 * n / tot) for o, n in c.most_common(26)))
