#!/usr/bin/env python
"""Reduce an `ncu --page raw --csv` dump of one icp_pair_kernel launch to what the repo tracks:
  profiles/<tag>_icp_pair_full_metrics.csv   the metrics the design discussion cites
  profiles/traffic.json                      dram bytes + warp instructions per launch (read by bench.py)
usage: extract_ncu.py <raw.csv> <tag>"""
import csv, json, os, sys
raw, tag = sys.argv[1:3]
here = os.path.dirname(os.path.abspath(__file__))
rows = list(csv.reader(open(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
pref = ('ID', 'Kernel Name', 'Block Size', 'Grid Size', 'gpu__time', 'dram__', 'sm__inst_executed', 'smsp__inst_executed',
        'smsp__issue', 'sm__warps', 'launch__', 'smsp__thread_inst', 'sm__cycles', 'l1tex__data_pipe_lsu_wavefronts_mem_shared',
        'sm__throughput', 'smsp__average_warp', 'smsp__warps_eligible', 'lts__t_bytes', 'smsp__pcsamp')
with open(os.path.join(here, tag + '_icp_pair_full_metrics.csv'), 'w') as f:
    w = csv.writer(f); w.writerow(['metric', 'unit', 'value'])
    for h, u, v in zip(hdr, units, vals):
        if h.startswith(pref): w.writerow([h, u, v])
d, u = dict(zip(hdr, vals)), dict(zip(hdr, units))
scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
b = lambda k: float(d[k]) * scale[u[k]]
t = {"kernel": d.get('Kernel Name', 'icp_pair_kernel'),
     "capture": "profiles/%s_icp_pair_full_metrics.csv (ncu --set full, bench.py C2 batch, one launch)" % tag,
     "dram_bytes_read": b('dram__bytes_read.sum'), "dram_bytes_write": b('dram__bytes_write.sum'),
     "warp_instr_per_launch": float(d['smsp__inst_executed.sum']),
     "issue_active_pct": float(d['smsp__issue_active.avg.pct_of_peak_sustained_active']),
     "gpu_time_ms_under_ncu": float(d['gpu__time_duration.sum']) * {'ms': 1, 'us': 1e-3, 'ns': 1e-6, 'msecond': 1, 'usecond': 1e-3, 'nsecond': 1e-6}.get(u['gpu__time_duration.sum'], 1)}
t["dram_bytes_per_launch"] = t["dram_bytes_read"] + t["dram_bytes_write"]
t["pairs_per_launch"] = 4096          # bench.py default (C2); bench.py scales both figures for other batch sizes
json.dump(t, open(os.path.join(here, 'traffic.json'), 'w'), indent=1)
print(json.dumps(t))
