#!/usr/bin/env python
"""Reduce an `ncu --page raw --csv` dump of one icp_pair_kernel launch to what the repo tracks:
  profiles/<tag>_icp_pair_full_metrics.csv   the metrics the design discussion cites
  profiles/traffic.json                      dram bytes + warp instructions per launch (read by bench.py)
  (stamped with the SHA-256 of the library the capture was taken from: bench.py reports the derived figures as stale
   when another build is loaded)
usage: extract_ncu.py <raw.csv> <tag> [pairs-in-the-launch] [file holding the library's sha256]"""
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
import hashlib, subprocess
root = os.path.dirname(here)
kname = d.get('Kernel Name', 'icp_pair_kernel').split('(')[0].split('::')[-1]
units = int(sys.argv[3]) if len(sys.argv) > 3 else 4096      # work units (pairs) of the captured launch
k = {"capture": "profiles/%s_icp_pair_full_metrics.csv (ncu --set full, one launch of %d pairs)" % (tag, units),
     "dram_bytes_read": b('dram__bytes_read.sum'), "dram_bytes_write": b('dram__bytes_write.sum'),
     "warp_instr_per_launch": float(d['smsp__inst_executed.sum']),
     "issue_active_pct": float(d['smsp__issue_active.avg.pct_of_peak_sustained_active']),
     "gpu_time_ms_under_ncu": float(d['gpu__time_duration.sum']) * {'ms': 1, 'us': 1e-3, 'ns': 1e-6, 'msecond': 1, 'usecond': 1e-3, 'nsecond': 1e-6}.get(u['gpu__time_duration.sum'], 1),
     "units_per_launch": units}
k["dram_bytes_per_launch"] = k["dram_bytes_read"] + k["dram_bytes_write"]
# the library the capture was taken from: sha256 recorded on the GPU box next to the report (argv[4]), else the in-tree file
sha = open(sys.argv[4]).read().strip() if len(sys.argv) > 4 else hashlib.sha256(open(os.path.join(root, 'slam_b200', 'lib', 'libccicp2d_b200.so'), 'rb').read()).hexdigest()
tp = os.path.join(here, 'traffic.json')
sys.path.insert(0, root)
from slam_b200.build import source_sha256
src = source_sha256()                  # sources + flags in the tree now: run this right after the capture, before editing them
try:
    t = json.load(open(tp))
    # captures of another library are dropped -- unless that library was built from the same sources and flags (nvcc does
    # not reproduce binaries bit for bit): then they stay, and this capture notes the binary it came from
    if t.get("lib_sha256") != sha and t.get("src_sha256") != src: t = {}
except Exception:
    t = {}
t.setdefault("lib_sha256", sha)
t["src_sha256"] = src
k["lib_sha256"] = sha
try: t["commit"] = subprocess.run(["git", "-C", root, "rev-parse", "--short", "HEAD"], stdout=subprocess.PIPE, text=True).stdout.strip()
except Exception: pass
t.setdefault("kernels", {})[kname] = k
json.dump(t, open(tp, 'w'), indent=1)
print(json.dumps(t))
