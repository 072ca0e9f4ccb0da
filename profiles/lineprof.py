#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` dump (SASS level) per CUDA source line, using the line table
nvdisasm -g prints for the same cubin.  usage: lineprof.py <src.csv> <nvdisasm -g -c output> <kernel substring>"""
import csv, re, sys, collections
src_csv, dis, kname = sys.argv[1:4]
# 1) SASS instruction order -> (file, line, inline chain)
lines = open(dis, errors="ignore").read().split("\n")
in_k = False; cur = ("?", 0); order = []
for ln in lines:
    if ln.startswith("//-") and ".text." in ln:
        in_k = kname in ln
        continue
    if not in_k: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        if "inlined at" in m.group(3) and False: pass
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
        order.append(cur)
# 2) ncu rows in the same order
rows = list(csv.reader(open(src_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
ci = hdr.index("Instructions Executed"); cs = hdr.index("# Samples"); ct = hdr.index("Thread Instructions Executed")
if len(body) != len(order): print("warning: %d ncu rows vs %d sass instrs" % (len(body), len(order)))
agg = collections.defaultdict(lambda: [0, 0, 0])
for r, key in zip(body, order):
    a = agg[key]; a[0] += int(r[ci] or 0); a[1] += int(r[cs] or 0); a[2] += int(r[ct] or 0)
tot = sum(a[0] for a in agg.values()); tots = sum(a[1] for a in agg.values())
print("total warp instr %d, samples %d" % (tot, tots))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    print("%-26s:%4d  inst %5.1f%%  samples %5.1f%%  thr/inst %4.1f" % (key[0], key[1], 100.0 * a[0] / tot, 100.0 * a[1] / max(tots, 1), a[2] / max(a[0], 1)))
