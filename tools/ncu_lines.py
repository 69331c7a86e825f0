#!/usr/bin/env python
"""Per-source-line sample table from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`.
usage: ncu_lines.py file.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None; hdr = None; data = []
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; si = hdr.index('# Samples'); ie = hdr.index('Instructions Executed'); continue
    if hdr and r[0] and r[0].isdigit() and len(r) > si:
        try: data.append((int(r[si]), int(r[ie]), cur_file, int(r[0]), r[1].strip()[:100]))
        except ValueError: pass
tot = sum(d[0] for d in data); toti = sum(d[1] for d in data)
print('samples', tot, 'warp-insts', toti)
for d in sorted(data, reverse=True)[:top]:
    print(f"{100*d[0]/tot:5.1f}% smp {100*d[1]/max(toti,1):5.1f}% inst  {d[2]}:{d[3]:<4d} {d[4]}")
