#!/bin/bash
TAG=${1:-r02l}
mkdir -p gpurun_out
python tools/debug_fcit.py 2>&1 | grep -v DEBUG | tail -5 | cut -c1-300
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 300 python tools/bench_planner.py > gpurun_out/${TAG}_bench_planner.json 2> /dev/null; cat gpurun_out/${TAG}_bench_planner.json
python tools/bench_configs.py --only C5 --no-cpu --vpm 72 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('C5 vpm72 %.3f ms/step build %.0f us' % (d['ms_per_step'], d['build_us']))"
python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --no-extras 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('bench value %.4g build %.3f ms' % (d['value'], d['world']['build_ms']))"
C5="python tools/bench_configs.py --only C5 --steps 4 --vpm 72 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/${TAG}_c5_launches.csv $C5 > /dev/null 2>&1
python - <<PY
import csv,collections
rows=list(csv.reader(open("gpurun_out/${TAG}_c5_launches.csv")))
hdr=[i for i,r in enumerate(rows) if "Kernel Name" in r]
if hdr:
    h=rows[hdr[0]]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
    d=collections.defaultdict(list)
    for r in rows[hdr[0]+1:]:
        if len(r)>vi:
            v=float(r[vi].replace(",","")); v = v/1000 if r[ui]=="ns" else v
            d[r[ki][:40]].append(v)
    for k,v in sorted(d.items(), key=lambda kv:-sum(kv[1])): print("%-42s n=%3d mean %.1f us" % (k,len(v),sum(v)/len(v)))
PY
