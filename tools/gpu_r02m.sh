#!/bin/bash
TAG=${1:-r02m}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log; tail -6 gpurun_out/${TAG}_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json")); r=d["roofline"]; f=d.get("free_workload",{})
print("N=1 value %.4g ms/step %.3f p1 %.4f p2 %.4f build %.3f | free: %.4g | e2e %.4g | cpu %.4g (1 core %.4g) | frac %.3f" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["world"]["build_ms"], f.get("value",0), d["e2e"]["value"], d["cpu_baseline"]["value"], d["cpu_baseline"]["one_core"]["value"], r["frac"]))
print("binding", r.get("binding_resource")); print("executed", r.get("executed"))
PY
python tools/bench_configs.py --only C5 --no-cpu --vpm 72 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('C5 vpm72 %.3f ms/step build %.0f us' % (d['ms_per_step'], d['build_us']))"
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
