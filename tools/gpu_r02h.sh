#!/bin/bash
TAG=${1:-r02h}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python tools/bench_configs.py > gpurun_out/${TAG}_bench_configs.jsonl 2> gpurun_out/${TAG}_bench_configs.err; echo "configs exit $?"; tail -3 gpurun_out/${TAG}_bench_configs.err
python tools/bench_configs.py --only C5 --vpm 72 --no-cpu >> gpurun_out/${TAG}_bench_configs.jsonl 2>> gpurun_out/${TAG}_bench_configs.err
cut -c1-1500 gpurun_out/${TAG}_bench_configs.jsonl
for v in 72 100; do python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e --no-extras --vpm $v > gpurun_out/${TAG}_bench_vpm$v.json 2>/dev/null; python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench_vpm$v.json")); r=d["roofline"]
print("vpm $v: value %.4g ms/step %.3f p1 %.4f p2 %.4f build %.3f items %.4f" % (d["value"], d["ms_per_step"], r["phase1_avg_launch_ms"], r["phase2_avg_launch_ms"], d["world"]["build_ms"], d["world"]["narrow_items_per_config"]))
PY
done
C5="python tools/bench_configs.py --only C5 --steps 4 --vpm 72 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_c5_launches.csv $C5 > /dev/null 2>&1
grep -E "k_count|k_scan|k_scatter|k_edt|memset|Memset|k_check_edges|k_resolve|k_c5" gpurun_out/${TAG}_c5_launches.csv | awk -F'","' '{print $5, $(NF)}' | sort | uniq -c | sort -rn | head -5
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
