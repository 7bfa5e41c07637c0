#!/bin/bash
# ncu launch list of one bench command.  Usage: bash tools/gpu_launchlist.sh <tag> <bench args...>
OUT=gpurun_out/$1; shift; mkdir -p $OUT
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches.csv python bench.py "$@" --no-cpu-baseline --no-e2e --steps 2 --warmup 1 > $OUT/ncu.log 2>&1; echo "exit $?"
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("$OUT/launches.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(',',''))
    except: continue
    a=agg.setdefault(r[ki][:70],[0,0.0]); a[0]+=1; a[1]+=v
tot=sum(a[1] for a in agg.values())
for k,a in sorted(agg.items(), key=lambda x:-x[1][1])[:12]: print(f"{k:72s} n={a[0]:3d} total={a[1]/1e6:9.3f} ms avg={a[1]/a[0]/1e6:8.3f} ms {100*a[1]/tot:5.1f}%")
PY
