#!/bin/bash
# short ncu launch list (kernel durations) of two benchmark iterations
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
python - <<'PY'
import csv, collections, re
lines=[l for l in open('gpurun_out/launches.csv') if not l.startswith('==')]
agg=collections.OrderedDict()
for row in csv.DictReader(lines):
    k=re.sub(r'\(.*','',row['Kernel Name']); agg.setdefault(k,[]).append(float(row['Metric Value'].replace(',','')))
for k,v in agg.items(): print(f"{k:45s} n={len(v):4d} mean={sum(v)/len(v)/1e3:9.1f} us")
PY
