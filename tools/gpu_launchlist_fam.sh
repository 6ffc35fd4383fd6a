#!/bin/bash
# ncu launch list (kernel durations) for the buoyancy / backward-step configs
for w in ${1:-buoyancy step}; do
CMD="python tools/bench_families.py $w"
$CMD > gpurun_out/fam_$w.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 150 --csv --log-file gpurun_out/launches_$w.csv $CMD > gpurun_out/ncu_fam_$w.log 2>&1
cat gpurun_out/fam_$w.log
python - $w <<'PY'
import csv, collections, re, sys
lines=[l for l in open('gpurun_out/launches_%s.csv' % sys.argv[1]) if not l.startswith('==')]
agg=collections.OrderedDict()
for row in csv.DictReader(lines):
    k=re.sub(r'\(.*','',row['Kernel Name']); agg.setdefault(k,[]).append(float(row['Metric Value'].replace(',','')))
for k,v in agg.items(): print(f"{k:45s} n={len(v):4d} mean={sum(v)/len(v)/1e3:9.1f} us  total={sum(v)/1e3:9.1f}")
PY
done
