CMD="python tools/bench_families.py implicit"
$CMD > gpurun_out/plain_lod.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 30 --csv --log-file gpurun_out/lod_launches.csv $CMD > /dev/null 2>&1
python - <<'PY'
import csv
lines=[l for l in open('gpurun_out/lod_launches.csv') if not l.startswith('==')]
for row in list(csv.DictReader(lines))[:27]:
    print(row['Kernel Name'][:40], row['Metric Name'], row['Metric Value'], row['Metric Unit'])
PY
