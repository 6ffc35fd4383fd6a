#!/bin/bash
# quick tuning sweep of the line-solver chunk size on the GPU box
for m in 13 10 8 7 6 5; do
  echo "M=$m"; CFD_B200_TDMA_M=$m python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(' ms/step %.3f  it/s %.1f  sweep launch %.1f us frac %.3f  layout %s' % (d['ms_per_step'], d['iters_per_s'], d['roofline']['launch_ms']*1e3, d['roofline']['frac'], d['config']['line_layout']))
    else: print(l.rstrip()[:200])
"
done
