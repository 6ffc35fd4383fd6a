#!/bin/bash
# one full ncu capture each of a general / TEACH assembly kernel and the nc=6 line solver (families bench)
mkdir -p gpurun_out
CMD="python tools/bench_families.py buoyancy"
ncu --set full --clock-control none --import-source on -k regex:'k_gen_assemble_v' -s 2 -c 1 -f -o gpurun_out/prof_gen_v $CMD > gpurun_out/ncu_full_gen_v.log 2>&1
echo "rc=$?"
CMD="python tools/bench_families.py step"
ncu --set full --clock-control none --import-source on -k regex:'k_teach_assemble_u' -s 2 -c 1 -f -o gpurun_out/prof_teach_u $CMD > gpurun_out/ncu_full_teach_u.log 2>&1
echo "rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_tdma_tma' -s 30 -c 1 -f -o gpurun_out/prof_teach_tdma $CMD > gpurun_out/ncu_full_teach_tdma.log 2>&1
echo "rc=$?"
ls -la gpurun_out | tail -6
