#!/bin/bash
# Round-2 ncu captures of the kernels outside the headline iteration: the SIP solve (1025^2, one launch = one 600-iteration
# solve) and the scalar-equation / property kernels of the turbulent backward step.  Launch list of the turbulent iteration.
set -u
mkdir -p gpurun_out
CMD1="python tools/bench_fdcavity.py 1025"
$CMD1 > gpurun_out/fam_plain1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_sip_solve' -s 2 -c 1 -f -o gpurun_out/prof_r02_sip $CMD1 > gpurun_out/ncu_sip.log 2>&1
echo "ncu sip rc=$?"; tail -2 gpurun_out/ncu_sip.log
CMD2="python tools/bench_families.py step_turb"
$CMD2 > gpurun_out/fam_plain2.log 2>&1; cat gpurun_out/fam_plain2.log | cut -c1-300
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_step_turb.csv $CMD2 > gpurun_out/ncu_list2.log 2>&1
echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_teach_scalar|k_teach_props|k_teach_assemble_u_fast|k_tdma_flowc' -s 40 -c 6 -f -o gpurun_out/prof_r02_turb $CMD2 > gpurun_out/ncu_turb.log 2>&1
echo "ncu turb rc=$?"; tail -2 gpurun_out/ncu_turb.log
ls -la gpurun_out/*.ncu-rep
