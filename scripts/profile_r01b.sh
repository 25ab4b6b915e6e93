#!/bin/bash
# ncu evidence for the half/newton kernels (defaults of bench.py at N=1): launch list + --set full
set -u
mkdir -p gpurun_out
python scripts/profile_run.py lj 25 > gpurun_out/plain_lj_half.log 2>&1 || { cat gpurun_out/plain_lj_half.log; exit 1; }
cat gpurun_out/plain_lj_half.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_lj_half.csv python scripts/profile_run.py lj 25 > gpurun_out/ncu_l1.log 2>&1
python scripts/profile_run.py eam 12 > gpurun_out/plain_eam_half.log 2>&1 || { cat gpurun_out/plain_eam_half.log; exit 1; }
cat gpurun_out/plain_eam_half.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_eam_half.csv python scripts/profile_run.py eam 12 > gpurun_out/ncu_l2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_lj|k_build' -s 1 -c 3 -o gpurun_out/prof_lj_half python scripts/profile_run.py lj 3 > gpurun_out/ncu_f1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_eam' -s 3 -c 3 -o gpurun_out/prof_eam_half python scripts/profile_run.py eam 3 > gpurun_out/ncu_f2.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_*half.csv
