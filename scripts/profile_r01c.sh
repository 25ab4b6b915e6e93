#!/bin/bash
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for bl in 4 8 16; do B200MD_BUILD_LANES=$bl B200MD_PHASE_TIMING=1 python scripts/profile_run.py lj 100 2>&1 | tail -1 | sed "s|^|build lanes $bl |"; done
for bl in 4 8; do B200MD_BUILD_LANES=$bl B200MD_PHASE_TIMING=1 python scripts/profile_run.py eam 100 2>&1 | tail -1 | sed "s|^|build lanes $bl |"; done
python scripts/profile_run.py lj 3 > gpurun_out/plain_c.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_build|k_lj' -c 3 -o gpurun_out/prof_build python scripts/profile_run.py lj 3 > gpurun_out/ncu_c.log 2>&1
python bench.py --no-cpu-baseline > gpurun_out/bench_lj_c.json 2> gpurun_out/bench_c.err; tail -c 1500 gpurun_out/bench_lj_c.json
python bench.py --workload eam --no-cpu-baseline > gpurun_out/bench_eam_c.json 2>> gpurun_out/bench_c.err; tail -c 1500 gpurun_out/bench_eam_c.json
