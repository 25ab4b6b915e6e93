#!/bin/bash
python scripts/chain_run.py 40 > gpurun_out/plain_chain.log 2>&1 || { cat gpurun_out/plain_chain.log; exit 1; }
cat gpurun_out/plain_chain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_chain.csv python scripts/chain_run.py 40 4 > gpurun_out/ncu_chain.log 2>&1
tail -2 gpurun_out/ncu_chain.log
