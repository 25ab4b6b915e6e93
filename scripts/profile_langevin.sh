#!/bin/bash
# chain + fix langevin: the new parity test, the bench line, and an ncu launch list (2 048 000 atoms, 12 steps)
(timeout 60 python -m pytest tests/test_langevin.py -m gpu -q -p no:cacheprovider > gpurun_out/r01z_langevin2.log 2>&1; echo "exit $?" >> gpurun_out/r01z_langevin2.log)
tail -n 3 gpurun_out/r01z_langevin2.log
timeout 120 python bench.py --workload chain --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r01z_bench_chain_langevin.json 2> gpurun_out/r01z_bench_chain_langevin.err
tail -c 1500 gpurun_out/r01z_bench_chain_langevin.json
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r01z_launches_chain_langevin.csv python scripts/chain_run.py 12 4 langevin > gpurun_out/r01z_ncu_chain_langevin.log 2>&1
tail -n 2 gpurun_out/r01z_ncu_chain_langevin.log
