#!/bin/bash
# What round 1 finished after its GPU budget ended; run these first in round 2.
#   1 GPU : bash scripts/round2_first.sh one
#   2 GPUs: bash scripts/round2_first.sh two        (gpurun --gpus 2)
#   8 GPUs: bash scripts/round2_first.sh eight      (gpurun --gpus 8)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
case "$1" in
one)
  timeout 120 python -m pytest tests/test_zz_optin.py -m gpu -q -p no:cacheprovider > gpurun_out/r2_optin.log 2>&1; tail -n 3 gpurun_out/r2_optin.log
  for f in 0 1; do
    B200MD_FUSE_NVE=$f timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r2_bench_lj_fuse$f.json 2> gpurun_out/r2_bench_lj_fuse$f.err
    python -c "import json;d=json.loads(open('gpurun_out/r2_bench_lj_fuse$f.json').read().splitlines()[-1]);print('fuse=$f', d['value']/1e6, d['roofline']['phases_ms_per_step'])"
  done ;;
two)
  timeout 90 $TR --nproc-per-node 2 --master-port 29741 scripts/mgpu_parity.py 1 half chain > gpurun_out/r2_parity_half.log 2>&1; echo "half rc=$?"
  grep -E "step (0|100)|chain|rror" gpurun_out/r2_parity_half.log | head -8
  timeout 90 $TR --nproc-per-node 2 --master-port 29742 scripts/mgpu_parity.py 1 full chain > gpurun_out/r2_parity_full.log 2>&1; echo "full rc=$?"
  grep -E "step (0|100)|chain|rror" gpurun_out/r2_parity_full.log | head -8
  for h in direct staged; do
    B200MD_HALO=$h timeout 200 $TR --nproc-per-node 2 --master-port 29743 bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench2_$h.log 2>&1
    tail -1 gpurun_out/r2_bench2_$h.log | python -c "import json,sys;d=json.loads(sys.stdin.read());print('$h', d['value']/1e6, d['config'].get('list_trial_ms_per_step'), d['roofline']['phases_ms_per_step'])"
  done ;;
eight)
  for wl in lj eam; do for h in direct staged; do
    B200MD_HALO=$h timeout 300 $TR --nproc-per-node 8 --master-port 29751 bench.py --gpus 8 --workload $wl --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench8_${wl}_$h.log 2>&1
    tail -1 gpurun_out/r2_bench8_${wl}_$h.log | python -c "import json,sys;d=json.loads(sys.stdin.read());print('$wl $h', d['value']/1e6, d['config'].get('list_trial_ms_per_step'), d['roofline']['phases_ms_per_step'])"
  done; done ;;
*) echo "usage: $0 one|two|eight" ;;
esac
