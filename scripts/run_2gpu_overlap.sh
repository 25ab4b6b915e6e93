#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
show() { python -c "import json,sys;d=json.load(open('$1'));print('$2 N=%d list=%s %.1f Matom-steps/s %.3f ms/step'%(d['n_gpus'],d['roofline']['list'],d['value']/1e6,d['ms_per_step']),{k:round(v,3) for k,v in d['roofline']['phases_ms_per_step'].items()})"; }
$TR --master-port 29701 scripts/mgpu_parity.py 1 half 2>&1 | grep -E "^(lj|eam|   rank0)|rror"
$TR --master-port 29702 scripts/mgpu_parity.py 1 full 2>&1 | grep -E "^(lj|eam)|rror" | grep "step 100"
for wl in lj eam; do for ov in 1 0; do
  B200MD_OVERLAP=$ov $TR --master-port 29711 bench.py --gpus 2 --workload $wl --list half --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/o2_${wl}_$ov.json; show gpurun_out/o2_${wl}_$ov.json "$wl overlap=$ov"
done; done
B200MD_SPLIT=0.3 $TR --master-port 29712 bench.py --gpus 2 --workload lj --list half --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/o2_s3.json; show gpurun_out/o2_s3.json "lj split=0.3"
B200MD_SPLIT=0.7 $TR --master-port 29713 bench.py --gpus 2 --workload lj --list half --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/o2_s7.json; show gpurun_out/o2_s7.json "lj split=0.7"
$TR --master-port 29714 bench.py --gpus 2 --workload lj --list full --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/o2_full.json; show gpurun_out/o2_full.json "lj full"
for ov in 1 0; do B200MD_OVERLAP=$ov $TR --master-port 29721 scripts/strong_eam.py 5 half 2>/dev/null | grep "eam/alloy" | cut -c1-120; done
