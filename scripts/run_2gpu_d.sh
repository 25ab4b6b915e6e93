#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
show() { python -c "import json,sys;d=json.load(open('$1'));print('$2 N=%d list=%s %.1f Matom-steps/s %.3f ms/step'%(d['n_gpus'],d['roofline']['list'],d['value']/1e6,d['ms_per_step']),{k:round(v,3) for k,v in d['roofline']['phases_ms_per_step'].items()})"; }
$TR --master-port 29701 scripts/mgpu_parity.py 1 half 2>&1 | grep -E "^(lj|eam)|rror" | grep -E "step (0|100)|rror"
$TR --master-port 29702 scripts/mgpu_parity.py 1 full 2>&1 | grep -E "^(lj|eam)|rror" | grep -E "step 100|rror"
$TR --master-port 29711 bench.py --gpus 2 --workload lj --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/d2_lj.json; show gpurun_out/d2_lj.json lj
$TR --master-port 29712 bench.py --gpus 2 --workload lj --list half --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/d2_ljh.json; show gpurun_out/d2_ljh.json lj
$TR --master-port 29713 bench.py --gpus 2 --workload eam --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/d2_eam.json; show gpurun_out/d2_eam.json eam
$TR --master-port 29714 bench.py --gpus 2 --workload chain --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/d2_chain.json; show gpurun_out/d2_chain.json chain
