#!/bin/bash
# 8-GPU box: parity on 8 ranks, weak-scaling bench at N=4,8 (N=1,2 for reference), strong scaling of C5
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29601 scripts/mgpu_parity.py 2>&1 | grep -E "^(lj|eam|   rank0)"
for n in 1 2 4 8; do
  if [ $n = 1 ]; then python bench.py --gpus 1 --no-cpu-baseline > gpurun_out/scale_lj_$n.json
  else $TR --nproc-per-node $n --master-port $((29610+n)) bench.py --gpus $n 2>/dev/null | tail -1 > gpurun_out/scale_lj_$n.json; fi
  python -c "import json;d=json.load(open('gpurun_out/scale_lj_$n.json'));print('lj weak N=%d %.1f Matom-steps/s %.3f ms/step comm %.3f e2e %.1f'%(d['n_gpus'],d['value']/1e6,d['ms_per_step'],d['roofline']['phases_ms_per_step']['comm'],d['e2e']['value']/1e6))"
done
for n in 1 2 4 8; do
  if [ $n = 1 ]; then python scripts/strong_eam.py 5
  else $TR --nproc-per-node $n --master-port $((29630+n)) scripts/strong_eam.py 5 2>/dev/null | grep "eam/alloy"; fi
done
$TR --nproc-per-node 8 --master-port 29650 bench.py --gpus 8 --workload eam 2>/dev/null | tail -1 > gpurun_out/scale_eam_8.json
python -c "import json;d=json.load(open('gpurun_out/scale_eam_8.json'));print('eam weak N=8 %.1f Matom-steps/s'%(d['value']/1e6))"
