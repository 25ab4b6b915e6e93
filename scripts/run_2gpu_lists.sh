#!/bin/bash
# 2 GPUs: half/newton list (reverse_comm of f and rho over NCCL) parity, then half vs full timing
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
$TR --master-port 29701 scripts/mgpu_parity.py 1 half 2>&1 | grep -E "^(lj|eam|   rank0)|rror"
for wl in lj eam; do for lk in full half; do
  $TR --master-port 29711 bench.py --gpus 2 --workload $wl --list $lk --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/n2_${wl}_${lk}.json
  python -c "import json;d=json.load(open('gpurun_out/n2_${wl}_${lk}.json'));print('N=2 $wl $lk %.1f Matom-steps/s %.3f ms/step'%(d['value']/1e6,d['ms_per_step']),d['roofline']['phases_ms_per_step'])"
done; done
for lk in full half; do python scripts/strong_eam.py 5 $lk; $TR --master-port 29721 scripts/strong_eam.py 5 $lk 2>/dev/null | grep "eam/alloy"; done
for wl in lj eam; do for lk in full half; do
  python bench.py --gpus 1 --workload $wl --list $lk --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/n1_${wl}_${lk}.json
  python -c "import json;d=json.load(open('gpurun_out/n1_${wl}_${lk}.json'));print('N=1 $wl $lk %.1f Matom-steps/s %.3f ms/step e2e %.1f'%(d['value']/1e6,d['ms_per_step'],d['e2e']['value']/1e6),d['roofline']['phases_ms_per_step'], 'frac %.3f step_frac %.3f'%(d['roofline']['frac'],d['roofline']['step_frac']))"
done; done
