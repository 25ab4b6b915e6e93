#!/bin/bash
# 8-GPU box after the half-list / occupancy retune: parity on 8 ranks (half list), weak scaling
# in.lj and in.eam (both list kinds at N=8), strong scaling C5
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
show() { python -c "import json,sys;d=json.load(open('$1'));print('$2 N=%d list=%s %.1f Matom-steps/s %.3f ms/step'%(d['n_gpus'],d['roofline']['list'],d['value']/1e6,d['ms_per_step']),{k:round(v,3) for k,v in d['roofline']['phases_ms_per_step'].items()})"; }
$TR --nproc-per-node 8 --master-port 29601 scripts/mgpu_parity.py 1 half 2>&1 | grep -E "^(lj|eam|   rank0)|rror"
for wl in lj eam; do
  python bench.py --gpus 1 --workload $wl --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/c_${wl}_1.json; show gpurun_out/c_${wl}_1.json $wl
  for lk in half full; do
    $TR --nproc-per-node 8 --master-port 29611 bench.py --gpus 8 --workload $wl --list $lk --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/c_${wl}_8_$lk.json; show gpurun_out/c_${wl}_8_$lk.json $wl
  done
done
python scripts/strong_eam.py 5 half
$TR --nproc-per-node 8 --master-port 29631 scripts/strong_eam.py 5 half 2>/dev/null | grep "eam/alloy"
$TR --nproc-per-node 4 --master-port 29632 scripts/strong_eam.py 5 half 2>/dev/null | grep "eam/alloy"
$TR --nproc-per-node 2 --master-port 29633 scripts/strong_eam.py 5 half 2>/dev/null | grep "eam/alloy"
for n in 2 4; do $TR --nproc-per-node $n --master-port $((29640+n)) bench.py --gpus $n --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/c_lj_$n.json; show gpurun_out/c_lj_$n.json lj; done
