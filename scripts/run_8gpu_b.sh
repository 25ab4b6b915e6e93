#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for ov in 1 0; do
  B200MD_OVERLAP=$ov $TR --nproc-per-node 8 --master-port 2961$ov bench.py --gpus 8 2>/dev/null | tail -1 > gpurun_out/s8_lj_ov$ov.json
  python -c "import json;d=json.load(open('gpurun_out/s8_lj_ov$ov.json'));print('lj weak N=8 overlap=$ov %.1f Matom-steps/s %.3f ms/step'%(d['value']/1e6,d['ms_per_step']))"
done
B200MD_OVERLAP=1 $TR --nproc-per-node 4 --master-port 29620 bench.py --gpus 4 2>/dev/null | tail -1 > gpurun_out/s4_lj_ov1.json
python -c "import json;d=json.load(open('gpurun_out/s4_lj_ov1.json'));print('lj weak N=4 overlap=1 %.1f Matom-steps/s %.3f ms/step'%(d['value']/1e6,d['ms_per_step']))"
for n in 2 4 8; do $TR --nproc-per-node $n --master-port $((29630+n)) scripts/strong_eam.py 5 2>&1 | grep -E "eam/alloy|rror" | tail -2; done
B200MD_OVERLAP=0 $TR --nproc-per-node 8 --master-port 29640 scripts/strong_eam.py 5 2>&1 | grep -E "eam/alloy|rror" | tail -2
