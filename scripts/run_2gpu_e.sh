#!/bin/bash
# guarded 2-rank sanity check (every command under its own timeout)
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
timeout 55 $TR --master-port 29701 scripts/mgpu_parity.py 1 half > gpurun_out/e_parity.log 2>&1; echo "parity rc=$?"
grep -E "^(lj|eam)|rror" gpurun_out/e_parity.log | grep -E "step (0|100)|rror" | head -8
timeout 55 $TR --master-port 29711 bench.py --gpus 2 --workload eam --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/e_bench.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/e_bench.log | python -c "import json,sys;d=json.loads(sys.stdin.read());print('eam N=2 %s %.1f Matom-steps/s'%(d['roofline']['list'],d['value']/1e6),d['roofline']['phases_ms_per_step'])" 2>&1 | tail -1
