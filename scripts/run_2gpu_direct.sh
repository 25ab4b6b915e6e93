#!/bin/bash
# 2-rank check of the direct halo (B200MD_HALO=direct): thermo must equal the reference logs like the staged path;
# the 32 000-atom loop time is latency-dominated, so staged vs direct shows the per-step halo cost
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
B200MD_HALO=direct timeout 45 $TR --master-port 29721 scripts/mgpu_parity.py 1 half > gpurun_out/direct_parity_half.log 2>&1; echo "direct half rc=$?"
grep -E "^(lj|eam)|rror|rank0" gpurun_out/direct_parity_half.log | grep -E "step (0|100)|rror|rank0" | head -12
timeout 40 $TR --master-port 29731 scripts/mgpu_parity.py 1 half > gpurun_out/staged_parity_half.log 2>&1; echo "staged half rc=$?"
grep -E "rank0|rror" gpurun_out/staged_parity_half.log | head -4
