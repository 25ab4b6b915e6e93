#!/bin/bash
for bl in 1 2 4; do
  B200MD_BUILD_LANES=$bl python scripts/chain_run.py 100 4 2>&1 | tail -1 | sed "s|^|build lanes $bl |"
done
for bl in 1 2 4 8; do
  B200MD_BUILD_LANES=$bl B200MD_PHASE_TIMING=1 python scripts/profile_run.py lj 100 2>&1 | tail -1 | sed "s|^|build lanes $bl |" | cut -c1-150
  B200MD_BUILD_LANES=$bl B200MD_PHASE_TIMING=1 python scripts/profile_run.py eam 100 2>&1 | tail -1 | sed "s|^|build lanes $bl |" | cut -c1-150
done
