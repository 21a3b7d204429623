#!/bin/bash
# round 1, session m: unbalanced / balanced scaling sweep with two CTAs per SM (B2OT_OCC=2) and 16 warps (B2OT_WARPS=16)
mkdir -p gpurun_out
for cfg in "B2OT_DUMMY=1" "B2OT_OCC=2" "B2OT_WARPS=16"; do
  echo "== $cfg" | tee -a gpurun_out/r1m.log
  env $cfg python scripts/bench_unbalanced.py 65536 60 2>&1 | tail -1 | tee -a gpurun_out/r1m.log
done
