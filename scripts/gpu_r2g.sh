#!/bin/bash
# ncu --set full of the hybrid batch kernel, both domains (B = 1184 problems = 8 waves, 40 iterations in one launch)
mkdir -p gpurun_out/r2g
B=1184 python scripts/bench_batch.py > gpurun_out/r2g/plain.log 2>&1 || { tail gpurun_out/r2g/plain.log; exit 1; }
for k in 0 1; do
ncu --set full --clock-control none --import-source on -k regex:batch_hyb -s $((2 + k * 0)) -c 1 -f -o gpurun_out/r2g/hyb_$k \
  python - <<PY > gpurun_out/r2g/ncu_$k.log 2>&1
import sys, torch
sys.path.insert(0, '.')
from pot_b200 import batch as otb
B, n, m = 1184, 256, 256
g = torch.Generator(device='cuda').manual_seed(0)
M = torch.rand(B, n, m, device='cuda', generator=g)
a = torch.full((B, n), 1.0 / n, device='cuda'); b = torch.full((B, m), 1.0 / m, device='cuda')
mode, tr = (('knopp', 1), ('log', 2))[$k]
for _ in range(4):
    otb._run(mode, M, tr, 0.1, M, a, b, 40, 0.0, True, False, 1, sync_n_iters=False)
torch.cuda.synchronize()
PY
tail -2 gpurun_out/r2g/ncu_$k.log
done
ls -la gpurun_out/r2g
