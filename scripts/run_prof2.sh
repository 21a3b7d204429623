#!/bin/bash
mkdir -p gpurun_out
python scripts/prof_case.py --mode log --n 16384 --m 16384 --iters 3 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fused_sweep -s 4 -c 1 -f -o gpurun_out/prof_log16k python scripts/prof_case.py --mode log --n 16384 --m 16384 --iters 3 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
python scripts/prof_case.py --mode knopp --n 65536 --m 65536 --iters 3 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fused_sweep -s 4 -c 1 -f -o gpurun_out/prof_knopp64k python scripts/prof_case.py --mode knopp --n 65536 --m 65536 --iters 3 > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/ncu3.log
