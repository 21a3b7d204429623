#!/bin/bash
mkdir -p gpurun_out
{
for st in 2 3; do B2OT_STAGES=$st python scripts/prof_case.py --n 16384 --m 16384 --iters 100 --label "stages=$st"; done
B2OT_L2_HINT=0 python scripts/prof_case.py --n 16384 --m 16384 --iters 100 --label "nohint"
B2OT_MAX_CLUSTER=16 python scripts/prof_case.py --n 16384 --m 65536 --iters 50 --label "cs16"
python scripts/prof_case.py --n 16384 --m 65536 --iters 50 --label "cs8"
python scripts/prof_case.py --n 32768 --m 8192 --iters 100 --label "cs1"
python scripts/prof_case.py --n 65536 --m 4096 --iters 100 --label "cs1-vpt4"
} > gpurun_out/prof1_timing.log 2>&1
cat gpurun_out/prof1_timing.log | grep -E "PERF|Error|error"
python scripts/prof_case.py --n 16384 --m 16384 --iters 3 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fused_sweep -s 4 -c 2 -f -o gpurun_out/prof_knopp16k python scripts/prof_case.py --n 16384 --m 16384 --iters 3 > gpurun_out/ncu1.log 2>&1
tail -5 gpurun_out/ncu1.log
