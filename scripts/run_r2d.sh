#!/bin/bash
mkdir -p gpurun_out
B2OT_LIB=$PWD/pot_b200/libb2ot_epi16.so timeout 900 python -m pytest tests/test_empirical_gpu.py -m gpu -q 2>&1 | tail -3 | tee gpurun_out/r2d.log
for lib in libb2ot_epi16.so libb2ot.so libb2ot_epi16.so libb2ot.so; do
  echo "== $lib" | tee -a gpurun_out/r2d.log
  B2OT_LIB=$PWD/pot_b200/$lib SIZES=16384,32768,65536 timeout 300 python scripts/bench_emp.py 2>&1 | grep -v simt | cut -c1-110 | tee -a gpurun_out/r2d.log
done
