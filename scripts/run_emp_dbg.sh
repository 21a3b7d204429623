#!/bin/bash
for d in 2 10 18 26 6 14 30; do echo "== f16 dbg $d"; B2OT_EMP_DBG=$d SIZES=32768 timeout 300 python scripts/bench_emp.py 2>&1 | grep -v simt | cut -c50-100; done
