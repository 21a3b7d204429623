#!/bin/bash
# A/B of the online-cost sweep variants (scripts/build_variants.sh emp ...): parity tests + ms per iteration each
mkdir -p gpurun_out/r2i
for v in ${VLIST:-v0 v1 v2 v3 v4 v5 v6 v7}; do
  export B2OT_LIB=pot_b200/libb2ot_$v.so
  [ -f $B2OT_LIB ] || continue
  timeout 300 python -m pytest tests/test_empirical_gpu.py -m gpu -q -k "tc" > gpurun_out/r2i/pytest_$v.log 2>&1
  echo "$v pytest rc=$? $(tail -1 gpurun_out/r2i/pytest_$v.log)"
  SIZES=32768,65536 VARIANTS=tc timeout 300 python scripts/bench_emp.py > gpurun_out/r2i/bench_$v.log 2>&1
  python - <<PY
import json
for l in open("gpurun_out/r2i/bench_$v.log"):
    if l.startswith("{"):
        d = json.loads(l); print("$v", d["n"], "ms/iter %.3f" % d["ms_per_iter"], "finite", d["finite"], d["info"])
PY
done
unset B2OT_LIB; [ -n "$SKIP_ALL" ] && exit 0
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2i/pytest_all.log 2>&1; echo "pytest all rc=$?"; tail -3 gpurun_out/r2i/pytest_all.log
