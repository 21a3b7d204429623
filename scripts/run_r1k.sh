#!/bin/bash
# round 1, session k: ping-pong log sweep, new default (slices <= 4096 columns, two CTAs per SM) vs B2OT_PP_OCC=1, several shapes
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  echo "== $name" | tee -a gpurun_out/r1k_pp.log
  env "$@" python scripts/bench_config2.py 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in d.items() if k in ('sweep_ms', 'ms_per_iteration', 'api_call_ms', 'frac', 'marginal_err_last', 'sweep')})" | tee -a gpurun_out/r1k_pp.log
}
for shape in "16384 16384" "16384 4096" "16384 2048" "8192 8192" "4096 32768" "8192 65536" "65536 65536"; do
  set -- $shape
  N=$1 M=$2 run "n=$1 m=$2 new-default" B2OT_DUMMY=1
  N=$1 M=$2 run "n=$1 m=$2 occ1" B2OT_PP_OCC=1
done
python -m pytest tests/test_dense_gpu.py -x -q -m gpu > gpurun_out/r1k_tests.log 2>&1; echo "dense tests rc=$?" | tee -a gpurun_out/r1k_pp.log
tail -2 gpurun_out/r1k_tests.log
