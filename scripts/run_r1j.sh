#!/bin/bash
# round 1, session j: ping-pong log sweep with wider clusters / two CTAs per SM (experiment switches)
mkdir -p gpurun_out
run() { # name, env...
  name=$1; shift
  echo "== $name" | tee -a gpurun_out/r1j_pp.log
  env "$@" python scripts/bench_config2.py 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in d.items() if k in ('sweep_ms', 'ms_per_iteration', 'frac', 'marginal_err_last', 'sweep')})" | tee -a gpurun_out/r1j_pp.log
}
run default B2OT_DUMMY=1
run cluster4_occ1 B2OT_PP_CLUSTER=4
run cluster4_occ2 B2OT_PP_CLUSTER=4 B2OT_PP_OCC=2
run cluster8_occ2 B2OT_PP_CLUSTER=8 B2OT_PP_OCC=2
run cluster8_occ1 B2OT_PP_CLUSTER=8
run cluster2_stages2 B2OT_PP_STAGES=2
N=65536 run n65536_default B2OT_DUMMY=1
N=65536 run n65536_cluster16_occ2 B2OT_PP_CLUSTER=16 B2OT_PP_OCC=2
for cfg in "B2OT_PP_CLUSTER=4 B2OT_PP_OCC=2" "B2OT_PP_CLUSTER=8 B2OT_PP_OCC=2"; do
  env $cfg python -m pytest tests/test_dense_gpu.py -x -q -m gpu -k "log" > gpurun_out/r1j_tests.log 2>&1; echo "tests [$cfg] rc=$?" | tee -a gpurun_out/r1j_pp.log
  tail -2 gpurun_out/r1j_tests.log
done
