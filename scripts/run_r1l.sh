#!/bin/bash
# round 1, session l: ping-pong log sweep after the 64-source fix: parity (whole GPU suite) + config 2 / 65536 timings
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r1l_all_tests.log 2>&1; echo "all gpu tests rc=$?" | tee -a gpurun_out/r1l_pp.log
tail -3 gpurun_out/r1l_all_tests.log
run() { # name, env...
  name=$1; shift
  echo "== $name" | tee -a gpurun_out/r1l_pp.log
  env "$@" python scripts/bench_config2.py 2>&1 | tail -1 | tee gpurun_out/r1l_last.json | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in d.items() if k in ('sweep_ms', 'ms_per_iteration', 'api_call_ms', 'it_per_s', 'frac', 'marginal_err_last', 'sweep')})" | tee -a gpurun_out/r1l_pp.log
}
N=16384 run "n=m=16384 default" B2OT_DUMMY=1
cp gpurun_out/r1l_last.json gpurun_out/r1l_config2.json
N=16384 run "n=m=16384 occ1" B2OT_PP_OCC=1
N=65536 run "n=m=65536 default" B2OT_DUMMY=1
N=65536 run "n=m=65536 occ1" B2OT_PP_OCC=1
N=8192 M=65536 run "n=8192 m=65536 default (per-rank shape at 8 GPUs)" B2OT_DUMMY=1
N=16384 M=4096 run "n=16384 m=4096 default" B2OT_DUMMY=1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:log_pp_sweep_kernel -s 30 -c 1 -f -o gpurun_out/r1l_logpp python scripts/bench_config2.py > gpurun_out/r1l_ncu.log 2>&1; echo "ncu rc=$?" | tee -a gpurun_out/r1l_pp.log
