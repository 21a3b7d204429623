#!/bin/bash
# end-of-round evidence on one GPU: full GPU test suite, smoke, every bench config + the reference arm, launch list of the
# default bench command, ncu --set full (DRAM traffic) of the headline sweep and of the config-2 sweep
mkdir -p gpurun_out/r2y
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2y/pytest_all.log 2>&1; echo "pytest all rc=$?"; tail -2 gpurun_out/r2y/pytest_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2y/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2y/smoke.log
for c in 3 2 1 5 4; do
  timeout 900 python bench.py --config $c > gpurun_out/r2y/bench_c$c.json 2> gpurun_out/r2y/bench_c$c.err; echo "bench c$c rc=$?"
  cut -c1-200 gpurun_out/r2y/bench_c$c.json
done
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r2y/ref_c3.json 2> gpurun_out/r2y/ref_c3.err; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2y/launches_bench_n1.csv \
  python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --no-torch-ref > gpurun_out/r2y/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_sweep -s 3 -c 1 -f -o gpurun_out/r2y/knopp_65536 \
  python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --no-torch-ref > gpurun_out/r2y/ncu_knopp.log 2>&1; echo "ncu knopp rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:log_pp_sweep -s 3 -c 1 -f -o gpurun_out/r2y/log_16384 \
  python bench.py --config 2 --steps 4 --warmup 3 --no-e2e --no-cpu --no-torch-ref > gpurun_out/r2y/ncu_log.log 2>&1; echo "ncu log rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:resident_knopp -s 1 -c 1 -f -o gpurun_out/r2y/resident_c1 \
  python bench.py --config 1 --steps 200 --warmup 3 --no-e2e --no-cpu --no-torch-ref > gpurun_out/r2y/ncu_res.log 2>&1; echo "ncu resident rc=$?"
ls -la gpurun_out/r2y | head -30
