#!/bin/bash
# round 1, session t: single-GPU validation of the round's final state: whole GPU suite, smoke, default bench,
# reference arm, ncu launch list of the bench command
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r1t_all_tests.log 2>&1; echo "all gpu tests rc=$?" | tee gpurun_out/r1t.log
tail -3 gpurun_out/r1t_all_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1t_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/r1t.log
tail -2 gpurun_out/r1t_smoke.log
python scripts/diag_api_overhead.py 2>&1 | grep "config 1" | tee -a gpurun_out/r1t.log
python bench.py > gpurun_out/r1t_bench.json 2> gpurun_out/r1t_bench.err; echo "bench rc=$?" | tee -a gpurun_out/r1t.log
python bench.py --impl reference > gpurun_out/r1t_bench_ref.json 2> gpurun_out/r1t_bench_ref.err; echo "bench ref rc=$?" | tee -a gpurun_out/r1t.log
cat gpurun_out/r1t_bench.json | cut -c1-2500
cat gpurun_out/r1t_bench_ref.json | cut -c1-600
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1t_launches_bench_n1.csv python bench.py --steps 20 --warmup 3 > gpurun_out/r1t_ncu_bench.log 2>&1; echo "ncu launch list rc=$?" | tee -a gpurun_out/r1t.log
