#!/bin/bash
# Round-1 re-validation on a fresh box: GPU tests, smoke, both bench arms, ncu launch list of the bench command,
# empirical (tcgen05) timing + one full ncu capture of its sweep kernel.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 300 gpurun_out/bench_ref.json
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
python bench.py --no-e2e --no-cpu > gpurun_out/bench_noe2e.json 2> gpurun_out/bench_noe2e.err; echo "bench2 exit $?"; cut -c1-260 gpurun_out/bench_noe2e.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo "ncu list exit $?"
SIZES=16384,65536 python scripts/bench_emp.py > gpurun_out/bench_emp.jsonl 2> gpurun_out/bench_emp.err; echo "emp exit $?"; cat gpurun_out/bench_emp.jsonl
SIZES=16384 ncu --set full --clock-control none --import-source on -k regex:emp_tc_lse -s 4 -c 1 -f -o gpurun_out/prof_emp_tc_16k python scripts/bench_emp.py > gpurun_out/ncu_emp.log 2>&1; echo "ncu emp exit $?"
ls -la gpurun_out | tail -20
