#!/bin/bash
# compute-sanitizer (memcheck, racecheck, synccheck) over small shapes of every kernel family:
# fused sweep with clusters of 1 / 2 / 8 (DSMEM st.async + mbarrier), log ping-pong sweep, finalize, batch kernel,
# tcgen05 empirical sweep, batch helpers, and the 2-rank peer exchange (two processes on one GPU).
out=gpurun_out/sanitize; mkdir -p $out
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # name tool pytest-args...
  name=$1; tool=$2; shift 2
  timeout 900 $CS --tool $tool --print-limit 20 --log-file $out/${name}_${tool}.log python -m pytest -x -q -m gpu "$@" > $out/${name}_${tool}.pytest.log 2>&1
  echo "$name $tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $out/${name}_${tool}.log | tail -1) | $(tail -1 $out/${name}_${tool}.pytest.log)"
}
for tool in memcheck racecheck synccheck; do
  run dense_shapes $tool tests/test_dense_gpu.py -k "test_oracle_parity_shapes and float32 and (257 or 37-8196 or 50-20000 or 33-65536)"
  run dense_golden $tool tests/test_dense_gpu.py -k "test_knopp_golden_fixed or test_log_golden_fixed or test_numerical_guard"
  run batch $tool tests/test_batch_gpu.py -k "test_solve_batch_golden or test_log_exact_column_pass"
  run emp_tc $tool tests/test_empirical_gpu.py -k "test_tc_golden or (test_tc_vs_oracle and 513)"
  run batchops $tool tests/test_batchops_gpu.py -k "test_batchops_golden"
done
run peer2 memcheck tests/test_multigpu.py -k test_two_ranks_one_gpu
ls -la $out | head -40
