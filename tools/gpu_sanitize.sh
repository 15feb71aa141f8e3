#!/bin/bash
# compute-sanitizer passes over the hot path on the GPU box: memcheck, racecheck, initcheck and synccheck over smoke() and over the
# chain / hard-RANSAC / verify / kNN-tile-edge parity tests (the kernels with cross-CTA locks, mbarriers and TMEM).  Summaries under gpurun_out/.
mkdir -p gpurun_out/sanitizer
CS=/usr/local/cuda/bin/compute-sanitizer
TESTS="tests/test_gpu_parity.py::test_ransac_se3_hard_cases tests/test_gpu_parity.py::test_match_tile_edges_both_knn_kernels tests/test_gpu_parity.py::test_verify_batch_matches_per_pair_oracle tests/test_sharded_verify.py::test_verify_sharded_one_rank_equals_verify_batch_and_pipelines tests/test_loopdetect.py"
for tool in memcheck racecheck initcheck synccheck; do
  timeout 600 $CS --tool $tool --print-limit 20 --error-exitcode 9 python __graft_entry__.py smoke > gpurun_out/sanitizer/smoke_$tool.log 2>&1
  echo "smoke $tool rc=$?" >> gpurun_out/sanitizer/summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|smoke ok" gpurun_out/sanitizer/smoke_$tool.log | tail -3 >> gpurun_out/sanitizer/summary.txt
done
for tool in memcheck racecheck; do
  timeout 1500 $CS --tool $tool --print-limit 20 --error-exitcode 9 python -m pytest $TESTS -m gpu -x -q > gpurun_out/sanitizer/tests_$tool.log 2>&1
  echo "tests $tool rc=$?" >> gpurun_out/sanitizer/summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitizer/tests_$tool.log | tail -3 >> gpurun_out/sanitizer/summary.txt
done
cat gpurun_out/sanitizer/summary.txt
