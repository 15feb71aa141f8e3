#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (both arms), ncu launch list + full captures. Outputs under gpurun_out/.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
NCU_BENCH="python bench.py --steps 1 --warmup 1 --no-cpu --core-only --frames 66 --sequences 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 2500 --csv --log-file gpurun_out/launches.csv $NCU_BENCH > gpurun_out/ncu_bench.log 2>&1
# -s 3: launches 0-2 of an extraction kernel are priming / warm-up (launch 2 reads host images), 3 is the device-resident step
for k in k_fast_cells k_orient_describe k_knn2_ k_gauss7 k_pyramid_resize k_pyramid_base k_quadtree k_se3_hypotheses k_se3_prepare; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/full_$k $NCU_BENCH > gpurun_out/ncu_full_$k.log 2>&1
done
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/bench.json
