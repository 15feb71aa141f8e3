#!/bin/bash
# ncu --set full capture of the kernels named on the command line (one launch each), bench workload, one sequence
mkdir -p gpurun_out
NCU_BENCH="python bench.py --steps 1 --warmup 1 --no-cpu --core-only --frames 66 --sequences 1"
for k in "$@"; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o gpurun_out/full_$k $NCU_BENCH > gpurun_out/ncu_full_$k.log 2>&1
  echo "$k rc=$?"
done
