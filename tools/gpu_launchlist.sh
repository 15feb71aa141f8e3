#!/bin/bash
# per-launch ncu metrics of the kernels matching $1 (bench workload, one sequence); csv -> gpurun_out/ll_$2.csv
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,launch__grid_size,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum \
  --clock-control none -k regex:$1 -s ${3:-16} -c ${4:-16} --csv --log-file gpurun_out/ll_$2.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu --frames 66 --sequences 1 > gpurun_out/ll_$2.log 2>&1
echo rc=$?
