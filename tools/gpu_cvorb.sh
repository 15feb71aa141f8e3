#!/bin/bash
# cv::ORB branch alone: isolated per-kernel times, ncu launch list, full captures of its kernels
mkdir -p gpurun_out
timeout 300 python tools/cvorb_run.py > gpurun_out/cvorb_run.log 2>&1; cat gpurun_out/cvorb_run.log
CVORB_STEPS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/cvorb_launches.csv python tools/cvorb_run.py > gpurun_out/cvorb_ncu.log 2>&1
for k in k_cv_pyr_level k_cv_fast k_cv_select k_cv_blur; do
  CVORB_STEPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 5 -c 1 -f -o gpurun_out/cvfull_$k python tools/cvorb_run.py > gpurun_out/cvfull_$k.log 2>&1
done
ls -la gpurun_out | tail -12
