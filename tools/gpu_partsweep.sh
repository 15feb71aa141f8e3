#!/bin/bash
# sweep of the SM split between chains (latency partition) and extraction/kNN (throughput partition)
mkdir -p gpurun_out
for cfg in "0 8" "16 8" "24 8" "32 8" "48 8" "32 12" "32 16" "24 16" "32 1"; do
  set -- $cfg
  timeout 300 python bench.py --steps 4 --no-cpu --latency-sms $1 --sequences $2 2>gpurun_out/part.err | python -c "
import json,sys
d=json.loads(sys.stdin.read())
c=d['config']
print('lat_sms=$1 S=%d value=%.0f e2e=%.0f single=%s part=%s ok=%.3f' % (c['sequences_in_flight'], d['value'], d['e2e']['value'], c['single_sequence_frames_per_s'], c['sm_partition'], c['tracked_ok_frac']))
" || tail -3 gpurun_out/part.err
done
