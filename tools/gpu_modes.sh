#!/bin/bash
# diagnostic: ceilings of the two halves of the pipeline (extraction only / chains only) beside the full step
mkdir -p gpurun_out
for cfg in "$@"; do
  set -- $cfg
  timeout 300 python bench.py --steps 4 --no-cpu --mode $1 --sequences $2 --latency-sms $3 2>gpurun_out/modes.err | python -c "
import json,sys
d=json.loads(sys.stdin.read())
c=d['config']
print('mode=$1 S=%d lat=$3 value=%.0f e2e=%.0f single=%s ms_per_step=%.2f' % (c['sequences_in_flight'], d['value'], d['e2e']['value'], c['single_sequence_frames_per_s'], d['ms_per_step']))
" || tail -3 gpurun_out/modes.err
done
