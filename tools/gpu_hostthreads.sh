#!/bin/bash
mkdir -p gpurun_out
for cfg in "$@"; do
  set -- $cfg
  timeout 300 python bench.py --steps 4 --no-cpu --sequences $1 --host-threads $2 2>gpurun_out/ht.err | python -c "
import json,sys
d=json.loads(sys.stdin.read())
c=d['config']
print('S=%d T=$2 value=%.0f e2e=%.0f single=%s ok=%.3f' % (c['sequences_in_flight'], d['value'], d['e2e']['value'], c['single_sequence_frames_per_s'], c['tracked_ok_frac']))
" || tail -3 gpurun_out/ht.err
done
