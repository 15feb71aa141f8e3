#!/bin/bash
# sweep: sequences in flight x speculative first chunk
mkdir -p gpurun_out
for cfg in "$@"; do
  set -- $cfg
  timeout 300 python bench.py --steps 4 --no-cpu --sequences $1 --first-chunk $2 2>gpurun_out/sweep.err | python -c "
import json,sys
d=json.loads(sys.stdin.read())
c=d['config']
print('S=%d fc=%s value=%.0f e2e=%.0f single=%s iters=%.2f ok=%.3f' % (c['sequences_in_flight'], '$2', d['value'], d['e2e']['value'], c['single_sequence_frames_per_s'], c['mean_ransac_iterations'], c['tracked_ok_frac']))
" || tail -3 gpurun_out/sweep.err
done
