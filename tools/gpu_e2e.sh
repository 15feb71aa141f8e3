#!/bin/bash
mkdir -p gpurun_out
for flags in "" "--host-zero-copy"; do
  timeout 300 python bench.py --steps 4 --no-cpu $flags 2>gpurun_out/e2e.err | python -c "
import json,sys
d=json.loads(sys.stdin.read())
c=d['config']
print('flags=[$flags] value=%.0f e2e=%.0f ok=%.3f' % (d['value'], d['e2e']['value'], c['tracked_ok_frac']))
" || tail -3 gpurun_out/e2e.err
done
