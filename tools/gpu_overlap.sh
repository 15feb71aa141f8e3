#!/bin/bash
# blur beside the quadtree (option overlap_blur): parity test + headline with and without
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "orb" 2>&1 | tail -2
for o in 0 1 0 1; do
  timeout 300 python bench.py --no-cpu --steps 8 --overlap-blur $o > gpurun_out/bench_o$o.json 2> gpurun_out/bench_o$o.err || tail -c 500 gpurun_out/bench_o$o.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_o$o.json") if l.startswith("{")][-1])
print("overlap $o value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "single", d.get("single_sequence"))
PY
done
