#!/bin/bash
# extraction contexts sharing K streams (rgbdfe_share_stream): headline for several K
for k in 0 4 8 2; do
  timeout 300 python bench.py --no-cpu --steps 8 --extract-streams $k > gpurun_out/bench_s$k.json 2> gpurun_out/bench_s$k.err || tail -c 600 gpurun_out/bench_s$k.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_s$k.json") if l.startswith("{")][-1])
print("extract streams $k value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ok", d.get("tracking_ok"))
PY
done
