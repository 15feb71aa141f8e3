#!/bin/bash
# kNN split sweep (diagnostic): headline + kNN time for several wave targets
for w in 6 3 12; do
  timeout 300 python bench.py --no-cpu --steps 6 --knn-waves $w > gpurun_out/bench_w$w.json 2> gpurun_out/bench_w$w.err || tail -c 500 gpurun_out/bench_w$w.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_w$w.json") if l.startswith("{")][-1])
print("waves $w value", round(d["value"]), "verify", round(d["verify"]["value"]), "knn ms", round(d["roofline"]["ms_per_launch"]["k_knn2_umma"],4))
PY
done
