#!/bin/bash
# parity tests of the matcher + a short bench; prints the headline numbers and per-kernel times
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --no-cpu > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err || tail -c 800 gpurun_out/bench_q.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_q.json") if l.startswith("{")][-1])
print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "verify", round(d["verify"]["value"]), "single", d.get("single_sequence"))
print({k: (round(v, 4) if v else v) for k, v in d["roofline"]["ms_per_launch"].items()})
PY
