#!/bin/bash
# one short C2 profile-mode line: step time and per-kernel times
python bench.py --profile-mode --steps 20 --warmup 3 > gpurun_out/prof_mode.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/prof_mode.json"))
print(d["ms_per_step"], d["ms_per_step_profiled"])
print({k: round(v["ms"] / v["calls"], 4) for k, v in d["kernels"].items()})
PY
