#!/bin/bash
# tests + a short device-resident bench (no e2e / CPU legs); prints the phase table
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python bench.py --steps 10 --no-e2e --no-cpu-baseline 2>gpurun_out/quick.err | tail -1 > gpurun_out/quick.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/quick.json"))
print("reads/s %.3e  ms/step %.3f  vote frac %.3f  launches %d" % (d["value"], d["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"]))
print({k: round(v, 3) for k, v in d["roofline"]["phase_ms_per_step"].items()})
PY
tail -3 gpurun_out/quick.err
