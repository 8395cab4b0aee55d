#!/bin/bash
# Short device-resident bench of every library in variants/ (and the in-tree one): phase table per variant.
for lib in consensuscruncher_b200/libccgpu.so variants/*.so; do
  [ -f "$lib" ] || continue
  CCGPU_LIB=$PWD/$lib python bench.py --steps 10 --no-e2e --no-cpu-baseline --sharded none --sweep none 2>gpurun_out/var.err | tail -1 > gpurun_out/var.json
  python - "$lib" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/var.json"))
    ph = d["roofline"]["phase_ms_per_step"]
    print("%-40s ms/step %.3f  %s" % (sys.argv[1], d["ms_per_step"], {k: round(v, 3) for k, v in ph.items()}))
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/var.err").read()[-300:])
PY
done
