#!/bin/bash
# Short device-resident bench of every library in variants/ (and the in-tree one): phase table per variant.
for lib in consensuscruncher_b200/libccgpu.so variants/*.so; do
  [ -f "$lib" ] || continue
  CCGPU_LIB=$PWD/$lib python bench.py --steps 10 --no-e2e --no-cpu-baseline 2>gpurun_out/var.err | tail -1 > gpurun_out/var.json
  python - "$lib" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/var.json"))
    ph = d["roofline"]["phase_ms_per_step"]
    print("%-40s ms/step %.3f  vote %.3f  sort %.3f segment %.3f join %.3f dcs %.3f" % (sys.argv[1], d["ms_per_step"], ph["vote"], ph["sort"], ph["segment"], ph["join"], ph["dcs"]))
except Exception as e:
    print(sys.argv[1], "FAILED", e, open("gpurun_out/var.err").read()[-300:])
PY
done
