#!/bin/bash
O=gpurun_out
B="python bench.py --steps 20 --warmup 5 --headline-only --no-cpu"
for z in 0 1 0 1 0 1; do
  SLB_BENCH_BIND=$z $B > $O/c21_z$z.json 2> $O/c21_z$z.err
  python - <<PY
import json
d = json.load(open("gpurun_out/c21_z$z.json")); e = d["e2e"]
print("bind $z value %.1f M  e2e %.1f M  serial %.1f M" % (d["value"] / 1e6, e["value"] / 1e6, e["serial_value"] / 1e6), ["%.0f" % (v / 1e6) for v in e["segment_values"]])
PY
done
