#!/bin/bash
O=gpurun_out
python -m pytest tests -x -q -m gpu -k "host" > $O/c18_tests.log 2>&1; tail -2 $O/c18_tests.log
B="python bench.py --steps 20 --warmup 5 --headline-only --no-cpu"
for d in 6 8 10 12 16 7 9; do
  SLB_BENCH_DEPTH=$d $B > $O/c18_d$d.json 2> $O/c18_d$d.err
done
python - <<'PY'
import json
for w in (6, 7, 8, 9, 10, 12, 16):
    try:
        d = json.load(open("gpurun_out/c18_d%d.json" % w)); e = d["e2e"]
        print("depth", w, "value %.1f M  e2e %.1f M  serial %.1f M" % (d["value"] / 1e6, e["value"] / 1e6, e["serial_value"] / 1e6))
    except Exception as ex:
        print(w, "failed", ex)
PY
