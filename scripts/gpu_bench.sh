#!/bin/bash
O=gpurun_out
( time python bench.py --steps 20 --warmup 5 > $O/c8_bench.json 2> $O/c8_bench.err ) 2> $O/c8_time.txt; echo "bench rc=$?"; tail -3 $O/c8_time.txt; tail -3 $O/c8_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/c8_bench.json"))
print("value %.1f M frac %.3f e2e %.1f M two %s" % (d["value"] / 1e6, d["roofline"]["frac"], d["e2e"]["value"] / 1e6, d.get("two_stream_overlap")))
print({k: v for k, v in d["configs"]["side_effects"].items() if k != "what"})
PY
