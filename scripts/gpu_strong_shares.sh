#!/bin/bash
O=gpurun_out
python bench_configs.py strong:mixed_26:8192 strong:mixed_26:65536 strong:big_256:512 strong:big_256:4096 > $O/c9_strong.jsonl 2> $O/c9.err; tail -2 $O/c9.err
python - <<'PY'
import json
for l in open("gpurun_out/c9_strong.jsonl"):
    d = json.loads(l)
    print(d["workload"], d["total_envs"], "%.5f ms" % d["ms_per_step"], d["two_stream_overlap"], d["ring_batches_per_gpu"])
PY
