#!/bin/bash
O=gpurun_out
python -m pytest tests/test_gpu_round2.py -x -q -m gpu -k "two_batches" > $O/c14_tests.log 2>&1; tail -3 $O/c14_tests.log
(echo '{"SLB_SIDE_WARPS": 4}'; python bench_configs.py side_effects
 echo '{"SLB_SIDE_WARPS": 2}'; SLB_SIDE_WARPS=2 python bench_configs.py side_effects
 echo '{"SLB_SIDE_WARPS": 1}'; SLB_SIDE_WARPS=1 python bench_configs.py side_effects) > $O/c14_side_ab.jsonl 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/c14_side_ab.jsonl"):
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    print({k: d[k] for k in d if k in ("SLB_SIDE_WARPS", "seconds", "episodes_per_s", "emd_seconds")})
PY
