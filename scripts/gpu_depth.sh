#!/bin/bash
O=gpurun_out
python -m pytest tests -x -q -m gpu -k "emd or no_alloc or side_effect_score" > $O/c23_tests.log 2>&1; tail -2 $O/c23_tests.log
SLB_EMD_BITS=0 python -m pytest tests -x -q -m gpu -k "emd or no_alloc" > $O/c23_tests_nobits.log 2>&1; tail -1 $O/c23_tests_nobits.log
(echo '{"variant": "cached minimum + flow bits"}'; python bench_configs.py side_effects
 echo '{"variant": "cached minimum only (SLB_EMD_BITS=0)"}'; SLB_EMD_BITS=0 python bench_configs.py side_effects) > $O/c23_emd.jsonl 2>&1
python scripts/profile_kernels.py emd 2>&1 | tail -2
SLB_EMD_BITS=0 python scripts/profile_kernels.py emd 2>&1 | tail -2
python - <<'PY'
import json
for l in open("gpurun_out/c23_emd.jsonl"):
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    print({k: d[k] for k in d if k in ("variant", "seconds", "emd_seconds", "emd_problems", "emd_max")})
PY
