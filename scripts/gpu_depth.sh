#!/bin/bash
O=gpurun_out
(echo '{"variant": "4 CTAs/SM, kfirst 8 (default)"}'; python bench_configs.py side_effects
 echo '{"variant": "4 CTAs/SM, kfirst 6"}'; SLB_SIDE_KFIRST=6 python bench_configs.py side_effects
 echo '{"variant": "5 CTAs/SM (89 regs), kfirst 6"}'; SLB_SIDE_KFIRST=6 SLB_LIB_PATH=scripts/_alt/libsafelife_b200_side5.so python bench_configs.py side_effects
 echo '{"variant": "5 CTAs/SM (89 regs), kfirst 8"}'; SLB_LIB_PATH=scripts/_alt/libsafelife_b200_side5.so python bench_configs.py side_effects
 echo '{"variant": "5 CTAs/SM (89 regs), kfirst 7"}'; SLB_SIDE_KFIRST=7 SLB_LIB_PATH=scripts/_alt/libsafelife_b200_side5.so python bench_configs.py side_effects) > $O/c22_side.jsonl 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/c22_side.jsonl"):
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    print({k: d[k] for k in d if k in ("variant", "seconds", "episodes_per_s", "max_keys")})
PY
SLB_SIDE_KFIRST=6 SLB_LIB_PATH=scripts/_alt/libsafelife_b200_side5.so python -m pytest tests -x -q -m gpu -k "side_effect or soak" > $O/c22_tests.log 2>&1; tail -2 $O/c22_tests.log
