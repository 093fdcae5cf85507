#!/bin/bash
# full GPU suite + the driver's two bench commands on the current build
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/c6_tests.log 2>&1; tail -3 $O/c6_tests.log
python bench.py --impl reference --steps 20 --warmup 5 > $O/c6_bench_ref.json 2> $O/c6_bench_ref.err; echo "ref rc=$?"
( time python bench.py --steps 20 --warmup 5 > $O/c6_bench.json 2> $O/c6_bench.err ) 2> $O/c6_bench_time.txt; echo "bench rc=$?"; tail -3 $O/c6_bench_time.txt
python - <<'PY'
import json
d = json.load(open("gpurun_out/c6_bench.json"))
print("value %.1f M frac %.3f e2e %.1f M" % (d["value"] / 1e6, d["roofline"]["frac"], d["e2e"]["value"] / 1e6))
for k, v in d["strong"].items():
    print("strong", k, "%.4f ms" % v["ms_per_step"], "%.3f" % v["hbm_frac_algorithmic_per_gpu"])
for k, v in d["configs"].items():
    print(k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a in ("ms_per_step", "hbm_frac_algorithmic", "hbm_frac", "alu_pipe_frac", "ms_per_call", "episodes_per_s", "seconds", "env_steps_per_s", "us_single_board")})
PY
