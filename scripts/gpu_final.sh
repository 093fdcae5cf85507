#!/bin/bash
# what the driver runs at round end, on the current build: GPU suite, smoke(), both bench arms
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/g_tests.log 2>&1; tail -2 $O/g_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/g_smoke.log 2>&1; tail -1 $O/g_smoke.log
python bench.py --impl reference --steps 20 --warmup 5 > $O/g_bench_ref.json 2> $O/g_bench_ref.err; echo "ref rc=$?"
( time python bench.py --steps 20 --warmup 5 > $O/g_bench.json 2> $O/g_bench.err ) 2> $O/g_time.txt; echo "bench rc=$?"; tail -3 $O/g_time.txt
python - <<'PY'
import json
d = json.load(open("gpurun_out/g_bench.json"))
e = d["e2e"]
print("value %.1f M frac %.3f e2e %.1f M (one-sample spread %s) serial %.1f M two-stream %.2f G" % (d["value"] / 1e6, d["roofline"]["frac"], e["value"] / 1e6, ["%.0f" % (v / 1e6) for v in e["segment_values"]], e["serial_value"] / 1e6, d["two_stream_overlap"]["value"] / 1e9))
r = json.load(open("gpurun_out/g_bench_ref.json")); print("reference %.3f M on %d threads" % (r["value"] / 1e6, r["cpu_baseline"]["cores"]))
for k, v in d["strong"].items(): print("strong", k, "%.5f ms" % v["ms_per_step"], v["two_stream_overlap"])
c = d["configs"]; print("big %.4f ms %.3f | ps %.1f M alu %.3f | side %.4f s emd %.3f s" % (c["big_256"]["ms_per_step"], c["big_256"]["hbm_frac_algorithmic"], c["prune_spawn_16384"]["env_steps_per_s"] / 1e6, c["prune_spawn_16384"]["alu_pipe_frac"], c["side_effects"]["seconds"], c["side_effects"]["emd_seconds"]))
PY
