#!/bin/bash
# A/B: 3 CTAs per SM (80 registers) for the 256x256 step kernel; 2 warps per CTA for the 26x26 kernels incl. e2e
O=gpurun_out
T="tests/test_gpu_round2.py::test_big_board_fused_step_at_length tests/test_gpu_round2.py::test_sliding_window_step_every_row_width tests/test_gpu_parity.py::test_full_size_256x256_4096"
python -m pytest $T -x -q -m gpu > $O/c5_parity_default.log 2>&1; echo "default: $(tail -1 $O/c5_parity_default.log)"
SLB_LIB_PATH=scripts/_alt/libsafelife_b200_c3.so python -m pytest $T -x -q -m gpu > $O/c5_parity_c3.log 2>&1; echo "c3: $(tail -1 $O/c5_parity_c3.log)"
(echo '{"lib": "default: 2 CTAs per SM, 128 registers"}'; python bench_configs.py big:4096 big:512 big:1024
 echo '{"lib": "3 CTAs per SM, 80 registers"}'; SLB_LIB_PATH=scripts/_alt/libsafelife_b200_c3.so python bench_configs.py big:4096 big:512 big:1024) > $O/c5_ab_ctas.jsonl 2>&1
cut -c1-150 $O/c5_ab_ctas.jsonl
python bench.py --steps 20 --warmup 5 --headline-only --no-cpu > $O/c5_bench_w4.json 2> $O/c5_bench_w4.err
cp scripts/_alt/libsafelife_b200_w2.so safelife_k2_b200/libsafelife_b200.so
python bench.py --steps 20 --warmup 5 --headline-only --no-cpu > $O/c5_bench_w2.json 2> $O/c5_bench_w2.err
python - <<'PY'
import json
for w in ("w4", "w2"):
    d = json.load(open("gpurun_out/c5_bench_%s.json" % w))
    print(w, "value %.1f M  e2e %.1f M  serial %.1f M  single %.1f M" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6, d["e2e"]["serial_value"] / 1e6, d["single_batch"]["value"] / 1e6))
PY
