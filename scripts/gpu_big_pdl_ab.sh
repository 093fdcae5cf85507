#!/bin/bash
O=gpurun_out
T="tests/test_gpu_round2.py::test_big_board_fused_step_at_length tests/test_gpu_round2.py::test_sliding_window_step_every_row_width tests/test_gpu_parity.py::test_full_size_256x256_4096"
for v in 23 55; do
  SLB_BIG_OPT=$v python -m pytest $T -x -q -m gpu > $O/c11_parity_$v.log 2>&1; echo "opt=$v parity: $(tail -1 $O/c11_parity_$v.log)"
done
rm -f $O/c11_ab.jsonl
for v in 23 55 23 55; do
  echo "{\"SLB_BIG_OPT\": $v}" >> $O/c11_ab.jsonl
  SLB_BIG_OPT=$v python bench_configs.py big:4096 big:512 big:1024 >> $O/c11_ab.jsonl 2>&1
done
cut -c1-140 $O/c11_ab.jsonl
