#!/bin/bash
# One gpurun call: parity of every compiled variant of the 256x256 step kernel (SLB_BIG_OPT), then A/B timing.
O=gpurun_out
T="tests/test_gpu_round2.py::test_big_board_fused_step_at_length tests/test_gpu_round2.py::test_sliding_window_step_every_row_width tests/test_gpu_parity.py::test_full_size_256x256_4096"
for v in ${VARIANTS:-31 16 17 19 23 29}; do
  SLB_BIG_OPT=$v python -m pytest $T -x -q -m gpu > $O/big_parity_$v.log 2>&1; echo "opt=$v parity: $(tail -1 $O/big_parity_$v.log)"
done
rm -f $O/big_ab.jsonl
for v in 0 ${VARIANTS:-31 16 17 19 23 29}; do
  echo "{\"SLB_BIG_OPT\": $v}" >> $O/big_ab.jsonl
  SLB_BIG_OPT=$v python bench_configs.py big:4096 big:512 >> $O/big_ab.jsonl 2>&1
done
cut -c1-150 $O/big_ab.jsonl
