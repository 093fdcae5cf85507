#!/bin/bash
# ncu of the new 256x256 step kernel + A/B of warps per CTA of the 26x26 step kernel (alternative builds)
O=gpurun_out
P="python scripts/profile_kernels.py"
NF="ncu --set full --clock-control none --import-source on"
python -m pytest tests/test_gpu_round2.py -x -q -m gpu -k "sliding or big_board" > $O/c4_parity.log 2>&1; tail -1 $O/c4_parity.log
$P big_256 > $O/c4_prof_plain.log 2>&1 && $NF -k regex:big_rows_step -s 3 -c 2 -o $O/r02_bigstep $P big_256 > $O/c4_ncu.log 2>&1
python profiles/tools/ncu_summary.py $O/r02_bigstep.ncu-rep > $O/r02_bigstep_ncu.txt 2>&1
python profiles/tools/line_profile.py $O/r02_bigstep.ncu-rep big_rows_step safelife_k2_b200/libsafelife_b200.so > $O/r02_bigstep_lines.txt 2>&1
head -14 $O/r02_bigstep_ncu.txt
C="step:mixed:8192 step:mixed:16384 step:prune-spawn:16384 step:append-still:4096 step:mixed:65536"
(echo '{"lib": "default (4 warps per CTA)"}'; python bench_configs.py $C
 echo '{"lib": "2 warps per CTA"}'; SLB_LIB_PATH=scripts/_alt/libsafelife_b200_w2.so python bench_configs.py $C
 echo '{"lib": "1 warp per CTA"}'; SLB_LIB_PATH=scripts/_alt/libsafelife_b200_w1.so python bench_configs.py $C) > $O/c4_ab_warps.jsonl 2>&1
cut -c1-140 $O/c4_ab_warps.jsonl
