#!/bin/bash
# One gpurun call: the no-alloc test alone and after the whole suite, then the ncu evidence of profiles/r02
# (each ncu command directly after a plain run of the same command line that exited 0; summaries are made on
# the box, the big .ncu-rep files stay there except the 256x256 one).
O=gpurun_out
B="python bench.py --steps 20 --warmup 5 --headline-only --no-cpu"
P="python scripts/profile_kernels.py"
NF="ncu --set full --clock-control none --import-source on"
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/z_alone.log 2>&1; tail -3 $O/z_alone.log
python -m pytest tests -x -q -m gpu > $O/z_all.log 2>&1; tail -4 $O/z_all.log
# A/B: batches of several waves cut into wave-sized launches (SLB_RES_CHUNK), no profiler
(python bench_configs.py step:mixed:4096 step:mixed:8192 step:mixed:16384 step:prune-spawn:16384 step:mixed:65536;
 SLB_RES_CHUNK=4096 python bench_configs.py step:mixed:8192 step:mixed:16384 step:prune-spawn:16384 step:mixed:65536 step:append-still:8192;
 SLB_RES_CHUNK=-1 python bench_configs.py step:mixed:8192 step:mixed:16384 step:prune-spawn:16384;
 SLB_RES_CHUNK=2048 python bench_configs.py step:mixed:8192 step:mixed:16384 step:prune-spawn:16384) > $O/ab2_chunk.jsonl 2>&1
cut -c1-200 $O/ab2_chunk.jsonl
$B > $O/p_bench_plain.json 2> $O/p_bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches.csv $B > $O/p_ncu_list.log 2>&1
$B > /dev/null 2>&1 && $NF -k regex:res_env_rollout -s 40 -c 3 -o $O/r02_step $B > $O/p_ncu_step.log 2>&1
$P > $O/p_prof_plain.log 2>&1 && $NF -k regex:big_rows_env -s 3 -c 2 -o $O/r02_bigrows $P > $O/p_ncu_bigrows.log 2>&1
$P > /dev/null 2>&1 && $NF -k regex:advance_pairs -s 1 -c 6 -o $O/r02_pairs $P > $O/p_ncu_pairs.log 2>&1
$P > /dev/null 2>&1 && $NF -k regex:res_env_rollout -s 4 -c 2 -o $O/r02_ps $P > $O/p_ncu_ps.log 2>&1
$P > /dev/null 2>&1 && $NF -k regex:side_effect_fast -c 2 -o $O/r02_side $P > $O/p_ncu_side.log 2>&1
$P > /dev/null 2>&1 && $NF -k regex:emd_solve -c 1 -o $O/r02_emd $P > $O/p_ncu_emd.log 2>&1
SLB_BIG_ROWS=0 $P big_256 > /dev/null 2>&1 && SLB_BIG_ROWS=0 $NF -k regex:big_env_kernel -s 3 -c 2 -o $O/r02_bigold $P big_256 > $O/p_ncu_bigold.log 2>&1
tail -2 $O/p_prof_plain.log
LIB=safelife_k2_b200/libsafelife_b200.so
for r in step bigrows pairs ps side emd bigold; do
  if [ -f $O/r02_$r.ncu-rep ]; then
    python profiles/tools/ncu_summary.py $O/r02_$r.ncu-rep > $O/r02_${r}_ncu.txt 2>&1
    ls -la $O/r02_$r.ncu-rep
  else echo "no report: $r"; tail -3 $O/p_ncu_$r.log; fi
done
python profiles/tools/line_profile.py $O/r02_bigrows.ncu-rep big_rows_env $LIB > $O/r02_bigrows_lines.txt 2>&1
python profiles/tools/line_profile.py $O/r02_bigold.ncu-rep big_env_kernel $LIB > $O/r02_bigold_lines.txt 2>&1
python profiles/tools/line_profile.py $O/r02_pairs.ncu-rep advance_pairs $LIB > $O/r02_pairs_lines.txt 2>&1
python profiles/tools/line_profile.py $O/r02_step.ncu-rep res_env_rollout $LIB > $O/r02_step_lines.txt 2>&1
python profiles/tools/line_profile.py $O/r02_emd.ncu-rep emd_solve $LIB > $O/r02_emd_lines.txt 2>&1
rm -f $O/r02_step.ncu-rep $O/r02_pairs.ncu-rep $O/r02_ps.ncu-rep $O/r02_side.ncu-rep $O/r02_emd.ncu-rep $O/r02_bigold.ncu-rep
head -12 $O/r02_bigrows_ncu.txt
