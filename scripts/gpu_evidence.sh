#!/bin/bash
# final-build evidence: GPU suite, the launch list of the driver's bench command, ncu --set full of the headline
# step kernel, of the prune-spawn step and of the 256x256 step kernel (each after a plain run of the same command)
O=gpurun_out
B="python bench.py --steps 20 --warmup 5 --headline-only --no-cpu"
P="python scripts/profile_kernels.py"
NF="ncu --set full --clock-control none --import-source on"
python -m pytest tests -x -q -m gpu > $O/e_tests.log 2>&1; tail -2 $O/e_tests.log
$B > $O/e_bench_plain.json 2> $O/e_bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/e_launches.csv $B > $O/e_ncu_list.log 2>&1
$B > /dev/null 2>&1 && $NF -k regex:res_env_rollout -s 40 -c 3 -o $O/e_step $B > $O/e_ncu_step.log 2>&1
$P step_prune_spawn big_256 > $O/e_prof_plain.log 2>&1 && $NF -k regex:big_rows_step -s 3 -c 2 -o $O/e_bigstep $P big_256 > $O/e_ncu_big.log 2>&1
$P step_prune_spawn > /dev/null 2>&1 && $NF -k regex:res_env_rollout -s 4 -c 2 -o $O/e_ps $P step_prune_spawn > $O/e_ncu_ps.log 2>&1
LIB=safelife_k2_b200/libsafelife_b200.so
for r in step bigstep ps; do
  python profiles/tools/ncu_summary.py $O/e_$r.ncu-rep > $O/e_${r}_ncu.txt 2>&1
done
python profiles/tools/line_profile.py $O/e_step.ncu-rep res_env_rollout $LIB > $O/e_step_lines.txt 2>&1
python profiles/tools/line_profile.py $O/e_ps.ncu-rep res_env_rollout $LIB > $O/e_ps_lines.txt 2>&1
python profiles/tools/line_profile.py $O/e_bigstep.ncu-rep big_rows_step $LIB > $O/e_bigstep_lines.txt 2>&1
rm -f $O/e_step.ncu-rep $O/e_bigstep.ncu-rep
head -12 $O/e_step_ncu.txt; head -12 $O/e_bigstep_ncu.txt
