#!/bin/bash
# Round-2 profiling recipe (run under gpurun on one B200; outputs in gpurun_out/, summaries are copied to profiles/).
#   tmp_data/lap_cache.npz = warm-start lap (written by bench.py --lap-cache on an earlier run), so that the untimed lap replay is skipped.
set -u
B="python bench.py --steps 2 --warmup 3 --no-secondary --no-cpu-baseline --in-flight 1 --lap-cache tmp_data/lap_cache.npz"
$B > gpurun_out/r02_prof_plain.json 2> gpurun_out/r02_prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/r02_launches.csv $B > /dev/null 2>&1
for k in k_eval_st k_eval_pm 'k_step_a$' 'k_step_b$' k_update; do
  n=$(echo $k | tr -d '$')
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -o gpurun_out/r02_$n $B > gpurun_out/r02_ncu_$n.log 2>&1
  ncu -i gpurun_out/r02_$n.ncu-rep --page raw --csv > gpurun_out/r02_${n}_raw.csv 2> /dev/null
  case $n in k_eval_pm|k_update) rm -f gpurun_out/r02_$n.ncu-rep;; esac    # gpurun brings back at most 64 MiB
done
for k in k_step_a_wide k_step_b_wide; do
  ncu --set full --clock-control none -k regex:$k -s 150 -c 1 -o gpurun_out/r02_$k $B > gpurun_out/r02_ncu_$k.log 2>&1
  ncu -i gpurun_out/r02_$k.ncu-rep --page raw --csv > gpurun_out/r02_${k}_raw.csv 2> /dev/null
  rm -f gpurun_out/r02_$k.ncu-rep
done
R="python bench.py --workload rollout --steps 1 --warmup 3 --rollout-batch 262144 --no-cpu-baseline"
$R > gpurun_out/r02_prof_rollout.json 2> /dev/null && \
ncu --set full --clock-control none -k regex:k_rollout_st_jac -s 2 -c 1 -o gpurun_out/r02_k_rollout $R > gpurun_out/r02_ncu_rollout.log 2>&1
ncu -i gpurun_out/r02_k_rollout.ncu-rep --page raw --csv > gpurun_out/r02_k_rollout_raw.csv 2> /dev/null
rm -f gpurun_out/r02_k_rollout.ncu-rep
du -sh gpurun_out; ls -la gpurun_out/r02_*
