CMD="python bench.py --workload rollout --steps 1 --warmup 0 --no-cpu-baseline"
$CMD > gpurun_out/t43_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_rollout_st_jac -s 1 -c 1 -o gpurun_out/r01_k_rollout_jac_k2 -f $CMD > gpurun_out/t43_ncu.log 2>&1
true
