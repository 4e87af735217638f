CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz"
$CMD > gpurun_out/t48_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_step_b$ -s 3 -c 1 -o gpurun_out/r01_k_step_b_final -f $CMD > gpurun_out/t48_ncu_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_update -s 3 -c 1 -o gpurun_out/r01_k_update -f $CMD > gpurun_out/t48_ncu_u.log 2>&1
true
