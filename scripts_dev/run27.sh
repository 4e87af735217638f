CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz"
$CMD > gpurun_out/t27_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_step_a -s 3 -c 1 -o gpurun_out/r01_k_step_a -f $CMD > gpurun_out/t27_ncu_a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_step_b -s 3 -c 1 -o gpurun_out/r01_k_step_b -f $CMD > gpurun_out/t27_ncu_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_eval_st -s 3 -c 1 -o gpurun_out/r01_k_eval_st_v3 -f $CMD > gpurun_out/t27_ncu_e.log 2>&1
true
