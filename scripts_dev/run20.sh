CMD="python bench.py --steps 1 --warmup 3 --batch 16384 --no-cpu-baseline --lap-cache /tmp/lap.npz"
$CMD > gpurun_out/t20_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_step -s 8 -c 1 -o gpurun_out/r01_k_step_v3 $CMD > gpurun_out/t20_ncu.log 2>&1
true
