VC_PASS_TRACE=1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t17_bench_base.log 2>&1
for v in mb3 mb4; do
  VC_LIB_PATH=$PWD/vehicle_control_b200/variants/libvc_$v.so python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t17_bench_$v.log 2>&1
done
CMD="python bench.py --steps 1 --warmup 3 --batch 16384 --no-cpu-baseline --lap-cache /tmp/lap.npz"
$CMD > gpurun_out/t17_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_eval_st -s 8 -c 1 -o gpurun_out/r01_k_eval_st_v2 $CMD > gpurun_out/t17_ncu.log 2>&1
true
