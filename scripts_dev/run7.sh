python -m pytest tests -m gpu -x -q > gpurun_out/t7_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/t7_pytest.log
VC_PASS_TRACE=1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t7_bench_base.log 2>&1
for v in mb4 ev3 ev4; do
  VC_LIB_PATH=$PWD/vehicle_control_b200/variants/libvc_$v.so python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t7_bench_$v.log 2>&1
done
true
