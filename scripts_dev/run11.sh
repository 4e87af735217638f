python -m pytest tests -m gpu -x -q > gpurun_out/t11_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/t11_pytest.log
python bench.py --steps 3 --warmup 3 --lap-cache /tmp/lap.npz > gpurun_out/t11_bench_nmpc.log 2>&1
for v in ev2 ev3 ev4; do
  VC_LIB_PATH=$PWD/vehicle_control_b200/variants/libvc_$v.so python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t11_bench_$v.log 2>&1
done
CMD="python bench.py --steps 1 --warmup 3 --batch 16384 --no-cpu-baseline --lap-cache /tmp/lap.npz"
$CMD > gpurun_out/t11_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/t11_launches.csv $CMD > gpurun_out/t11_ncu.log 2>&1
true
