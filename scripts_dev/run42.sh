python -m pytest tests -m gpu -q -x > gpurun_out/t42_pytest.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz > gpurun_out/t42_bench_new.log 2>&1
VC_LIB_PATH=$PWD/vehicle_control_b200/variants/pre_update.so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz > gpurun_out/t42_bench_old.log 2>&1
true
