for v in b_mb3 b_mb4; do
VC_LIB_PATH=$PWD/vehicle_control_b200/variants/$v.so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz > gpurun_out/t28_$v.log 2>&1
done
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz > gpurun_out/t28_base.log 2>&1
true
