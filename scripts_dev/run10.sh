python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t10_bench.log 2>&1
python scripts_dev/slow_cases.py 8192 > gpurun_out/t10_slow.log 2> gpurun_out/t10_slow_trace.log
true
