VC_PASS_TRACE=1 VC_PROFILE=1 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz > gpurun_out/t26_trace.log 2>&1
true
