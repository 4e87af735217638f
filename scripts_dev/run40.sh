python -m pytest tests -m gpu -q -x > gpurun_out/t40_pytest.log 2>&1
for w in 0 296 148 592 1184; do
VC_WIDE_MAX=$w python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz > gpurun_out/t40_bench_w$w.log 2>&1
done
true
