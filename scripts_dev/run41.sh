python -m pytest tests -m gpu -q > gpurun_out/t41_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/t41_pytest.log
python __graft_entry__.py --smoke > gpurun_out/t41_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/t41_smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/t41_bench_nmpc.log 2>&1
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz"
$CMD > gpurun_out/t41_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r01_launches_v3.csv $CMD > gpurun_out/t41_ncu.log 2>&1
true
