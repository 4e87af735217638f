python -m pytest tests -m gpu -q -x > gpurun_out/t21_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/t21_pytest.log
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t21_bench_split.log 2>&1
VC_SPLIT=0 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t21_bench_nosplit.log 2>&1
true
