python -m pytest tests -m gpu -q -x > gpurun_out/t18_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/t18_pytest.log
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t18_bench.log 2>&1
true
