python -m pytest tests -m gpu -q > gpurun_out/t15_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/t15_pytest.log
python __graft_entry__.py --smoke > gpurun_out/t15_smoke.log 2>&1
python bench.py --steps 3 --warmup 3 > gpurun_out/t15_bench_nmpc.log 2>&1
python bench.py --workload rollout --steps 3 --warmup 3 > gpurun_out/t15_bench_rollout.log 2>&1
true
