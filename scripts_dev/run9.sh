python -m pytest tests -m gpu -x -q > gpurun_out/t9_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/t9_pytest.log
VC_PASS_TRACE=1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t9_bench_base.log 2>&1
VC_GIVEUP=6 VC_MAXDAMP=2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t9_bench_g6.log 2>&1
VC_RESTORE=3 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t9_bench_r3.log 2>&1
true
