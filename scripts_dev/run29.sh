python -m pytest tests -m gpu -q > gpurun_out/t29_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/t29_pytest.log
python __graft_entry__.py --smoke > gpurun_out/t29_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/t29_smoke.log
python bench.py --steps 5 --warmup 3 > gpurun_out/t29_bench_nmpc.log 2>&1
python bench.py --workload rollout --steps 5 --warmup 3 > gpurun_out/t29_bench_rollout.log 2>&1
python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/t29_bench_ref.log 2>&1
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --lap-cache gpurun_out/lap_cache.npz"
$CMD > gpurun_out/t29_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r01_launches_v2.csv $CMD > gpurun_out/t29_ncu.log 2>&1
true
