python -m pytest tests -m gpu -x -q > gpurun_out/t8_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/t8_pytest.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz > gpurun_out/t8_bench_base.log 2>&1
python bench.py --workload rollout --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/t8_bench_rollout.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --batch 16384 --no-cpu-baseline --lap-cache /tmp/lap.npz"
$CMD > gpurun_out/t8_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_step -s 8 -c 1 -o gpurun_out/r01_k_step_v2 $CMD > gpurun_out/t8_ncu1.log 2>&1
true
