for k in 1 2 4; do
  VC_JAC_K=$k python bench.py --workload rollout --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('K=$k', round(d['value']/1e6,1), 'M state-steps/s', round(d['ms_per_step'],2), 'ms', round(d['roofline']['frac'],4))" >> gpurun_out/t24_jac.log
done
VC_JAC_K=2 python -m pytest tests/test_gpu_dynamics.py tests/test_gpu_batch.py -m gpu -q -k "rollout or c5" > gpurun_out/t24_pytest_k2.log 2>&1
true
