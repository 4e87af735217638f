#!/bin/bash
# Times tuning builds against each other on one box: for every tmp_data/libs/lib_<name>.so (VC_LIB_PATH) the per-kernel
# times of full-load passes (scripts/eval_time_probe.py) and the bench-parity tests.  usage: variants.sh name...
mkdir -p gpurun_out
for v in "$@"; do
  lib=$PWD/tmp_data/libs/lib_$v.so
  VC_LIB_PATH=$lib timeout 120 python scripts/eval_time_probe.py 65536 2> gpurun_out/var_$v.trace > /dev/null
  t=$(grep "pass   [2-4] " gpurun_out/var_$v.trace | awk '{a+=$5; b+=$7; c+=$9; n++} END {printf "eval_st %.3f eval_pm %.3f step %.3f ms (mean of %d full passes)", a/n, b/n, c/n, n}')
  p=$(VC_LIB_PATH=$lib timeout 120 python -m pytest tests/test_gpu_bench_parity.py -q -m gpu 2>&1 | tail -1)
  echo "$v: $t | $p" | tee -a gpurun_out/variants.log
done
