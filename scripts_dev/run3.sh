VC_DEBUG=1 python scripts_dev/case_debug.py c3_194 st3_98 obs1_165 giant3_68 giant3_0 shoe2_134 > gpurun_out/t3_cases.log 2>&1
VC_STALL=100000 python scripts_dev/case_debug.py c3_194 st3_98 obs1_165 giant3_68 giant3_0 shoe2_134 2>&1 | grep "====" > gpurun_out/t3_cases_nostall.log
python -m pytest tests -m gpu -x -q -k "not full_lap" > gpurun_out/t3_pytest.log 2>&1
VC_PASS_TRACE=1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/t3_bench.log 2>&1
true
