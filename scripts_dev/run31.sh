python scripts_dev/fail_cases.py cascaded_obstacles_shoe:cascaded race1_shoe:singletrack,cascaded > gpurun_out/t31_fail.log 2>&1
python -m pytest tests/test_variants.py -m gpu -q -s > gpurun_out/t31_variants.log 2>&1
true
