python -m pytest tests -m gpu -q -x > gpurun_out/t36_pytest.log 2>&1
python scripts/lap_parity_all.py --out gpurun_out/lap_parity_all.json > gpurun_out/t36_laps.log 2>&1
python scripts/c4_batched_lap.py 16384 1200 > gpurun_out/t36_c4.json 2> gpurun_out/t36_c4.log
true
