python scripts/lap_parity_all.py --out gpurun_out/lap_parity_retry.json cascaded_obstacles_shoe race1_shoe race_obstacles_shoe cascaded_giantObstacle1_ippodromo > gpurun_out/t33_laps.log 2>&1
python -m pytest tests -m gpu -q -x > gpurun_out/t33_pytest.log 2>&1
true
