timeout 2400 python scripts/lap_parity_all.py --out gpurun_out/lap_parity_all.json > gpurun_out/t30_laps.log 2>&1
true
