python scripts_dev/lap_compare.py cascaded3_ippodromo/cascaded singletrack3_ippodromo/singletrack cascaded_obstacles1_ippodromo/cascaded cascaded_giantObstacle3_ippodromo/cascaded race2_shoe/cascaded cascaded1_ippodromo/cascaded > gpurun_out/t2_lapcmp.log 2>&1
python scripts_dev/lap_debug.py 1 35 > gpurun_out/t2_lapdebug.log 2>&1
VC_PASS_TRACE=1 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/t2_passtrace.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/t2_launches.csv python bench.py --steps 1 --warmup 3 --batch 16384 --no-cpu-baseline > gpurun_out/t2_ncu.log 2>&1
true
