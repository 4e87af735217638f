python scripts_dev/rungs_compare.py > gpurun_out/t22_rungs.log 2>&1
LAPS="cascaded3_ippodromo/cascaded singletrack3_ippodromo/singletrack cascaded_obstacles1_ippodromo/cascaded cascaded_giantObstacle3_ippodromo/cascaded race2_shoe/cascaded cascaded1_ippodromo/cascaded singletrack_shoe/singletrack"
for r in 2 3; do
  echo "#### warm rungs $r" >> gpurun_out/t22_laps.log
  VC_WARM_RUNGS=$r python scripts_dev/lap_compare.py $LAPS 2>&1 | grep -v "Warn\|^NMPC\|^   tick" | cut -c1-230 >> gpurun_out/t22_laps.log
done
for r in 0 2 3; do
VC_MU0_RUNGS=$r python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap.npz 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']; print('bench rungs $r', round(d['value']), round(d['ms_per_step'],1), c['status_counts'], round(c['mean_iters'],2), c['iters_percentiles'], c['passes'])" >> gpurun_out/t22_laps.log
done
true
