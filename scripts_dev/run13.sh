LAPS="cascaded3_ippodromo/cascaded singletrack3_ippodromo/singletrack cascaded_obstacles1_ippodromo/cascaded cascaded_giantObstacle3_ippodromo/cascaded race2_shoe/cascaded singletrack_shoe/singletrack"
CASES="c3_194 st3_98 obs1_165 giant3_68 shoe2_134"
run() {
  name=$1; shift
  echo "######## $name: $@" >> gpurun_out/t13.log
  env "$@" python scripts_dev/c4_debug.py 35 2>&1 | grep "^tick" | cut -c1-200 >> gpurun_out/t13.log
  env "$@" python scripts_dev/case_debug.py $CASES 2>&1 | grep "==== .*status" | cut -c1-150 >> gpurun_out/t13.log
  env "$@" python scripts_dev/lap_compare.py $LAPS 2>&1 | grep -v "Warn\|^NMPC\|^   tick" | cut -c1-200 >> gpurun_out/t13.log
  env "$@" python bench.py --steps 1 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap_$name.npz 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']; print('bench', round(d['value']), round(d['ms_per_step'],1), c['status_counts'], round(c['mean_iters'],2), c['iters_percentiles'], c['lap_ticks'])" >> gpurun_out/t13.log
}
rm -f gpurun_out/t13.log
run A_restore4_check VC_RESTORE=4
run B_stall4_restore8_check VC_STALL=4 VC_RESTORE=8
run C_stall4_restore4_check VC_STALL=4 VC_RESTORE=4
true
