LAPS="cascaded3_ippodromo/cascaded singletrack3_ippodromo/singletrack cascaded_obstacles1_ippodromo/cascaded cascaded_giantObstacle3_ippodromo/cascaded race2_shoe/cascaded cascaded1_ippodromo/cascaded singletrack_shoe/singletrack"
CASES="c3_194 st3_98 obs1_165 giant3_68 shoe2_134"
run() {
  name=$1; shift
  echo "######## $name: $@" >> gpurun_out/t14.log
  env "$@" python scripts_dev/c4_debug.py 35 2>&1 | grep "^tick" | head -2 | cut -c1-200 >> gpurun_out/t14.log
  env "$@" python scripts_dev/case_debug.py $CASES 2>&1 | grep "==== .*status" | cut -c1-150 >> gpurun_out/t14.log
  env "$@" python scripts_dev/lap_compare.py $LAPS 2>&1 | grep -v "Warn\|^NMPC\|^   tick" | cut -c1-200 >> gpurun_out/t14.log
}
rm -f gpurun_out/t14.log
run default
run nogiveup VC_GIVEUP=100000
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --lap-cache /tmp/lap_def.npz 2>&1 | tail -1 > gpurun_out/t14_bench.log
true
