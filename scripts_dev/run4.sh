LAPS="cascaded3_ippodromo/cascaded singletrack3_ippodromo/singletrack cascaded_obstacles1_ippodromo/cascaded cascaded_giantObstacle3_ippodromo/cascaded race2_shoe/cascaded cascaded1_ippodromo/cascaded singletrack_shoe/singletrack"
CASES="c3_194 st3_98 obs1_165 giant3_68 giant3_0 shoe2_134"
run() {  # name, env...
  name=$1; shift
  echo "######## $name: $@" >> gpurun_out/t4_robust.log
  env "$@" python scripts_dev/case_debug.py $CASES 2>&1 | grep "==== .*status" >> gpurun_out/t4_robust.log
  env "$@" python scripts_dev/lap_compare.py $LAPS 2>&1 | grep -v "Warn\|^NMPC" >> gpurun_out/t4_robust.log
}
rm -f gpurun_out/t4_robust.log
run nostall VC_STALL=100000
run restore4 VC_STALL=100000 VC_RESTORE=4
run restore8 VC_STALL=100000 VC_RESTORE=8
run restore4_nosoc VC_STALL=100000 VC_RESTORE=4 VC_MAXSOC=0
run stall4_restore8 VC_STALL=4 VC_RESTORE=8
true
