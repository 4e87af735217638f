LAPS="cascaded3_ippodromo/cascaded singletrack3_ippodromo/singletrack cascaded_obstacles1_ippodromo/cascaded cascaded_giantObstacle3_ippodromo/cascaded race2_shoe/cascaded cascaded1_ippodromo/cascaded singletrack_shoe/singletrack"
CASES="c3_194 st3_98 obs1_165 giant3_68 giant3_0 shoe2_134"
run() {  # name, env...
  name=$1; shift
  echo "######## $name: $@" >> gpurun_out/t5_robust.log
  env "$@" python scripts_dev/case_debug.py $CASES 2>&1 | grep "==== .*status" | cut -c1-200 >> gpurun_out/t5_robust.log
  env "$@" python scripts_dev/lap_compare.py $LAPS 2>&1 | grep -v "Warn\|^NMPC\|^   tick" | cut -c1-260 >> gpurun_out/t5_robust.log
}
rm -f gpurun_out/t5_robust.log
run zinit VC_STALL=100000 VC_RESTORE=4 VC_ZINIT=0.001
run zinit_nosoc VC_STALL=100000 VC_RESTORE=4 VC_ZINIT=0.001 VC_MAXSOC=0
run zcap1 VC_STALL=100000 VC_RESTORE=4 VC_ZCAP=1
run restore3_soc1 VC_STALL=100000 VC_RESTORE=3 VC_MAXSOC=1
true
