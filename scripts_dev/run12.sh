python scripts_dev/c4_debug.py 35 > gpurun_out/t12_c4.log 2>&1
VC_GIVEUP=100000 VC_MAXDAMP=8 python scripts_dev/c4_debug.py 35 > gpurun_out/t12_c4_old.log 2>&1
VC_RESTORE=0 VC_STALL=4 VC_GIVEUP=100000 VC_MAXDAMP=8 python scripts_dev/c4_debug.py 35 > gpurun_out/t12_c4_older.log 2>&1
python scripts_dev/c4_debug.py 25 > gpurun_out/t12_c4_25.log 2>&1
true
