#!/bin/bash
# usage: robust_suite.sh  -> KAT summary + laps with/without obstacles (env vars pass through)
python scripts_dev/nmpc_debug.py tick 2>&1 | grep -E "tick[0-9] " | sed -E 's/oracle obj.*//' | awk '{print $1, $3,$4,$5,$6,$7,$8}' 
for o in 0 1; do python scripts_dev/lap_debug.py $o 35 2>&1 | head -n 12 | grep -v "^\[vc\]" | cut -c1-200; done
