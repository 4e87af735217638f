#!/bin/bash
# tuning build of the library with extra nvcc flags: build_variant.sh name [-D...]  -> tmp_data/libs/lib_<name>.so
name=$1; shift
mkdir -p tmp_data/libs /tmp/vb_$name
for f in vc_kernels vc_solver vc_api_host vc_track_build; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --fmad=true -diag-suppress 177,550 "$@" \
       -c vehicle_control_b200/csrc/$f.cu -o /tmp/vb_$name/$f.o > /tmp/vb_$name/$f.log 2>&1 &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o tmp_data/libs/lib_$name.so /tmp/vb_$name/*.o && echo "built $name"
grep -h -A2 "Function properties for _ZN2vc9k_eval_st" /tmp/vb_$name/vc_solver.log | tail -2
