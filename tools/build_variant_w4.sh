#!/bin/bash
# A/B build that recompiles only the W=4 translation unit (C1/C2/C5-100) and links it with the other objects of
# the last full build: tools/build_variant_w4.sh <name> [extra nvcc flags] -> build/variants/liborlg_<name>.so
set -e
name=$1; shift
cd /root/repo/optical_rl_gym_qot_aware_b200/csrc
mkdir -p /root/repo/build/variants
obj=/root/repo/build/variants/orlg_w4_$name.o
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=false -std=c++17 -Xcompiler -fPIC -DORLG_FAST_AB "$@" -c -o $obj orlg_w4.cu
others=$(ls /root/repo/build/*.o | grep -v "/orlg_w4.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o /root/repo/build/variants/liborlg_$name.so $obj $others
ls -la /root/repo/build/variants/liborlg_$name.so
