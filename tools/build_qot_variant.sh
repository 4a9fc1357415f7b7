#!/bin/bash
# variant of liborlg.so that differs in orlg.cu only (qot / gsnr kernels): tools/build_qot_variant.sh <name> [-D...]
set -e
name=$1; shift
cd /root/repo/optical_rl_gym_qot_aware_b200/csrc
mkdir -p /root/repo/build/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=false -std=c++17 -Xcompiler -fPIC "$@" -c -o /root/repo/build/variants/orlg_$name.o orlg.cu
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o /root/repo/build/variants/liborlg_$name.so /root/repo/build/variants/orlg_$name.o \
  /root/repo/build/orlg_w2.o /root/repo/build/orlg_w4.o /root/repo/build/orlg_w10.o /root/repo/build/orlg_phy_w2.o /root/repo/build/orlg_phy_w4.o /root/repo/build/orlg_phy_w10.o
