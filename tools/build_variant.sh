#!/bin/bash
# build a kernel variant of liborlg.so into build/variants/liborlg_<name>.so : scripts_variant.sh <name> [extra nvcc flags]
set -e
name=$1; shift
cd /root/repo/optical_rl_gym_qot_aware_b200/csrc
out=/root/repo/build/variants/$name; mkdir -p $out
for u in orlg orlg_w2 orlg_w4 orlg_w10 orlg_phy_w2 orlg_phy_w4 orlg_phy_w10; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=false -std=c++17 -Xcompiler -fPIC "$@" -c -o $out/$u.o $u.cu &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o /root/repo/build/variants/liborlg_$name.so $out/*.o
rm -rf $out
ls -la /root/repo/build/variants/liborlg_$name.so
