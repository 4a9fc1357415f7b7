#!/bin/bash
# ncu capture of the QoT gate kernel on the headline workload: tools/profile_qot.sh <tag>
tag=${1:-r02a}
mkdir -p gpurun_out
python tools/exp_qot.py > gpurun_out/exp_qot_$tag.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:qot_kernel -s 1510 -c 1 -o gpurun_out/qot_$tag -f python tools/exp_qot.py > gpurun_out/ncu_qot_$tag.log 2>&1
ncu -i gpurun_out/qot_$tag.ncu-rep --page raw --csv > gpurun_out/qot_${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/qot_$tag.ncu-rep --page source --csv > gpurun_out/qot_${tag}_source.csv 2>/dev/null
tail -3 gpurun_out/ncu_qot_$tag.log
