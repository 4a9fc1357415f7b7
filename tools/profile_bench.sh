#!/bin/bash
# Profile of the default bench workload: plain run first, then the launch list of the same command, then one full
# capture of the dominant kernel.  tools/profile_bench.sh <tag> [kernel regex] [extra bench args...]
set -e
tag=${1:-r02}; kern=${2:-qot_kernel}; shift || true; shift || true
mkdir -p gpurun_out
ARGS="--steps 5 --warmup 3 --no-cpu-baseline --no-workloads --e2e-steps 2 $*"
python bench.py $ARGS > gpurun_out/${tag}_prof_plain.json 2> gpurun_out/${tag}_prof_plain.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py $ARGS > gpurun_out/${tag}_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:$kern -s 620 -c 1 -f -o gpurun_out/${tag}_prof \
  python bench.py $ARGS > gpurun_out/${tag}_ncu_f.log 2>&1
ncu -i gpurun_out/${tag}_prof.ncu-rep --page raw --csv > gpurun_out/${tag}_kernel_raw.csv 2>/dev/null
ls -la gpurun_out/${tag}_prof.ncu-rep
