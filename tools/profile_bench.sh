#!/bin/bash
# Profile of a bench workload: plain run first, then the launch list of the same command, then one full capture of the
# dominant kernel (the last full-size launch of it the list holds).
#   tools/profile_bench.sh <tag> [kernel regex] [extra bench args...]
set -e
tag=${1:-r02}; kern=${2:-qot_kernel}; shift || true; shift || true
mkdir -p gpurun_out
ARGS="--steps 5 --warmup 3 --no-cpu-baseline --no-workloads --e2e-steps 2 $*"
python bench.py $ARGS > gpurun_out/${tag}_prof_plain.json 2> gpurun_out/${tag}_prof_plain.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py $ARGS > gpurun_out/${tag}_ncu_l.log 2>&1
skip=$(python - "$kern" gpurun_out/${tag}_launches.csv <<'PY'
import csv, re, sys
kern, path = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(path)) if len(r) > 14 and r[0].isdigit() and re.search(kern, r[4])]
grid = [int(r[8].strip("()").split(",")[0]) for r in rows]
big = max(grid)
last = max(i for i, g in enumerate(grid) if g == big)
print(max(0, last - 1))
PY
)
echo "full capture: launch $skip of $kern"
ncu --set full --clock-control none --import-source on -k regex:$kern -s $skip -c 1 -f -o gpurun_out/${tag}_prof \
  python bench.py $ARGS > gpurun_out/${tag}_ncu_f.log 2>&1
ncu -i gpurun_out/${tag}_prof.ncu-rep --page raw --csv > gpurun_out/${tag}_kernel_raw.csv 2>/dev/null
ncu -i gpurun_out/${tag}_prof.ncu-rep --page source --csv --print-source sass > gpurun_out/${tag}_kernel_source.csv 2>/dev/null
ls -la gpurun_out/${tag}_prof.ncu-rep
[ "$(stat -c %s gpurun_out/${tag}_prof.ncu-rep)" -gt 20000000 ] && rm gpurun_out/${tag}_prof.ncu-rep  # gpurun brings back 64 MiB at most
