# round-1 profile of the default workload: plain run first, then the launch list, then one full capture
set -e
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_r1h.csv \
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:env_kernel -s 8 -c 3 -f -o gpurun_out/prof_r1h \
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2 > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out/prof_r1h.ncu-rep
