# ncu capture of the PhyRMSA step kernel on the C4 workload (16 384 envs keep the pre-roll short under ncu)
set -e
mkdir -p gpurun_out
python bench.py --workload c4_phyrmsa_us14_320 --envs 16384 --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2 > gpurun_out/c4_plain.json 2> gpurun_out/c4_plain.err
ncu --set full --clock-control none --import-source on -k regex:env_kernel -s 12 -c 2 -f -o gpurun_out/prof_c4 \
  python bench.py --workload c4_phyrmsa_us14_320 --envs 16384 --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2 > gpurun_out/ncu_c4.log 2>&1
ncu -i gpurun_out/prof_c4.ncu-rep --page raw --csv > gpurun_out/c4_kernel_raw.csv
