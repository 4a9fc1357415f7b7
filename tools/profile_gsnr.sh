# ncu capture of the GN-model kernel on the C3 workload (8 192 envs keep the replayed launches short)
set -e
mkdir -p gpurun_out
python bench.py --workload c3_rmcsa_jpn12_320 --envs 8192 --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2 > gpurun_out/gsnr_plain.json 2> gpurun_out/gsnr_plain.err
ncu --set full --clock-control none --import-source on -k regex:gsnr_kernel -s 3 -c 1 -f -o gpurun_out/prof_gsnr \
  python bench.py --workload c3_rmcsa_jpn12_320 --envs 8192 --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2 > gpurun_out/ncu_gsnr.log 2>&1
ncu -i gpurun_out/prof_gsnr.ncu-rep --page raw --csv > gpurun_out/gsnr_kernel_raw.csv
ncu -i gpurun_out/prof_gsnr.ncu-rep --page source --csv --print-source sass > gpurun_out/gsnr_kernel_source.csv 2>/dev/null || true
