# One gpurun call that refreshes everything under profiles/ for a capture tag (default r01j):
# GPU tests, the default bench line and its reference arm, the launch list + full-set capture of the step
# kernel (tools/profile_c2.sh), then every other workload.
tag=${1:-r01j}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${tag}_gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err; echo "ref rc=$?"
sed -i "s/r1h/$tag/g" tools/profile_c2.sh
timeout 900 bash tools/profile_c2.sh > gpurun_out/${tag}_profile.log 2>&1; echo "prof rc=$?"
ncu -i gpurun_out/prof_$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_env_kernel_raw.csv 2>/dev/null
for w in c1_rmsa_nsfnet_100 c3_rmcsa_jpn12_320 c4_phyrmsa_us14_320 c5_rmsa_spn_100 c5_rmsa_spn_320; do
  timeout 600 python bench.py --workload $w --steps 100 --warmup 10 --cpu-seconds 8 2>>gpurun_out/${tag}_wl.err | tail -1 > gpurun_out/${tag}_wl_$w.json
done
timeout 600 python bench.py --workload c3_rmcsa_jpn12_320 --no-gsnr --steps 100 --warmup 10 --cpu-seconds 8 2>>gpurun_out/${tag}_wl.err | tail -1 > gpurun_out/${tag}_wl_c3_rmcsa_jpn12_320_no_gsnr.json
tail -3 gpurun_out/${tag}_pytest.log; cat gpurun_out/${tag}_bench.json
