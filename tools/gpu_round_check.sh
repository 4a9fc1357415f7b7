mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r1i_gpus.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r1i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1i_pytest.log
timeout 600 python bench.py > gpurun_out/r1i_bench.json 2> gpurun_out/r1i_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1i_bench_reference.json 2> gpurun_out/r1i_bench_reference.err; echo "ref rc=$?"
sed -i 's/r1h/r1i/g' tools/profile_c2.sh
timeout 900 bash tools/profile_c2.sh > gpurun_out/r1i_profile.log 2>&1; echo "prof rc=$?"
ncu -i gpurun_out/prof_r1i.ncu-rep --page raw --csv > gpurun_out/r1i_env_kernel_raw.csv 2>/dev/null
tail -3 gpurun_out/r1i_pytest.log; cat gpurun_out/r1i_bench.json
