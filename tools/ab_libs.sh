# A/B of library builds on one box: tools/ab_libs.sh "<lib paths>" "<workloads>" [steps] [envs]
libs=$1; wls=${2:-c2_deeprmsa_nsfnet_100}; steps=${3:-300}; envs=${4:-65536}
for wl in $wls; do for lib in $libs; do
  echo "== $wl envs=$envs $lib"
  ORLG_LIB=$lib python bench.py --workload $wl --envs $envs --steps $steps --warmup 20 --no-cpu-baseline --e2e-steps 2 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.4g kernel_us=%.2f warm=%.4g rollout=%.4g' % (d['value'], d['roofline']['kernel_ms']*1e3, d['extra']['warm_l2_env_steps_per_s'], d['extra']['fused_rollout_env_steps_per_s']))"
done; done
