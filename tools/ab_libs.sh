# A/B of library builds on one box: tools/ab_libs.sh "<lib paths>" "<workloads>" [steps]
libs=$1; wls=${2:-c2_deeprmsa_nsfnet_100}; steps=${3:-500}
for wl in $wls; do for lib in $libs; do
  echo "== $wl $lib"
  ORLG_LIB=$lib python bench.py --workload $wl --steps $steps --warmup 20 --no-cpu-baseline --e2e-steps 5 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['kernel_ms'], d['extra'])"
done; done
