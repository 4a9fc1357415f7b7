set -x
ORLG_DIRECT=1 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
ORLG_DIRECT=0 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in c3_rmcsa_jpn12_320 c4_phyrmsa_us14_320 c5_rmsa_spn_100 c5_rmsa_spn_320 c2_deeprmsa_nsfnet_100 c1_rmsa_nsfnet_100; do
 for d in 0 1; do
  echo "== $wl direct=$d"
  ORLG_DIRECT=$d python bench.py --workload $wl --steps 200 --warmup 10 --no-cpu-baseline --e2e-steps 5 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['kernel_ms'], d['extra'])"
 done
done
