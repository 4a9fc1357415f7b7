# A/B of the staged tile size (ORLG_BD) and the direct variant on the default workload (B200, one gpurun call)
mkdir -p gpurun_out
for wl in c2_deeprmsa_nsfnet_100 c1_rmsa_nsfnet_100; do
for v in "ORLG_BD=128" "ORLG_BD=64" "ORLG_BD=32" "ORLG_DIRECT=1"; do
  echo "== $wl $v"
  env $v python bench.py --workload $wl --steps 200 --warmup 20 --no-cpu-baseline --e2e-steps 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['roofline']['kernel_ms'], d['extra'])"
done; done 2>&1 | tee gpurun_out/ab_bd.log
