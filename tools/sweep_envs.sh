#!/bin/bash
# BASELINE configs[4]: env-count sweep 1 K - 4 M envs on the SPN topology (and the headline workload up to what fits),
# one B200.  tools/sweep_envs.sh <tag>   -> gpurun_out/<tag>_sweep.jsonl
tag=${1:-r02}
out=gpurun_out/${tag}_sweep.jsonl; : > $out
for w in c5_rmsa_spn_100 c6_rmsa_us14_320_gn; do
  for n in 1024 4096 16384 65536 262144 1048576 4194304; do
    if [ $w = c6_rmsa_us14_320_gn ] && [ $n -gt 1048576 ]; then continue; fi   # 29 KB per env: 4 M envs do not fit 180 GB with the e2e shards
    timeout 600 python bench.py --workload $w --envs $n --steps 10 --warmup 3 --no-cpu-baseline --no-workloads --e2e-steps 5 >> $out 2>> gpurun_out/${tag}_sweep.err || echo "{\"workload\": \"$w\", \"envs\": $n, \"failed\": true}" >> $out
  done
done
wc -l $out
