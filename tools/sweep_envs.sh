# SURVEY 8d / BASELINE configs[4]: env-count sweep on the SPN topology (and the default workload), one GPU per call.
# usage: bash tools/sweep_envs.sh [tag]   -> gpurun_out/sweep_<tag>.jsonl
tag=${1:-r1}
mkdir -p gpurun_out
out=gpurun_out/sweep_$tag.jsonl; : > $out
for wl in c5_rmsa_spn_100 c5_rmsa_spn_320 c2_deeprmsa_nsfnet_100; do
  for n in 1024 4096 16384 65536 262144 1048576 4194304; do
    timeout 600 python bench.py --heap-capacity 1024 --workload $wl --envs $n --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 3 2>>gpurun_out/sweep_$tag.err | tail -1 >> $out
  done
done
python - <<PY
import json
print("| workload | envs | env-steps/s (L2 flushed) | warm L2 | fused rollout | step kernel us | roofline.frac | e2e |")
print("|---|---|---|---|---|---|---|---|")
for ln in open("$out"):
    try: d = json.loads(ln)
    except Exception: continue
    c = d["config"]; r = d["roofline"]; x = d["extra"]
    print(f"| {c['workload']} | {c['envs_per_gpu']} | {d['value']:.3g} | {x['warm_l2_env_steps_per_s']:.3g} | {x['fused_rollout_env_steps_per_s']:.3g} | {r['kernel_ms']*1e3:.1f} | {r['frac']:.3f} | {d['e2e']['value']:.3g} |")
PY
