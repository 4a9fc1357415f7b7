"""Gate statistics of a QoT workload from the oracle (test infrastructure): gate evaluations per step, share that miss
on the candidate's own GSNR, neighbours and neighbour hops per evaluation.  DESIGN.md 6f quotes them."""
import ctypes, os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "tests"), os.path.join(REPO, "oracle")]
import bench, oracle_binding as OB

name = sys.argv[1] if len(sys.argv) > 1 else "c6_rmsa_us14_320_gn"
n_env, steps = 64, 400
spec, w = bench.make_workload(name)
pool = OB.OraclePool(spec, 0, n_env)
pool.rollout(w["preroll"], w["hid"], os.cpu_count())
L = OB.lib()
a = (ctypes.c_longlong * 4)(); b = (ctypes.c_longlong * 4)()
L.oracle_qot_stats(a)
pool.rollout(steps, w["hid"], os.cpu_count())
L.oracle_qot_stats(b)
d = [b[i] - a[i] for i in range(4)]
tot = n_env * steps
cnt = pool.counters()
print(name, {"gate_evals_per_step": d[0] / tot, "candidate_miss_share": d[1] / max(1, d[0]),
             "neighbours_per_full_eval": d[2] / max(1, d[0] - d[1]), "neighbour_hops_per_full_eval": d[3] / max(1, d[0] - d[1]),
             "blocking": 1 - cnt[:, 1].sum() / cnt[:, 0].sum(), "queue": float(cnt[:, 9].mean())})
