"""Scratch experiment: throughput of the GN-model QoT gate workload, graph vs loop timing."""
import os, sys, time
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "tests")]
import golden_util as G
from optical_rl_gym_qot_aware_b200 import spec as S
from optical_rl_gym_qot_aware_b200.engine import Engine

n = int(os.environ.get("N", 65536))
load = float(os.environ.get("LOAD", 100))
adaptive = os.environ.get("ADAPT", "1") == "1"
env_id = os.environ.get("ENV", "RMSA-v0")
if env_id == "RMSA-v0":
    spec = S.make_spec("RMSA-v0", G.topology("us14"), seed=0x5EED, auto_reset=True, qot="gn", qot_adaptive=adaptive,
                       load=load, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=1000)
else:
    spec = S.make_spec("RMCSA-v0", G.topology("jpn12_k5"), seed=0x5EED, auto_reset=True, qot="gn", qot_adaptive=adaptive,
                       load=250, mean_service_holding_time=25, num_spectrum_resources=320, num_spatial_resources=7,
                       worst_xt=-84.7, episode_length=1000)
eng = Engine(spec, n)
print("hbm bytes/env", eng.hbm_bytes / n)
eng.reset(True)
t0 = time.time(); eng.rollout(S.H_SAP_GN_FF, 1500); torch.cuda.synchronize(); print("preroll s", time.time() - t0)
cnt = eng.counters()
print("acc", float(cnt[:, 1].sum()) / float(cnt[:, 0].sum()), "queue", float(cnt[:, 9].float().mean()), "fault", int(cnt[:, 10].max()))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=eng.device)
for g in (True, False, True):
    ms, first = eng.timed_steps(20, heuristic_id=S.H_SAP_GN_FF, flush=flush, use_graph=g)
    print("graph" if g else "loop", "sum ms", ms.sum(), "median", np.median(ms), "gate kernel median", np.median(first),
          "steps/s", n * 20 / (ms.sum() * 1e-3))
ms, first = eng.timed_steps(20, heuristic_id=S.H_SAP_GN_FF, flush=None, use_graph=True)
print("warm", np.median(ms), np.median(first))
print("check", int(eng.qot_check().abs().sum()), "fault", int(eng.counters()[:, 10].max()))
# the agent loop: heuristic launch, then the step with its answer (verdict launch takes the proposal)
act = torch.zeros((n, eng.action_dim), dtype=torch.int32, device=eng.device)
rew = torch.zeros(n, dtype=torch.int32, device=eng.device)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
th, ts = [], []
for k in range(20):
    ev[0].record(); eng.heuristic(S.H_SAP_GN_FF, act); ev[1].record(); eng.step(act, rew, None, None, None); ev[2].record()
    torch.cuda.synchronize()
    th.append(ev[0].elapsed_time(ev[1])); ts.append(ev[1].elapsed_time(ev[2]))
print("agent loop: heuristic ms", np.median(th), "step (verdict + step kernel) ms", np.median(ts))
print("check", int(eng.qot_check().abs().sum()), "fault", int(eng.counters()[:, 10].max()))
