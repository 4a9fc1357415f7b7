"""Soak of the QoT gate's running sums: long rollouts, then orlg_qot_check (every row of every running service against a
from-scratch evaluation, bit for bit) and the fault bits.   python tools/soak_qot.py [rounds] [steps per round]"""
import os, sys, time
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "tests")]
import golden_util as G
from optical_rl_gym_qot_aware_b200 import spec as S
from optical_rl_gym_qot_aware_b200.engine import Engine

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 4
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
for env_id, topo, kw in (("RMSA-v0", "us14", dict(load=100, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=1000)),
                         ("RMCSA-v0", "jpn12_k5", dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                                       num_spatial_resources=7, worst_xt=-84.7, episode_length=1000))):
    spec = S.make_spec(env_id, G.topology(topo), seed=0x50AC, auto_reset=True, qot="gn", qot_adaptive=True, **kw)
    eng = Engine(spec, 65536)
    eng.reset(True)
    act = torch.zeros((65536, eng.action_dim), dtype=torch.int32, device=eng.device)
    rew = torch.zeros(65536, dtype=torch.int32, device=eng.device)
    for r in range(rounds):
        t0 = time.time()
        eng.rollout(S.H_SAP_GN_FF, steps)
        for _ in range(50):  # the agent loop too (proposal path)
            eng.heuristic(S.H_SAP_GN_FF, act)
            eng.step(act, rew, None, None, None)
        bad = int(eng.qot_check().abs().sum())
        cnt = eng.counters()
        print(env_id, "round", r, "steps", int(cnt[0, 0]), "bad rows", bad, "fault", int(cnt[:, 10].max()),
              "accepted", float(cnt[:, 1].sum()) / float(cnt[:, 0].sum()), f"{time.time() - t0:.1f}s", flush=True)
        assert bad == 0 and int(cnt[:, 10].max()) == 0
print("soak ok")
