"""The bench configuration at its full size (BASELINE configs[1]: DeepRMSA-v0, NSFNET, 100 slots, 65 536 envs),
checked through properties that do not need the oracle to run 65 536 envs:

* the one-launch step with the fused next-action hint, the two-launch form and the fused rollout are three
  schedules of the same computation: identical counters, rewards and occupancy for every env;
* envs are keyed by global id: three shards with offsets reproduce the single engine;
* invariants of the reference's bookkeeping hold for every env (rmsa_env.py:222-341, 643-695);
* 48 envs sampled across the batch (first / last tiles included) equal the CPU oracle bit for bit."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB
from optical_rl_gym_qot_aware_b200 import spec as S

pytestmark = pytest.mark.gpu

N_ENV, STEPS = 65536, 240


def _spec():
    probs = [0.01801802, 0.04004004, 0.05305305, 0.01901902, 0.04504505, 0.02402402, 0.06706707,
             0.08908909, 0.13813814, 0.12212212, 0.07607608, 0.12012012, 0.01901902, 0.16916917]
    return S.make_spec("DeepRMSA-v0", G.topology("nsfnet"), seed=0x5EED, request_source=S.SRC_PHILOX,
                       auto_reset=True, obs_kind=S.OBS_F32, j=1, mean_service_holding_time=7.5,
                       mean_service_inter_arrival_time=1 / 12.0, num_spectrum_resources=100, episode_length=50,
                       node_request_probabilities=probs)


def test_full_size_schedules_shards_invariants_and_oracle_sample():
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine
    from optical_rl_gym_qot_aware_b200.sharding import shard_bounds

    spec = _spec()
    hid = S.H_SAP_FF
    dev = torch.device("cuda", 0)

    def stepped(fused):
        eng = Engine(spec, N_ENV)
        eng.reset(True)
        act = torch.zeros((N_ENV, 1), dtype=torch.int32, device=dev)
        rew = torch.zeros(N_ENV, dtype=torch.int32, device=dev)
        done = torch.zeros(N_ENV, dtype=torch.uint8, device=dev)
        obs = eng.new_obs()
        total = torch.zeros(N_ENV, dtype=torch.int64, device=dev)
        eng.heuristic(hid, act)
        if fused:
            eng.set_next_action_hint(hid, act)
        for _ in range(STEPS):
            if not fused:
                eng.heuristic(hid, act)
            eng.step(act, rew, done, obs, None)
            total += rew
        out = (eng.counters().cpu().numpy(), eng.export_state().cpu().numpy(), total.cpu().numpy(),
               obs.cpu().numpy().copy())
        eng.close()
        return out

    c1, s1, r1, o1 = stepped(True)
    c2, s2, r2, o2 = stepped(False)
    assert np.array_equal(c1, c2) and np.array_equal(s1, s2) and np.array_equal(r1, r2) and np.array_equal(o1, o2)

    roll = Engine(spec, N_ENV)
    roll.reset(True)
    rs = torch.zeros(N_ENV, dtype=torch.int64, device=dev)
    roll.rollout(hid, STEPS, rs)
    c3, s3 = roll.counters().cpu().numpy(), roll.export_state().cpu().numpy()
    roll.close()
    assert np.array_equal(c1, c3) and np.array_equal(s1, s3) and np.array_equal(r1, rs.cpu().numpy())

    for r in range(3):  # global env ids: shards reproduce the single engine
        lo, hi = shard_bounds(N_ENV, 3, r)
        part = Engine(spec, hi - lo, global_env_offset=lo)
        part.reset(True)
        part.rollout(hid, STEPS)
        assert np.array_equal(part.counters().cpu().numpy(), c1[lo:hi])
        assert np.array_equal(part.export_state().cpu().numpy(), s1[lo:hi])
        part.close()

    # bookkeeping invariants, every env
    proc, acc, ep_proc, ep_acc, br_req, br_prov = (c1[:, i] for i in range(6))
    assert (c1[:, 10] == 0).all()                                   # no device fault
    assert (proc == STEPS + 1).all()                                # the pending request is counted (rmsa_env.py:673)
    assert ((acc >= 0) & (acc <= STEPS)).all() and ((ep_acc <= ep_proc) & (ep_proc <= 50)).all()
    assert (br_prov <= br_req).all() and (br_req >= 200 * proc).all() and (br_req <= 1200 * proc).all()
    assert (r1 == 2 * acc - STEPS).all()                            # reward +1 / -1 (deeprmsa_env.py:123-124)
    assert (c1[:, 11] == STEPS // 49).all()                         # an episode is episode_length - 1 steps (F7)
    used = (s1 == 0).reshape(N_ENV, -1).sum(1)
    qlen = c1[:, 9]
    assert ((qlen >= 0) & (qlen <= acc)).all()
    assert (used >= 4 * qlen).all() and (used <= 97 * 9 * qlen).all()  # 4..97 slots on 1..9 links per live service
    assert 0.3 < 1.0 - acc.sum() / (STEPS * N_ENV) < 0.9               # the config blocks about two thirds

    # oracle on a sample spread over the batch
    sample = sorted(set(np.linspace(0, N_ENV - 1, 40).astype(int).tolist() + [1, 127, 128, N_ENV - 129, N_ENV - 2]))
    for i in sample:
        _, oc, ost, ors = OB.rollout_many(spec, i, 1, STEPS, hid, 1, want_state=True)
        assert np.array_equal(oc[0, :10], c1[i, :10]) and int(ors[0]) == int(r1[i]) and np.array_equal(ost[0], s1[i]), i


@pytest.mark.parametrize("workload,steps", [("c1_rmsa_nsfnet_100", 200), ("c3_rmcsa_jpn12_320", 200),
                                            ("c4_phyrmsa_us14_320", 150), ("c5_rmsa_spn_320", 200)])
def test_other_bench_workloads_at_full_size(workload, steps):
    """The remaining bench workloads exactly as bench.py builds them, 65 536 envs: fused-hint stepping == fused
    rollout, two shards == the single engine, no fault bits, and a sample of envs == the CPU oracle."""
    import os
    import sys

    import torch

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from optical_rl_gym_qot_aware_b200.engine import Engine
    from optical_rl_gym_qot_aware_b200.sharding import shard_bounds

    spec, w = bench.make_workload(workload)
    hid = w["hid"]
    dev = torch.device("cuda", 0)
    eng = Engine(spec, N_ENV)
    eng.reset(True)
    act = torch.zeros((N_ENV, eng.action_dim), dtype=torch.int32, device=dev)
    rew = torch.zeros(N_ENV, dtype=torch.int32, device=dev)
    done = torch.zeros(N_ENV, dtype=torch.uint8, device=dev)
    obs = eng.new_obs() if eng.obs_dim else None
    total = torch.zeros(N_ENV, dtype=torch.int64, device=dev)
    eng.heuristic(hid, act)
    eng.set_next_action_hint(hid, act)
    for _ in range(steps):
        eng.step(act, rew, done, obs, None)
        total += rew
    c1, s1, r1 = eng.counters().cpu().numpy(), eng.export_state().cpu().numpy(), total.cpu().numpy()
    eng.close()
    assert (c1[:, 10] == 0).all() and (c1[:, 0] == steps + (0 if workload.startswith("c3") else 1)).all()

    for r in range(2):
        lo, hi = shard_bounds(N_ENV, 2, r)
        part = Engine(spec, hi - lo, global_env_offset=lo)
        part.reset(True)
        rs = torch.zeros(hi - lo, dtype=torch.int64, device=dev)
        part.rollout(hid, steps, rs)
        assert np.array_equal(part.counters().cpu().numpy(), c1[lo:hi])
        assert np.array_equal(part.export_state().cpu().numpy(), s1[lo:hi])
        assert np.array_equal(rs.cpu().numpy(), r1[lo:hi])
        part.close()

    for i in sorted(set(np.linspace(0, N_ENV - 1, 12).astype(int).tolist() + [31, 32, N_ENV - 33])):
        _, oc, ost, ors = OB.rollout_many(spec, i, 1, steps, hid, 1, want_state=True)
        nc = 9 if workload.startswith("c4") else 10  # PhyRMSA queues one entry per channel, the oracle one per service
        assert np.array_equal(oc[0, :nc], c1[i, :nc]) and int(ors[0]) == int(r1[i]) and np.array_equal(ost[0], s1[i]), i
