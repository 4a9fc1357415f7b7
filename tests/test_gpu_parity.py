"""GPU parity: the CUDA engine, called through the C-ABI, must reproduce
(a) every golden trace exported from the unmodified reference, bit for bit, and
(b) the CPU oracle on seeded Philox streams at larger env counts."""
import numpy as np
import pytest

import golden_util as G
from optical_rl_gym_qot_aware_b200 import spec as S

pytestmark = pytest.mark.gpu

HEUR_OF_CASE = {
    "rmsa_nsfnet_sapff": 1, "rmsa_nsfnet_spff": 0, "rmsa_nsfnet_llpff": 2,
    "rmsa_spn_sapff": 1, "rmsa_spn320_sapff": 1,
    "deeprmsa_nsfnet_kspff": 1, "deeprmsa_nsfnet_spff": 0,
    "rmcsa_jpn12_heur": 3, "rmcsa_nsfnet_heur": 3,
}


def _engine(spec, n, **kw):
    from optical_rl_gym_qot_aware_b200.engine import Engine

    return Engine(spec, n, **kw)


@pytest.mark.parametrize("layout", ["feature_major", "env_major"])
@pytest.mark.parametrize("case", G.CASES)
def test_engine_replays_golden(case, layout):
    import torch

    if layout == "env_major" and not case.startswith("deeprmsa"):
        pytest.skip("layout only matters for envs with an observation")
    g, spec = G.load_case(case)
    n_env = 3  # same trace in every env: exercises partially filled tiles too
    eng = _engine(spec, n_env, obs_layout=layout)
    reqs = g["requests"]
    eng.load_trace(np.repeat(reqs[:, None], n_env, axis=1))
    eng.reset(full=True)
    kind = spec.env_kind
    dev = eng.device
    n = len(g["actions"])
    reward = torch.zeros(n_env, dtype=torch.int32, device=dev)
    done = torch.zeros(n_env, dtype=torch.uint8, device=dev)
    alloc = torch.zeros((n_env, 4), dtype=torch.int32, device=dev)
    obs = eng.new_obs() if kind == 1 else None
    hact = torch.zeros((n_env, eng.action_dim), dtype=torch.int32, device=dev)
    if kind == 1:
        eng.observation(obs)
        assert np.array_equal(obs.cpu().numpy()[1], g["obs0"])
    for t in range(n):
        if case in HEUR_OF_CASE:
            eng.heuristic(HEUR_OF_CASE[case], hact)
            assert np.array_equal(hact.cpu().numpy()[2], g["actions"][t]), (case, t)
        act = torch.from_numpy(np.repeat(g["actions"][t][None], n_env, axis=0).astype(np.int32)).to(dev)
        eng.step(act, reward, done, obs, alloc)
        a = alloc.cpu().numpy()
        assert (a == a[0]).all()
        assert a[0, 0] == g["accepted"][t], (case, t)
        assert reward.cpu().numpy()[0] == g["reward"][t], (case, t)
        d = bool(done.cpu().numpy()[0])
        assert d == bool(g["done"][t]), (case, t)
        if a[0, 0]:
            path, slot, ns, core = g["alloc"][t]
            assert (a[0, 1] & 0xFFFF, a[0, 1] >> 16, a[0, 2], a[0, 3]) == (path, core, slot, ns), (case, t)
        if d:
            eng.reset(full=False)
        assert np.array_equal(eng.counters().cpu().numpy()[1, :8], g["counters"][t]), (case, t)
        if kind == 1:
            assert np.array_equal(obs.cpu().numpy()[2], g["obs"][t]), (case, t)
        rq = eng.requests()[0]
        assert rq == reqs[t + 1], (case, t)
    st = eng.export_state().cpu().numpy()
    assert np.array_equal(st[0], g["final_avail"]) and np.array_equal(st[2], g["final_avail"])
    assert int(eng.counters()[:, 10].max()) == 0  # no device fault
    eng.close()


PHILOX_CASES = [
    ("RMSA-v0", "nsfnet", dict(load=50, num_spectrum_resources=100, episode_length=70), 1, 1000, 300),
    ("RMSA-v0", "nsfnet", dict(load=80, num_spectrum_resources=100, episode_length=50), 2, 700, 200),
    ("RMSA-v0", "spn", dict(load=300, mean_service_holding_time=25, num_spectrum_resources=320), 1, 300, 300),
    ("RMSA-v0", "germany50", dict(load=300, mean_service_holding_time=25, num_spectrum_resources=64), 0, 200, 200),
    ("DeepRMSA-v0", "nsfnet", dict(mean_service_holding_time=7.5, mean_service_inter_arrival_time=1 / 12.0,
                                   episode_length=50, j=1), 1, 1000, 300),
    ("RMCSA-v0", "jpn12_k5", dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                  num_spatial_resources=7, episode_length=100), 3, 200, 300),
]


@pytest.mark.parametrize("env_id,topo,kw,hid,n_env,n_steps", PHILOX_CASES)
def test_rollout_matches_oracle_on_philox(env_id, topo, kw, hid, n_env, n_steps):
    """Fused rollout (heuristic + step, state on chip) == CPU oracle, env by env."""
    import torch

    import oracle_binding as OB

    spec = S.make_spec(env_id, G.topology(topo), seed=1234, request_source=S.SRC_PHILOX, auto_reset=True, **kw)
    _, o_cnt, o_state, o_rsum = OB.rollout_many(spec, 5, n_env, n_steps, hid, 8, want_state=True)
    eng = _engine(spec, n_env, global_env_offset=5)
    eng.reset(full=True)
    rsum = torch.zeros(n_env, dtype=torch.int64, device=eng.device)
    eng.rollout(hid, n_steps // 2, rsum)
    eng.rollout(hid, n_steps - n_steps // 2, rsum)
    cnt = eng.counters().cpu().numpy()
    assert np.array_equal(cnt[:, :9], o_cnt[:, :9])
    assert np.array_equal(cnt[:, 9], o_cnt[:, 9])  # queue length
    assert np.array_equal(cnt[:, 11], o_cnt[:, 11])  # episodes
    assert int(cnt[:, 10].max()) == 0
    assert np.array_equal(rsum.cpu().numpy(), o_rsum)
    assert np.array_equal(eng.export_state().cpu().numpy(), o_state)
    eng.close()


def test_step_mode_equals_rollout_mode():
    """gym-mode (heuristic kernel -> step kernel) and the fused rollout give identical state."""
    import torch

    spec = S.make_spec("DeepRMSA-v0", G.topology("nsfnet"), seed=7, mean_service_holding_time=7.5,
                       mean_service_inter_arrival_time=1 / 12.0, episode_length=50, auto_reset=True)
    n_env, n_steps = 500, 120
    a = _engine(spec, n_env)
    b = _engine(spec, n_env)
    a.reset(True)
    b.reset(True)
    act = torch.zeros((n_env, 1), dtype=torch.int32, device=a.device)
    obs = a.new_obs()
    for _ in range(n_steps):
        a.heuristic(1, act)
        a.step(act, None, None, obs, None)
    b.rollout(1, n_steps)
    assert torch.equal(a.counters(), b.counters())
    assert torch.equal(a.export_state(), b.export_state())
    obs_b = b.new_obs()
    b.observation(obs_b)
    assert torch.equal(obs, obs_b)


@pytest.mark.parametrize("env_id,topo,kw,hid", [
    ("DeepRMSA-v0", "nsfnet", dict(mean_service_holding_time=7.5, mean_service_inter_arrival_time=1 / 12.0,
                                   episode_length=50, j=2), 1),
    ("RMSA-v0", "spn", dict(load=100, mean_service_holding_time=25, num_spectrum_resources=100, episode_length=80), 2),
    ("RMCSA-v0", "nsfnet", dict(load=250, mean_service_holding_time=25, num_spectrum_resources=64,
                                num_spatial_resources=7, episode_length=90), 3),
])
def test_fused_next_action_hint_equals_heuristic_launch(env_id, topo, kw, hid):
    """orlg_set_next_action_hint: the action written by the step kernel for the new pending request
    is what a separate orlg_heuristic launch writes, also when it aliases the action buffer."""
    import torch

    spec = S.make_spec(env_id, G.topology(topo), seed=21, auto_reset=True, **kw)
    n_env, n_steps = 400, 150
    a, b = _engine(spec, n_env), _engine(spec, n_env)
    a.reset(True)
    b.reset(True)
    act_a = torch.zeros((n_env, a.action_dim), dtype=torch.int32, device=a.device)
    act_b = torch.zeros_like(act_a)
    a.heuristic(hid, act_a)
    a.set_next_action_hint(hid, act_a)  # aliased: read as action, overwritten with the next one
    for t in range(n_steps):
        b.heuristic(hid, act_b)
        assert torch.equal(act_a, act_b), t
        a.step(act_a)
        b.step(act_b)
    assert torch.equal(a.counters(), b.counters()) and torch.equal(a.export_state(), b.export_state())
    a.set_next_action_hint(hid, None)
    before = act_a.clone()
    a.step(act_a)
    assert torch.equal(act_a, before)


INFO_CASES = [c for c in G.CASES if c.startswith("info_")]


@pytest.mark.gpu
@pytest.mark.parametrize("case", INFO_CASES)
def test_engine_info_dict_matches_reference(case):
    """the step's info dict (rmsa_env.py:293-332) written by the step kernel — blocking rates,
    network compactness (+ difference), time-weighted link compactness / utilization means, fairness, per-bit-rate
    blocking — bit for bit against the golden trace of the unmodified reference."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    g, spec = G.load_case(case)
    spec.info_stats = True
    n_env = 3
    eng = Engine(spec, n_env)
    eng.load_trace(np.repeat(g["requests"][:, None], n_env, axis=1))
    eng.reset(True)
    info = eng.set_info_output()
    dev = eng.device
    reward = torch.zeros(n_env, dtype=torch.int32, device=dev)
    done = torch.zeros(n_env, dtype=torch.uint8, device=dev)
    act = torch.zeros((n_env, eng.action_dim), dtype=torch.int32, device=dev)
    for t in range(len(g["actions"])):
        act[:] = torch.as_tensor(g["actions"][t].astype(np.int32), device=dev)
        eng.step(act, reward, done)
        want = np.concatenate([g["info_main"][t], [g["info_fairness"][t]], g["info_bit_rate_blocking"][t]])
        got = info.cpu().numpy()
        assert np.array_equal(got[0], want) and np.array_equal(got[2], want), (case, t, np.flatnonzero(got[0] != want), got[0], want)
        if bool(done.cpu().numpy()[0]):
            eng.reset(False)
    assert int(eng.counters()[:, 10].max()) == 0
    eng.close()
