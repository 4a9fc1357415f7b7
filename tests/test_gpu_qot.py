"""GN-model QoT gate on the GPU (qot_kernel + step kernel through the C-ABI).

1. Replay of the traces in which the unmodified reference envs step behind a gate whose every GSNR is the
   reference's own calculate_osnr (oracle/refharness/gen_qot_golden.py): decisions, slots, rewards, final
   spectrum identical; GSNR within 1e-4 dB (north star).
2. Philox-driven envs against the oracle's literal gate, env by env.
3. The engine's running cross-channel sums equal a from-scratch evaluation bit for bit (orlg_qot_check)."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB
from optical_rl_gym_qot_aware_b200 import spec as S

pytestmark = pytest.mark.gpu
TOL_DB = 1e-4


def close(a, b, tol=TOL_DB):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    same = (np.isnan(a) & np.isnan(b)) | (np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b)))
    with np.errstate(invalid="ignore"):
        return bool(np.all(same | (np.abs(a - b) <= tol)))


@pytest.mark.parametrize("case", G.QOT_CASES)
def test_engine_gate_replays_reference_trace(case):
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    g, spec = G.load_case(case)
    n_env = 3
    eng = Engine(spec, n_env)
    eng.load_trace(np.repeat(g["requests"][:, None], n_env, axis=1))
    eng.reset(True)
    dev = eng.device
    by_heur = bool(int(g["by_heuristic"]))
    rew = torch.zeros(n_env, dtype=torch.int32, device=dev)
    done = torch.zeros(n_env, dtype=torch.uint8, device=dev)
    alloc = torch.zeros((n_env, 4), dtype=torch.int32, device=dev)
    hact = torch.zeros((n_env, eng.action_dim), dtype=torch.int32, device=dev)
    for t in range(len(g["actions"])):
        if by_heur:
            eng.heuristic(S.H_SAP_GN_FF, hact)
            assert np.array_equal(hact.cpu().numpy()[1], g["actions"][t]), (case, t)
            assert close(eng.qot_last().cpu().numpy()[1], g["heur_gsnr"][t]), (case, t)
        act = torch.from_numpy(np.repeat(g["actions"][t][None], n_env, axis=0).astype(np.int32)).to(dev)
        eng.step(act, rew, done, None, alloc)
        al = alloc.cpu().numpy()
        assert (al[:, 0] == g["accepted"][t]).all() and (al[:, 3] == g["slots"][t]).all(), (case, t, al)
        assert (rew.cpu().numpy() == g["rewards"][t]).all() and (done.cpu().numpy() == g["dones"][t]).all()
        assert close(eng.qot_last().cpu().numpy()[2], g["gsnr"][t]), (case, t, eng.qot_last().cpu().numpy()[2], g["gsnr"][t])
        if bool(done[0]):
            eng.reset(False)
        if t % 97 == 0:
            assert int(eng.qot_check().abs().sum()) == 0, (case, t)
    st = eng.export_state().cpu().numpy()
    assert np.array_equal(st[0].reshape(g["final_state"].shape), g["final_state"])
    assert int(eng.qot_check().abs().sum()) == 0
    assert int(eng.counters()[:, 10].max()) == 0


CONFIGS = [
    ("RMSA-v0", "us14", dict(load=60, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=100), True),
    ("RMSA-v0", "nsfnet", dict(load=40, mean_service_holding_time=25, num_spectrum_resources=100, episode_length=70), False),
    ("RMCSA-v0", "jpn12_k5", dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                  num_spatial_resources=7, worst_xt=-84.7, episode_length=100), True),
]


@pytest.mark.parametrize("env_id,topo,kw,adaptive", CONFIGS)
def test_engine_gate_matches_oracle_on_philox_streams(env_id, topo, kw, adaptive):
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec(env_id, G.topology(topo), seed=77, auto_reset=True, qot="gn", qot_adaptive=adaptive, **kw)
    n_env, n_steps = 40, 260
    eng = Engine(spec, n_env)
    eng.reset(True)
    dev = eng.device
    rew = torch.zeros(n_env, dtype=torch.int32, device=dev)
    alloc = torch.zeros((n_env, 4), dtype=torch.int32, device=dev)
    act = torch.zeros((n_env, eng.action_dim), dtype=torch.int32, device=dev)
    ora = [OB.OracleEnv(spec, i) for i in range(n_env)]
    for o in ora:
        o.reset(True)
    for t in range(n_steps):
        eng.step_auto(S.H_SAP_GN_FF, act, rew, None, None, alloc)
        a, al, r, gq = act.cpu().numpy(), alloc.cpu().numpy(), rew.cpu().numpy(), eng.qot_last().cpu().numpy()
        for i, o in enumerate(ora):
            oa = o.heuristic(S.H_SAP_GN_FF)
            assert np.array_equal(oa, a[i]), (t, i, oa, a[i])
            hq = o.qot_last()
            _, orr, _, oal = o.step(oa)
            assert orr == r[i] and np.array_equal(oal, al[i]), (t, i, oal, al[i])
            # the fused form evaluates the gate once: a blocked request reports the heuristic's last candidate
            want = o.qot_last() if oal[0] else hq
            assert close(want, gq[i]), (t, i, want, gq[i])
    cnt = eng.counters().cpu().numpy()
    for i, o in enumerate(ora):
        assert np.array_equal(o.counters()[:10], cnt[i, :10])
    assert np.array_equal(eng.export_state().cpu().numpy(), np.stack([o.state() for o in ora]))
    assert int(eng.qot_check().abs().sum()) == 0 and int(cnt[:, 10].max()) == 0
    acc = cnt[:, 1].sum() / cnt[:, 0].sum()
    assert 0.2 < acc < 0.999


def test_gate_with_another_policy_and_rollout():
    """A non-GN heuristic proposes, the gate decides (orlg_step_auto's 3-launch form); orlg_rollout on a gated engine."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec("RMSA-v0", G.topology("us14"), seed=5, auto_reset=True, qot="gn", load=80,
                       mean_service_holding_time=25, num_spectrum_resources=320, episode_length=100)
    n_env, n_steps = 64, 200
    eng = Engine(spec, n_env)
    eng.reset(True)
    rsum = torch.zeros(n_env, dtype=torch.int64, device=eng.device)
    eng.rollout(S.H_SAP_FF, n_steps, rsum)
    ora = [OB.OracleEnv(spec, i) for i in range(n_env)]
    tot = np.zeros(n_env, np.int64)
    for i, o in enumerate(ora):
        o.reset(True)
        for _ in range(n_steps):
            _, r, _, _ = o.step(o.heuristic(S.H_SAP_FF))
            tot[i] += r
    assert np.array_equal(rsum.cpu().numpy(), tot)
    assert np.array_equal(eng.export_state().cpu().numpy(), np.stack([o.state() for o in ora]))
    assert int(eng.qot_check().abs().sum()) == 0
    # the gate does block here: fewer acceptances than the ungated env on the same stream
    spec0 = S.make_spec("RMSA-v0", G.topology("us14"), seed=5, auto_reset=True, load=80,
                        mean_service_holding_time=25, num_spectrum_resources=320, episode_length=100)
    e0 = Engine(spec0, n_env)
    e0.reset(True)
    r0 = torch.zeros(n_env, dtype=torch.int64, device=eng.device)
    e0.rollout(S.H_SAP_FF, n_steps, r0)
    assert int(r0.sum()) > int(rsum.sum())


@pytest.mark.parametrize("env_id,topo,kw,adaptive", CONFIGS)
def test_heuristic_then_step_takes_the_proposal(env_id, topo, kw, adaptive):
    """The agent loop (orlg_heuristic 13, then orlg_step with its answer) gives what orlg_step_auto gives: the verdict
    launch takes the heuristic launch's verdict.  A different action, a reset or another step in between do not."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec(env_id, G.topology(topo), seed=91, auto_reset=True, qot="gn", qot_adaptive=adaptive, **kw)
    n_env, n_steps = 48, 220
    a_eng, b_eng = Engine(spec, n_env), Engine(spec, n_env)
    a_eng.reset(True)
    b_eng.reset(True)
    dev = a_eng.device
    mk = lambda *s, dt=torch.int32: torch.zeros(s, dtype=dt, device=dev)  # noqa: E731
    act_a, act_b, rew_a, rew_b = mk(n_env, a_eng.action_dim), mk(n_env, a_eng.action_dim), mk(n_env), mk(n_env)
    al_a, al_b = mk(n_env, 4), mk(n_env, 4)
    rng = np.random.default_rng(3)
    for t in range(n_steps):
        a_eng.step_auto(S.H_SAP_GN_FF, act_a, rew_a, None, None, al_a)
        b_eng.heuristic(S.H_SAP_GN_FF, act_b)
        assert torch.equal(act_a, act_b), t
        if t % 7 == 3:  # some envs get another action than the proposal: evaluated afresh (the twin steps it too)
            other = act_b.clone()
            pick = torch.from_numpy(rng.random(n_env) < 0.3).to(dev)
            other[pick, -1] = (other[pick, -1] + 1) % spec.num_slots
            b_eng.step(other, rew_b, None, None, al_b)
            assert int(b_eng.qot_check().abs().sum()) == 0
            # the twin cannot follow a diverged history: both start again (a reset also ends a proposal)
            a_eng.reset(True)
            b_eng.reset(True)
            continue
        if t % 11 == 5:  # a heuristic launch that is not followed by its step: the next proposal must be recomputed
            b_eng.heuristic(S.H_SAP_GN_FF, act_b)
        b_eng.step(act_b, rew_b, None, None, al_b)
        assert torch.equal(rew_a, rew_b) and torch.equal(al_a, al_b), t
        ok = al_a[:, 0].cpu().numpy() == 1  # a blocked request: the fused form reports the heuristic's last candidate, the step nothing
        assert close(a_eng.qot_last().cpu().numpy()[ok], b_eng.qot_last().cpu().numpy()[ok]), t
    assert torch.equal(a_eng.export_state(), b_eng.export_state())
    assert torch.equal(a_eng.counters(), b_eng.counters())
    assert int(a_eng.qot_check().abs().sum()) == 0 and int(b_eng.qot_check().abs().sum()) == 0


@pytest.mark.parametrize("env_id,topo,kw,adaptive", [CONFIGS[0], CONFIGS[2]])
def test_running_sums_survive_a_long_run(env_id, topo, kw, adaptive):
    """Thousands of provisions and releases per env (events applied with reductions that return nothing, accepted
    lightpaths handed from launch to launch): every running sum still equals a from-scratch evaluation, bit for bit."""
    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec(env_id, G.topology(topo), seed=1234, auto_reset=True, qot="gn", qot_adaptive=adaptive, **kw)
    eng = Engine(spec, 4096)
    eng.reset(True)
    for _ in range(3):
        eng.rollout(S.H_SAP_GN_FF, 1500)
        assert int(eng.qot_check().abs().sum()) == 0
    cnt = eng.counters()
    assert int(cnt[:, 10].max()) == 0 and int(cnt[0, 0]) >= 4500  # the pending request is counted on arrival
