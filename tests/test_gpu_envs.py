"""GPU tests of the gym-facing surface (make / BatchedEnv / OrlgVecEnv / GymEnv): the wrappers
must hand back exactly what the engine computes, in the reference's shapes."""
import numpy as np
import pytest

import golden_util as G
from optical_rl_gym_qot_aware_b200 import spec as S

pytestmark = pytest.mark.gpu

HEUR = {"rmsa_nsfnet_sapff": "shortest_available_path_first_fit",
        "deeprmsa_nsfnet_kspff": "shortest_available_path_first_fit",
        "rmcsa_jpn12_heur": "shortest_available_path_best_modulation_first_core_first_fit"}


def _kwargs(g):
    kw = {}
    for k in g["kwargs_keys"]:
        v = g["kw_" + str(k)]
        kw[str(k)] = v.item() if v.ndim == 0 else v
    return kw


@pytest.mark.parametrize("case", sorted(HEUR))
def test_scalar_gym_env_replays_reference_trace(case):
    """make(id, topology=...) -> scalar env; heuristic + step reproduce the golden trace 4-tuples."""
    import optical_rl_gym_qot_aware_b200 as pkg

    g, spec = G.load_case(case)
    env = pkg.make(G.ENV_NAME[int(g["env_kind"])], topology=spec.tables, request_source="replay", **_kwargs(g))
    env.load_trace(g["requests"])
    n = min(400, len(g["actions"]))
    if spec.env_kind == 1:
        assert np.array_equal(env.observation(), g["obs0"])
    for t in range(n):
        cs = env.current_service
        assert (cs.arrival_time, cs.holding_time, cs.bit_rate, cs.source_id, cs.destination_id) == tuple(
            g["requests"][t].tolist())
        a = env.heuristic(HEUR[case])
        assert np.array_equal(np.atleast_1d(a), g["actions"][t])
        obs, r, d, info = env.step(a)
        assert isinstance(r, int) and isinstance(d, bool)
        assert (r, d, info["accepted"]) == (int(g["reward"][t]), bool(g["done"][t]), int(g["accepted"][t]))
        if spec.env_kind == 1:
            assert np.array_equal(obs, g["obs"][t])
        if d:
            env.reset()
        assert env.episode_services_processed == int(g["counters"][t][2])
    av = env.topology.graph["available_slots"]
    assert av.shape == ((spec.num_cores,) if spec.env_kind == 2 else ()) + (spec.tables.n_edges, spec.num_slots)
    env.close()


def test_vecenv_matches_device_path_and_oracle():
    """OrlgVecEnv (host numpy, orlg_step_host) == BatchedEnv (device tensors) == CPU oracle."""
    import torch

    import optical_rl_gym_qot_aware_b200 as pkg
    import oracle_binding as OB

    kw = dict(seed=99, mean_service_holding_time=7.5, mean_service_inter_arrival_time=1 / 12.0, episode_length=30,
              j=2)
    n_env, n_steps = 300, 90
    topo = G.topology("nsfnet")
    vec = pkg.make("DeepRMSA-v0", topology=topo, num_envs=n_env, vec_env=True, **kw)
    dev = pkg.make("DeepRMSA-v0", topology=topo, num_envs=n_env, auto_reset=True, **kw)
    obs_h = vec.reset()
    obs_d = dev.reset()
    assert obs_h.shape == (n_env, 1 + 28 + 7 * 5) and obs_h.dtype == np.float64
    assert np.array_equal(obs_h, obs_d.cpu().numpy())
    rsum = np.zeros(n_env, np.int64)
    episodes = 0
    for t in range(n_steps):
        act = dev.heuristic("sap_ff")
        a_host = act.cpu().numpy()
        obs_h, rew_h, done_h, infos = vec.step(a_host)
        obs_d, rew_d, done_d, info_d = dev.step(act)
        assert np.array_equal(obs_h, obs_d.cpu().numpy()), t
        assert np.array_equal(rew_h, rew_d.cpu().numpy().astype(np.float32))
        assert np.array_equal(done_h, done_d.cpu().numpy().astype(bool))
        assert np.array_equal([i["accepted"] for i in infos], info_d["accepted"].cpu().numpy().astype(bool))
        rsum += rew_d.cpu().numpy()
        episodes += int(done_h.sum())
    assert episodes == n_env * (n_steps // 29)
    spec = S.make_spec("DeepRMSA-v0", topo, auto_reset=True, **kw)
    _, o_cnt, o_state, o_rsum = OB.rollout_many(spec, 0, n_env, n_steps, S.H_SAP_FF, 4, want_state=True)
    assert np.array_equal(dev.counters().cpu().numpy()[:, :9], o_cnt[:, :9])
    assert np.array_equal(rsum, o_rsum)
    assert np.array_equal(dev.available_slots().cpu().numpy(), o_state)
    st = dev.episode_stats(torch.from_numpy(rsum).to(dev.device))
    assert st["envs"] == n_env and st["fault"] == 0 and st["reward_sum"] == int(o_rsum.sum())
    assert st["services_processed"] == int(o_cnt[:, 0].sum())
    vec.close()
    dev.close()


def test_sharded_engines_equal_single_engine():
    """Global env ids: 2 shards (offsets 0 and n/2) == one engine holding all envs."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine
    from optical_rl_gym_qot_aware_b200.sharding import shard_bounds

    spec = S.make_spec("RMSA-v0", G.topology("spn"), seed=5, load=100, mean_service_holding_time=25,
                       num_spectrum_resources=100, episode_length=60, auto_reset=True)
    n, steps = 333, 150
    whole = Engine(spec, n)
    whole.reset(True)
    whole.rollout(S.H_SAP_FF, steps)
    cw, sw = whole.counters().cpu(), whole.export_state().cpu()
    for r in range(2):
        lo, hi = shard_bounds(n, 2, r)
        part = Engine(spec, hi - lo, global_env_offset=lo)
        part.reset(True)
        part.rollout(S.H_SAP_FF, steps)
        assert torch.equal(part.counters().cpu(), cw[lo:hi])
        assert torch.equal(part.export_state().cpu(), sw[lo:hi])
        part.close()
    whole.close()


def test_matrix_observation_matches_oracle():
    import torch

    import oracle_binding as OB
    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec("RMCSA-v0", G.topology("nsfnet"), seed=8, load=250, mean_service_holding_time=25,
                       num_spectrum_resources=64, num_spatial_resources=7, episode_length=100,
                       obs_kind=S.OBS_MATRIX_U8)
    eng = Engine(spec, 5, obs_layout="env_major")
    eng.reset(True)
    eng.rollout(S.H_SAP_BM_FC_FF, 80)
    obs = eng.new_obs()
    eng.observation(obs)
    for i in (0, 4):
        o = OB.OracleEnv(spec, i)
        o.reset(True)
        for _ in range(80):
            o.step(o.heuristic(S.H_SAP_BM_FC_FF))
        assert np.array_equal(obs[i].cpu().numpy(), o.matrix_observation())
    eng.close()


def test_phy_scalar_env_five_tuple_and_info_reward():
    """make("PhyRMSA-v0") -> 5-tuple step, reference-format heuristic actions, info
    metrics at the reference's point in the step, UseInfoReward."""
    import optical_rl_gym_qot_aware_b200 as pkg
    from optical_rl_gym_qot_aware_b200.envs import UseInfoReward

    g, spec = G.load_case("phy_us14_bmfa_cut_100ch")
    z = np.load(G.os.path.join(G.GOLDEN, f"phytab_{str(g['tables_key'])}.npz"))
    kw = _kwargs(g)
    kw.pop("num_spectrum_resources")
    kw.pop("bit_rate_selection")
    env = pkg.make("PhyRMSA-v0", topology=spec.tables, request_source="replay", modulation_level=z["phy_mod"],
                   connections_detail=z["phy_conn"], gsnr=z["phy_gsnr"], **kw)
    env.load_trace(g["requests"])
    wrapped = UseInfoReward(env, "number_cuts_total")
    for t in range(300):
        path, chans = env.heuristic("phy_aware_bmfa_rmsa")
        want = g["actions"][t]
        assert path == want[0] and [c[0] for c in chans] == [c for c in want[1:] if c >= 0], t
        if chans:  # the last channel may be partially used (phy_rmsa_env.py:1429)
            assert all(c[1] == c[3] and c[2] == 0 for c in chans[:-1]) and chans[-1][1] + chans[-1][2] == chans[-1][3]
        out = wrapped.step((path, chans)) if t % 2 else env.step((path, chans))
        assert len(out) == 5 and out[3] is False
        obs, r, d, _, info = out
        assert info["accepted"] == int(g["accepted"][t]) and d == bool(g["done"][t])
        assert info["number_cuts_total"] == g["stats"][t][6]
        assert abs(info["rss_total_metric"] - g["stats"][t][7]) < 1e-12
        assert r == (info["number_cuts_total"] if t % 2 else int(g["reward"][t]))
        assert set(obs) == {"topology", "current_service"}
        if d:
            env.reset()
    assert env.topology.graph["available_channels"].shape == (spec.tables.n_edges, spec.num_slots)
    env.close()


def test_scalar_env_info_dict_matches_reference():
    """make(..., info_stats=True): env.step's info carries the reference's entries (rmsa_env.py:293-332)."""
    import optical_rl_gym_qot_aware_b200 as pkg

    g, spec = G.load_case("info_rmsa_nsfnet_sapff")
    env = pkg.make("RMSA-v0", topology=spec.tables, request_source="replay", info_stats=True, **_kwargs(g))
    env.load_trace(g["requests"])
    keys = ("service_blocking_rate", "episode_service_blocking_rate", "bit_rate_blocking_rate",
            "episode_bit_rate_blocking_rate", "network_compactness", "network_compactness_difference",
            "avg_link_compactness", "avg_link_utilization")
    for t in range(300):
        _, r, d, info = env.step(env.heuristic("shortest_available_path_first_fit"))
        assert [info[k] for k in keys] == list(g["info_main"][t]), t
        assert info["fairness"] == g["info_fairness"][t]
        assert info["bit_rate_blocking_200"] == g["info_bit_rate_blocking"][t][0]
        if d:
            env.reset()
    env.close()


def test_phy_scalar_env_grooming_virtual_actions():
    """grooming=True through the gym surface: the heuristic returns the reference's virtual-layer action
    (idp + 20, [(channel, share, left over, capacity, True), ...]) (phy_rmsa_env.py:1380,1664-1670) and step takes it."""
    import optical_rl_gym_qot_aware_b200 as pkg

    g, spec = G.load_case("phy_us14_groom_bmfa_100ch")
    assert spec.phy_grooming == 1
    z = np.load(G.os.path.join(G.GOLDEN, f"phytab_{str(g['tables_key'])}.npz"))
    kw = _kwargs(g)
    kw.pop("num_spectrum_resources")
    kw.pop("bit_rate_selection")
    env = pkg.make("PhyRMSA-v0", topology=spec.tables, request_source="replay", modulation_level=z["phy_mod"],
                   connections_detail=z["phy_conn"], gsnr=z["phy_gsnr"], **kw)
    env.load_trace(g["requests"])
    virt = 0
    for t in range(400):
        path, chans = env.heuristic("phy_aware_bmfa_rmsa")
        want = g["actions"][t]
        assert path == want[0] and [c[0] for c in chans] == [c for c in want[1:] if c >= 0], t
        if path >= 20:
            virt += 1
            need = env.current_service.bit_rate / 100
            assert all(c[4] is True for c in chans) and sum(c[1] for c in chans) == need
            assert all(c[2] == 0 for c in chans[:-1]) and chans[-1][2] >= 0
        _, r, d, _, info = env.step((path, chans))
        assert info["accepted"] == int(g["accepted"][t]) and r == int(g["reward"][t]) and d == bool(g["done"][t])
        if d:
            env.reset()
    assert virt > 50
    env.close()


def test_path_only_first_fit_action_wrapper():
    """PathOnlyFirstFitAction (rmsa_env.py:974-1008): path 0 + device first fit == the SP-FF golden trace."""
    import torch

    import optical_rl_gym_qot_aware_b200 as pkg
    from optical_rl_gym_qot_aware_b200.envs import PathOnlyFirstFitAction

    g, spec = G.load_case("rmsa_nsfnet_spff")
    n_env = 2
    env = pkg.make("RMSA-v0", topology=spec.tables, num_envs=n_env, request_source="replay", **_kwargs(g))
    env.load_trace(np.repeat(g["requests"][:, None], n_env, axis=1))
    w = PathOnlyFirstFitAction(env)
    for t in range(400):
        act = w.action(torch.zeros(n_env, dtype=torch.int32)).cpu().numpy()
        assert np.array_equal(act[0], g["actions"][t]), (t, act, g["actions"][t])
        obs, r, d, info = w.step([0, spec.tables.k_paths])  # env 1 always rejects through the wrapper
        assert int(info["accepted"][0]) == int(g["accepted"][t]) and int(info["accepted"][1]) == 0
        if bool(d[0]):
            env.reset()
    env.close()


def test_host_wait_modes_give_the_same_step():
    """orlg_set_host_wait: sleeping on the blocking-sync event and spinning on the stream are two ways to wait
    for the same orlg_step_host_async work; host buffers must hold identical results afterwards."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec("DeepRMSA-v0", G.topology("nsfnet"), seed=11, mean_service_holding_time=7.5,
                       mean_service_inter_arrival_time=1 / 12.0, episode_length=50, auto_reset=True, obs_kind=S.OBS_F32)
    n, outs = 512, []
    for blocking in (False, True):
        eng = Engine(spec, n)
        eng.set_host_wait(blocking)
        eng.reset(True)
        st = torch.cuda.Stream()
        act = torch.zeros((n, 1), dtype=torch.int32).pin_memory()
        rew = torch.zeros(n, dtype=torch.int32).pin_memory()
        done = torch.zeros(n, dtype=torch.uint8).pin_memory()
        obs = torch.zeros((eng.obs_dim, n), dtype=torch.float32).pin_memory()
        trace = []
        for t in range(60):
            act[:, 0] = t % 6
            eng.step_host_async(act, rew, done, obs, None, stream=st)
            eng.step_host_wait(stream=st)
            trace.append((rew.numpy().copy(), done.numpy().copy(), obs.numpy().copy()))
        outs.append(trace)
        eng.close()
    for a, b in zip(*outs):
        assert all(np.array_equal(x, y) for x, y in zip(a, b))


@pytest.mark.gpu
def test_gymenv_views_carry_reference_style_heuristics():
    """Code written against the reference env's attribute surface (env.k_shortest_paths, current_service,
    get_number_slots, is_path_free, get_available_blocks, topology[n1][n2]["index"], topology.graph[...]) runs on a
    GymEnv: a test-local restatement of shortest-available-path first-fit (rmsa_env.py:901-913) and of DeepRMSA's
    K-SP-FF (deeprmsa_env.py:146-155) over those views picks the device heuristic's action at every step; the info
    dict's blocking rates are the reference's (counted before the next request)."""
    import golden_util as G
    from optical_rl_gym_qot_aware_b200 import make

    def sap_ff(env):  # the walk of rmsa_env.py:901-913 over the env's public surface
        cs = env.current_service
        for idp, path in enumerate(env.k_shortest_paths[cs.source, cs.destination]):
            n = env.get_number_slots(path)
            for s0 in range(0, env.topology.graph["num_spectrum_resources"] - n):
                if env.is_path_free(path, s0, n):
                    return (idp, s0)
        return (env.topology.graph["k_paths"], env.topology.graph["num_spectrum_resources"])

    env = make("RMSA-v0", topology=G.topology("nsfnet"), seed=5, load=60, num_spectrum_resources=100, episode_length=40)
    acc = 0
    for t in range(120):
        a = sap_ff(env)
        assert a == env.heuristic("sap_ff"), (t, a)
        proc, accepted = env.services_processed, env.services_accepted
        _, r, d, info = env.step(a)
        acc += r
        # rmsa_env.py:293-296: rates from the counters before _next_service counts the next request
        assert info["service_blocking_rate"] == (proc - (accepted + r)) / proc, (t, info)
        if d:
            env.reset()
    assert 0 < acc < 120
    avail = env.topology.graph["available_slots"]
    assert avail.shape == (22, 100) and env.get_available_slots(env.k_shortest_paths["1", "2"][0]).shape == (100,)
    env.close()

    def ksp_ff(env):  # deeprmsa_env.py:146-155
        for idp in range(env.k_paths):
            starts, _ = env.get_available_blocks(idp)
            if len(starts) > 0:
                return idp * env.j
        return env.k_paths * env.j

    env = make("DeepRMSA-v0", topology=G.topology("nsfnet"), seed=6, mean_service_holding_time=7.5,
               mean_service_inter_arrival_time=1 / 12.0, episode_length=50, j=1)
    for t in range(100):
        a = ksp_ff(env)
        assert a == env.heuristic("ksp_ff"), (t, a)
        _, _, d, _ = env.step(a)
        if d:
            env.reset()
    env.close()


@pytest.mark.gpu
def test_vecenv_infos_and_monitor_rates_under_auto_reset(tmp_path):
    """SB3 indexes infos[i]: one dict per env whatever the env count; the episode_* rates VecMonitor logs when an
    episode ends under auto-reset are those of the finished episode (the reference's info dict), not the counters
    after the in-kernel reset."""
    import golden_util as G
    import oracle_binding as OB
    from optical_rl_gym_qot_aware_b200 import make, spec as S
    from optical_rl_gym_qot_aware_b200.monitor import VecMonitor, load_results

    n = 130
    venv = make("RMSA-v0", topology=G.topology("nsfnet"), num_envs=n, vec_env=True, seed=7, load=60,
                num_spectrum_resources=100, episode_length=30)
    mon = VecMonitor(venv, str(tmp_path / "m"), info_keywords=("episode_service_blocking_rate",))
    mon.reset()
    ora = OB.OracleEnv(S.make_spec("RMSA-v0", G.topology("nsfnet"), seed=7, load=60, num_spectrum_resources=100,
                                   episode_length=30, auto_reset=True), 3)
    ora.reset(True)
    want = []
    for t in range(65):
        act = venv.env.heuristic("sap_ff").cpu().numpy()
        _, _, done, infos = mon.step(act)
        assert isinstance(infos, list) and len(infos) == n and {"accepted", "path", "initial_slot"} <= set(infos[3])
        c = ora.counters()
        _, r, d, _ = ora.step(ora.heuristic(S.H_SAP_FF))
        if d:  # the finished episode's rate, from the counters before the next request was counted
            want.append((c[2] - (c[3] + r)) / c[2])
        assert d == bool(done[3])
    mon.close()
    _, cols = load_results(str(tmp_path / "m") + ".monitor.csv")
    got = cols["episode_service_blocking_rate"]
    assert len(want) == 2 and len(got) == 2 * n
    assert 0.0 <= got.min() and got.max() < 1.0 and abs(got[3] - want[0]) < 1e-6 and abs(got[n + 3] - want[1]) < 1e-6
