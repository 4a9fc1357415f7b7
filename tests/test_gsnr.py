"""GN-model GSNR (examples/calculate_osnr.py:9-56).

Golden vectors: tests/golden/gsnr_*.npz hold numbers produced by the reference's own
calculate_osnr (oracle/refharness/gen_gsnr_golden.py) on network states reached by a stored
request/action trace.  The C oracle (CPU) and the CUDA kernel (GPU, through orlg_gsnr) replay
the trace and must agree with them; north star tolerance 1e-4 dB, observed ~1e-11."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB
from optical_rl_gym_qot_aware_b200 import spec as S

TOL_DB = 1e-4  # BASELINE.json north_star: "GSNR must agree within 1e-4 dB"


def _close(a, b, tol):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    with np.errstate(invalid="ignore"):
        return bool(np.all(both_inf | (np.abs(a - b) <= tol)))


@pytest.mark.parametrize("case", G.GSNR_CASES)
def test_oracle_gsnr_matches_reference_function(case):
    g, spec = G.load_case(case)
    gp = S.gsnr_params(spec.tables)
    env = OB.OracleEnv(spec, 0, trace=g["requests"])
    env.reset(full=True)
    qi = 0
    for t in range(len(g["actions"])):
        if qi < len(g["q_step"]) and g["q_step"][qi] == t:
            out = env.gsnr(gp, g["q_query"][qi])
            assert _close(out, g["q_out"][qi], 1e-9), (case, t, out, g["q_out"][qi])
            qi += 1
        _, _, d, _ = env.step(g["actions"][t])
        if d:
            env.reset(full=False)
    assert qi == len(g["q_step"]) > 20
    bad = env.gsnr(gp, [spec.tables.k_paths, 0, 4, 0])  # reject action -> NaN
    assert np.isnan(bad).all()


def test_gsnr_params_follow_the_reference_builder():
    t = G.topology("jpn12_k5")
    gp = S.gsnr_params(t)
    # examples/create_topology_gn.py:122-134: int(length // 80) + 1 equal spans
    assert all(int(c) == int(L // 80) + 1 for c, L in zip(gp.span_count, t.edge_length))
    assert np.allclose(gp.span_count * gp.span_length_km, t.edge_length, rtol=0, atol=1e-9)
    assert gp.span_length_km.max() <= 80.0
    # 0.2 dB/km power loss == exp(-2 alpha L)
    assert abs(10 * np.log10(np.exp(2 * gp.attenuation_per_m * 1000.0)) - 0.2) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("case", G.GSNR_CASES)
def test_engine_gsnr_matches_reference_function(case):
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    g, spec = G.load_case(case)
    n_env = 3
    eng = Engine(spec, n_env)
    eng.gsnr_configure()
    eng.load_trace(np.repeat(g["requests"][:, None], n_env, axis=1))
    eng.reset(True)
    done = torch.zeros(n_env, dtype=torch.uint8, device=eng.device)
    qi, worst = 0, 0.0
    for t in range(len(g["actions"])):
        if qi < len(g["q_step"]) and g["q_step"][qi] == t:
            q = torch.from_numpy(np.repeat(g["q_query"][qi][None], n_env, axis=0).astype(np.int32)).to(eng.device)
            out = eng.gsnr(q).cpu().numpy()
            assert _close(out[0], g["q_out"][qi], TOL_DB) and _close(out[2], g["q_out"][qi], TOL_DB), (case, t, out)
            fin = np.isfinite(g["q_out"][qi])
            worst = max(worst, float(np.abs(out[1][fin] - g["q_out"][qi][fin]).max()))
            qi += 1
        act = torch.from_numpy(np.repeat(g["actions"][t][None], n_env, axis=0).astype(np.int32)).to(eng.device)
        eng.step(act, None, done, None, None)
        if bool(done[0]):
            eng.reset(False)
    assert qi == len(g["q_step"])
    print(f"{case}: worst |dGSNR| = {worst:.3e} dB")
    bad = torch.tensor([[spec.tables.k_paths, 0, 4, 0]] * n_env, dtype=torch.int32, device=eng.device)
    assert torch.isnan(eng.gsnr(bad)).all()
    eng.close()


@pytest.mark.gpu
def test_engine_gsnr_matches_oracle_on_philox_envs():
    """Many different network states: heuristic candidates of 200 Philox-driven envs vs the oracle."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec("RMSA-v0", G.topology("spn"), seed=17, load=300, mean_service_holding_time=25,
                       num_spectrum_resources=320, episode_length=500)
    gp = S.gsnr_params(spec.tables)
    n_env, steps = 200, 260
    eng = Engine(spec, n_env)
    eng.gsnr_configure(gp)
    eng.reset(True)
    eng.rollout(S.H_SAP_FF, steps)
    act = torch.zeros((n_env, 2), dtype=torch.int32, device=eng.device)
    eng.heuristic(S.H_SAP_FF, act)
    a = act.cpu().numpy()
    rq = eng.requests()
    t = spec.tables
    q = np.zeros((n_env, 4), np.int32)
    for i in range(n_env):
        if a[i, 0] < t.k_paths:
            se = t.path_se[rq["src"][i] * t.n_nodes + rq["dst"][i], a[i, 0]]
            q[i] = (a[i, 0], a[i, 1], int(np.ceil(rq["bit_rate"][i] / (se * 12.5))) + 1, 0)
        else:
            q[i] = (-1, 0, 0, 0)
    out = eng.gsnr(torch.from_numpy(q).to(eng.device)).cpu().numpy()
    checked = 0
    for i in range(0, n_env, 5):
        o = OB.OracleEnv(spec, i)
        o.reset(True)
        for _ in range(steps):
            o.step(o.heuristic(S.H_SAP_FF))
        ref = o.gsnr(gp, q[i])
        if q[i, 0] < 0:
            assert np.isnan(out[i]).all() and np.isnan(ref).all()
        else:
            assert _close(out[i], ref, TOL_DB), (i, out[i], ref)
            checked += 1
    assert checked >= 20
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("env_id", ["RMSA-v0", "RMCSA-v0"])
def test_gsnr_for_actions_equals_explicit_queries(env_id):
    """orlg_gsnr_for_actions on the heuristic's own action tensor == orlg_gsnr on the queries built on the
    host from the same actions (slot count from the pending bit rate, rmsa_env.py:708-719 /
    rmcsa_env.py:753-765): bit-identical outputs, NaN for rejections."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    if env_id == "RMSA-v0":
        spec = S.make_spec(env_id, G.topology("spn"), seed=5, load=300, mean_service_holding_time=25,
                           num_spectrum_resources=320, episode_length=500)
        hid = S.H_SAP_FF
    else:
        spec = S.make_spec(env_id, G.topology("jpn12_k5"), seed=5, load=250, mean_service_holding_time=25,
                           num_spectrum_resources=320, num_spatial_resources=7, worst_xt=-84.7, episode_length=500)
        hid = S.H_SAP_BM_FC_FF
    n_env = 300
    eng = Engine(spec, n_env)
    eng.gsnr_configure()
    eng.reset(True)
    eng.rollout(hid, 400)
    act = torch.zeros((n_env, eng.action_dim), dtype=torch.int32, device=eng.device)
    eng.heuristic(hid, act)
    a = act.cpu().numpy()
    a[::17, 0] = spec.tables.k_paths  # a few rejection actions
    rq, t = eng.requests(), spec.tables
    q = np.full((n_env, 4), -1, np.int32)
    for i in range(n_env):
        if a[i, 0] >= t.k_paths:
            continue
        pr = rq["src"][i] * t.n_nodes + rq["dst"][i]
        if env_id == "RMSA-v0":
            se, core, start = int(t.path_se[pr, a[i, 0]]), 0, a[i, 1]
        else:
            se, core, start = int(t.mod_se[a[i, 1]]), a[i, 2], a[i, 3]
        q[i] = (a[i, 0] | (core << 16), start, int(np.ceil(rq["bit_rate"][i] / (se * spec.channel_width))) + 1, se)
    want = eng.gsnr(torch.from_numpy(q).to(eng.device)).cpu().numpy()
    got = eng.gsnr_for_actions(torch.from_numpy(a).to(eng.device)).cpu().numpy()
    assert np.isnan(got[::17]).all() and np.isfinite(got[:, 0]).sum() >= n_env // 2
    assert np.array_equal(got, want, equal_nan=True)
    # the warp-cooperative kernel (link bit sets, ballot-compacted interferers, one log per asinh pair, butterfly
    # sums, tabulated asinh pair) against the generic one it replaced (literal formula, sequential sums): 1e-7 dB
    import os

    os.environ["ORLG_GSNR_GENERIC"] = "1"
    try:
        generic = eng.gsnr_for_actions(torch.from_numpy(a).to(eng.device)).cpu().numpy()
    finally:
        os.environ.pop("ORLG_GSNR_GENERIC")
    assert np.array_equal(np.isnan(got), np.isnan(generic)) and np.array_equal(np.isinf(got), np.isinf(generic))
    fin = np.isfinite(got)
    assert np.abs(got[fin] - generic[fin]).max() < 1e-7  # the tabulated asinh pair uses exact grid distances
    eng.close()
