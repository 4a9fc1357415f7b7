"""Edge cases of the step path on the GPU: arbitrary (also out-of-range and negative) actions are rejections, not
errors, exactly as in the reference (rmsa_env.py:233-235,270-271; deeprmsa_env.py:53-56; rmcsa_env.py:222-227) and
as the CPU oracle treats them; a release queue that is too small raises the sticky fault bit instead of
corrupting memory; envs counts around the tile sizes."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB
from optical_rl_gym_qot_aware_b200 import spec as S

pytestmark = pytest.mark.gpu

GARBAGE = [
    ("RMSA-v0", "nsfnet", dict(load=60, num_spectrum_resources=100, episode_length=40)),
    ("DeepRMSA-v0", "nsfnet", dict(mean_service_holding_time=7.5, mean_service_inter_arrival_time=1 / 12.0,
                                   episode_length=30, j=2)),
    ("RMCSA-v0", "jpn12_k5", dict(load=200, mean_service_holding_time=25, num_spectrum_resources=320,
                                  num_spatial_resources=7, worst_xt=-84.7, episode_length=50)),
]


@pytest.mark.parametrize("env_id,topo,kw", GARBAGE)
@pytest.mark.parametrize("n_env", [1, 31, 129])
def test_arbitrary_actions_match_the_oracle(env_id, topo, kw, n_env):
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec(env_id, G.topology(topo), seed=77, request_source=S.SRC_PHILOX, auto_reset=True, **kw)
    eng = Engine(spec, n_env, global_env_offset=9)
    eng.reset(True)
    ad = eng.action_dim
    hi = {"RMSA-v0": [7, 110], "DeepRMSA-v0": [13], "RMCSA-v0": [7, 8, 9, 330]}[env_id]
    rng = np.random.default_rng(5)
    steps = 120
    acts = np.stack([rng.integers(-2, h, size=(steps, n_env)) for h in hi], axis=2).astype(np.int32)
    rew = torch.zeros(n_env, dtype=torch.int32, device=eng.device)
    done = torch.zeros(n_env, dtype=torch.uint8, device=eng.device)
    alloc = torch.zeros((n_env, 4), dtype=torch.int32, device=eng.device)
    got = []
    for t in range(steps):
        eng.step(torch.from_numpy(acts[t].reshape(n_env, ad)).to(eng.device), rew, done, None, alloc)
        got.append((rew.cpu().numpy().copy(), done.cpu().numpy().copy(), alloc.cpu().numpy()[:, 0].copy()))
    cnt = eng.counters().cpu().numpy()
    state = eng.export_state().cpu().numpy()
    assert int(cnt[:, 10].max()) == 0
    for i in sorted({0, n_env // 2, n_env - 1}):
        o = OB.OracleEnv(spec, 9 + i)
        o.reset(True)
        accepted = 0
        for t in range(steps):
            _, r, d, al = o.step(acts[t, i])
            assert (r, d, int(al[0])) == (int(got[t][0][i]), bool(got[t][1][i]), int(got[t][2][i])), (i, t)
            accepted += int(al[0])
        assert accepted > 0 and accepted < steps  # the garbage contains valid and invalid decisions
        assert np.array_equal(o.state(), state[i])
    eng.close()


def test_release_queue_overflow_sets_the_fault_bit():
    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec("RMSA-v0", G.topology("nsfnet"), seed=3, load=200, num_spectrum_resources=100,
                       episode_length=1000, request_source=S.SRC_PHILOX, auto_reset=True)
    eng = Engine(spec, 64, heap_capacity=8)  # far fewer entries than services that will be running
    eng.reset(True)
    eng.rollout(S.H_SAP_FF, 200)
    cnt = eng.counters().cpu().numpy()
    assert (cnt[:, 10] != 0).any(), "overflow must be reported through the sticky fault bits"
    eng.close()
    big = Engine(spec, 64)  # provable bound: never overflows
    big.reset(True)
    big.rollout(S.H_SAP_FF, 200)
    assert int(big.counters()[:, 10].max()) == 0
    big.close()
