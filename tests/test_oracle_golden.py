"""The CPU oracle replays every golden trace exported from the unmodified
reference and must reproduce it bit for bit (accept/block, reward, done,
allocation, counters, DeepRMSA f64 observation, final available_slots)."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB

HEUR_OF_CASE = {
    "rmsa_nsfnet_sapff": 1, "rmsa_nsfnet_spff": 0, "rmsa_nsfnet_llpff": 2,
    "rmsa_spn_sapff": 1, "rmsa_spn320_sapff": 1,
    "deeprmsa_nsfnet_kspff": 1, "deeprmsa_nsfnet_spff": 0,
    "rmcsa_jpn12_heur": 3, "rmcsa_nsfnet_heur": 3,
}


@pytest.mark.parametrize("case", G.CASES)
def test_oracle_replays_golden(case):
    g, spec = G.load_case(case)
    env = OB.OracleEnv(spec, 0, trace=g["requests"])
    env.reset(full=True)
    kind = spec.env_kind
    if kind == 1:
        assert np.array_equal(env.observation(), g["obs0"])
    n = len(g["actions"])
    for t in range(n):
        rq = env.request()
        assert rq == g["requests"][t]
        if case in HEUR_OF_CASE:  # the oracle's own heuristic picks the reference's action
            assert np.array_equal(env.heuristic(HEUR_OF_CASE[case]), g["actions"][t]), (case, t)
        obs, r, d, alloc = env.step(g["actions"][t], want_obs=(kind == 1))
        assert alloc[0] == g["accepted"][t], (case, t)
        assert r == g["reward"][t] and d == bool(g["done"][t]), (case, t)
        if alloc[0]:
            path, slot, ns, core = g["alloc"][t]
            assert (alloc[1] & 0xFFFF, alloc[1] >> 16, alloc[2], alloc[3]) == (path, core, slot, ns), (case, t)
        if d:
            env.reset(full=False)
        assert np.array_equal(env.counters()[:8], g["counters"][t]), (case, t)
        if kind == 1:
            assert np.array_equal(obs, g["obs"][t]), (case, t)
    assert np.array_equal(env.state(), g["final_avail"])
    env.close()
