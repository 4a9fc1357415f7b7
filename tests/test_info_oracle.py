"""The info dict of RMSAEnv.step (rmsa_env.py:276-332): the C oracle against golden traces from
the unmodified reference — blocking rates, network compactness and its difference, the time-weighted link
compactness / utilization means (numpy's pairwise np.mean), fairness and the per-bit-rate blocking, bit for bit."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB

INFO_CASES = [c for c in G.CASES if c.startswith("info_")]


def expected_info(g, t):
    return np.concatenate([g["info_main"][t], [g["info_fairness"][t]], g["info_bit_rate_blocking"][t]])


@pytest.mark.parametrize("case", INFO_CASES)
def test_oracle_info_matches_reference(case):
    g, spec = G.load_case(case)
    spec.info_stats = True
    env = OB.OracleEnv(spec, 0, trace=g["requests"])
    env.reset(full=True)
    for t in range(len(g["actions"])):
        _, r, d, alloc = env.step(g["actions"][t])
        assert alloc[0] == g["accepted"][t]
        got, want = env.info(), expected_info(g, t)
        assert got.shape == want.shape
        assert np.array_equal(got, want), (case, t, np.flatnonzero(got != want), got, want)
        if d:
            env.reset(full=False)
    env.close()
