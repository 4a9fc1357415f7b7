"""GN-model QoT gate (include/orlg.h "QoT gate"): the C oracle against traces in which the unmodified
reference envs step behind a gate whose every GSNR is a call of the reference's own calculate_osnr
(oracle/refharness/gen_qot_golden.py).  Decisions, slot counts, rewards and the final spectrum must be
identical; GSNR values within the north star's 1e-4 dB (observed ~1e-12)."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB
from optical_rl_gym_qot_aware_b200 import spec as S

TOL_DB = 1e-4


def close(a, b, tol=TOL_DB):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    same = (np.isnan(a) & np.isnan(b)) | (np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b)))
    with np.errstate(invalid="ignore"):
        return bool(np.all(same | (np.abs(a - b) <= tol)))


@pytest.mark.parametrize("case", G.QOT_CASES)
def test_oracle_gate_replays_reference_trace(case):
    g, spec = G.load_case(case)
    assert spec.qot_gate and spec.qot_adaptive == bool(int(g["adaptive"]))
    env = OB.OracleEnv(spec, 0, trace=g["requests"])
    env.reset(full=True)
    by_heur = bool(int(g["by_heuristic"]))
    n_eval = 0
    for t in range(len(g["actions"])):
        if by_heur:  # heuristic 13 must propose what the harness's walk over the reference env proposed
            a = env.heuristic(S.H_SAP_GN_FF)
            assert np.array_equal(a, g["actions"][t]), (case, t, a, g["actions"][t])
            assert close(env.qot_last(), g["heur_gsnr"][t]), (case, t, env.qot_last(), g["heur_gsnr"][t])
        _, r, d, alloc = env.step(g["actions"][t])
        assert alloc[0] == g["accepted"][t] and alloc[3] == g["slots"][t], (case, t, alloc)
        assert r == g["rewards"][t] and d == bool(g["dones"][t])
        assert close(env.qot_last(), g["gsnr"][t]), (case, t, env.qot_last(), g["gsnr"][t])
        n_eval += int(np.isfinite(g["gsnr"][t][0]))
        if d:
            env.reset(full=False)
    assert n_eval > 300
    assert np.array_equal(env.state().reshape(g["final_state"].shape), g["final_state"])


def test_thresholds_are_the_sd_fec_vector_of_the_gn_builder():
    spec = S.make_spec("RMSA-v0", G.topology("us14"), num_spectrum_resources=320, qot="gn", qot_adaptive=True)
    # examples/create_topology_gn.py:65 target_SNR_dB_vec, minimum_osnr of BPSK .. 64QAM (:68-110)
    assert list(spec.qot_thr_db[1:7]) == [3.71925646843142, 6.72955642507124, 10.8453935345953, 13.2406469649752,
                                          16.1608982942870, 19.0134649345090]
    assert list(spec.qot_se) == [6, 5, 4, 3, 2, 1]
    with pytest.raises(ValueError):
        S.make_spec("DeepRMSA-v0", G.topology("nsfnet"), qot="gn")
