"""The Philox request source against the reference's process (MT19937 cannot be matched bit for bit on
the device, so throughput runs use per-env Philox streams).  Same distributions?  Two samples at fixed seeds - the
reference's stream regenerated with CPython's random (tests/ref_trace.py, pinned by the stored traces) and the
Philox stream of the oracle (bit-identical to the CUDA engine's, tests/test_gpu_envs.py) - are compared with each
other and with the distributions `_next_service` draws from (rmsa_env.py:646-659, optical_network_env.py:191-208):
exponential inter-arrival and holding times, weighted source, destination over the renormalised weights, bit rates."""
import ctypes as C

import numpy as np

import golden_util as G
import oracle_binding as OB
from optical_rl_gym_qot_aware_b200 import spec as S
from ref_trace import REQ_DTYPE, reference_trace

PROBS = [0.01801802, 0.04004004, 0.05305305, 0.01901902, 0.04504505, 0.02402402, 0.06706707,
         0.08908909, 0.13813814, 0.12212212, 0.07607608, 0.12012012, 0.01901902, 0.16916917]


def philox_trace(spec, genv, n):
    L = OB.lib()
    os_ = OB.OracleSpec(spec)
    out = np.zeros(n, REQ_DTYPE)
    one = np.zeros(1, REQ_DTYPE)
    now = 0.0
    for r in range(n):
        L.oracle_philox_request(C.byref(os_.cfg), genv, r, C.c_double(now), one.ctypes.data)
        out[r] = one[0]
        now = float(one[0]["arrival"])
    return out


def chi2(obs, exp):
    exp = np.asarray(exp, np.float64) * obs.sum() / np.sum(exp)
    m = exp > 0
    return float(((obs[m] - exp[m]) ** 2 / exp[m]).sum()), int(m.sum()) - 1


def test_philox_source_draws_the_reference_distributions():
    spec = S.make_spec("RMSA-v0", G.topology("nsfnet"), seed=77, load=50, node_request_probabilities=PROBS)
    n = 120_000
    mt = reference_trace(spec, 77, n)
    ph = np.concatenate([philox_trace(spec, g, n // 4) for g in (0, 1, 2, 65535)])
    N = spec.tables.n_nodes
    p = np.asarray(PROBS) / np.sum(PROBS)
    # destination given source: weights with the source zeroed, renormalised (optical_network_env.py:201-203)
    joint = np.outer(p, p)
    np.fill_diagonal(joint, 0.0)
    joint = joint / (1.0 - p)[:, None]
    for name, tr in (("mt19937", mt), ("philox", ph)):
        iat = np.diff(np.concatenate([[0.0], tr["arrival"]])) if name == "mt19937" else None
        if name == "philox":  # four streams, each restarting its clock at 0
            iat = np.concatenate([np.diff(np.concatenate([[0.0], ph["arrival"][i * (n // 4):(i + 1) * (n // 4)]])) for i in range(4)])
        # exponential: mean and standard deviation both equal the configured mean (3 sigma of the sample mean ~ 0.9 %)
        assert abs(iat.mean() / spec.mean_iat - 1) < 0.01 and abs(iat.std() / spec.mean_iat - 1) < 0.02, (name, iat.mean())
        assert abs(tr["holding"].mean() / spec.mean_holding - 1) < 0.01 and abs(tr["holding"].std() / spec.mean_holding - 1) < 0.02, name
        assert (tr["src"] != tr["dst"]).all()
        x, dof = chi2(np.bincount(tr["src"], minlength=N), p)
        assert x < 40.0 and dof == N - 1, (name, "src", x)                # chi2(13): P(x > 40) ~ 1e-4
        pair = np.bincount(tr["src"].astype(np.int64) * N + tr["dst"], minlength=N * N)
        x, dof = chi2(pair, joint.reshape(-1))
        assert x < 260.0 and dof == N * N - N - 1, (name, "pair", x)      # chi2(181): P(x > 260) ~ 1e-4
        br = np.bincount((tr["bit_rate"] - 200) // 50, minlength=21)
        x, dof = chi2(br, np.ones(21))
        assert x < 50.0 and dof == 20, (name, "bit rate", x)              # chi2(20): P(x > 50) ~ 2e-4
    # the two samples against each other: homogeneity chi-square on the (src, dst) table
    a = np.bincount(mt["src"].astype(np.int64) * N + mt["dst"], minlength=N * N).astype(np.float64)
    b = np.bincount(ph["src"].astype(np.int64) * N + ph["dst"], minlength=N * N).astype(np.float64)
    m = (a + b) > 0
    x = float((((a[m] - b[m]) ** 2) / (a[m] + b[m])).sum())
    assert x < 260.0, x


def test_steady_state_blocking_of_philox_and_reference_streams_agree():
    """RMSA-v0 with SAP-FF on the two request streams (oracle; 40 envs x 3 000 steps each after 600 warm-up steps):
    the blocking probabilities agree within the sampling error."""
    kw = dict(load=50, num_spectrum_resources=100, episode_length=100000)
    tables = G.topology("nsfnet")
    rates = {}
    for src in ("mt19937", "philox"):
        acc = proc = 0
        for i in range(40):
            if src == "philox":
                spec = S.make_spec("RMSA-v0", tables, seed=5, request_source=S.SRC_PHILOX, **kw)
                env = OB.OracleEnv(spec, i)
            else:
                spec = S.make_spec("RMSA-v0", tables, request_source=S.SRC_REPLAY, **kw)
                env = OB.OracleEnv(spec, 0, trace=reference_trace(spec, 500 + i, 3700))
            env.reset(True)
            for t in range(3600):
                _, r, _, _ = env.step(env.heuristic(S.H_SAP_FF))
                if t >= 600:
                    acc += r
                    proc += 1
            env.close()
        rates[src] = 1.0 - acc / proc
    assert 0.3 < rates["philox"] < 0.8 and abs(rates["philox"] - rates["mt19937"]) < 0.02, rates
