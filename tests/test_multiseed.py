"""Multi-seed pins (>= 64 seeds x >= 10 000 steps per env id).

tests/golden/ms_*.npz hold, per seed, rolling 64-bit hashes of every decision of the unmodified reference driven by
its own heuristic (oracle/refharness/gen_multiseed_golden.py).  The request stream is regenerated from the seed with
CPython's random (tests/ref_trace.py, itself pinned below by the stored traces).  The oracle (CPU) and the CUDA engine
(GPU, all seeds as one batch through the C-ABI) must reproduce every checkpoint hash, the final counters and the
checksum of the final spectrum - rare branches (release-time ties, LLP-FF ties, 97-slot requests at the first-fit
boundary, a full network) included."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB
from optical_rl_gym_qot_aware_b200 import spec as S
from ref_trace import fold, reference_trace

HID = {"sapff": S.H_SAP_FF, "llpff": S.H_LLP_FF, "kspff": S.H_SAP_FF, "rmcsa": S.H_SAP_BM_FC_FF, "bmfa": S.H_PHY_BM_FA_CUT}


def test_regenerated_traces_equal_the_stored_ones():
    """ref_trace.reference_trace == the request traces the reference itself exported (21 goldens)."""
    n = 0
    for case in G.CASES + G.PHY_CASES + G.QOT_CASES + G.DEFRAG_CASES:
        g, spec = G.load_case(case)
        if "kw_seed" not in g:
            continue
        tr = reference_trace(spec, int(g["kw_seed"]), len(g["requests"]))
        assert all(np.array_equal(tr[k], g["requests"][k]) for k in tr.dtype.names), case
        n += 1
    assert n >= 20


def checksum(state):
    av = np.asarray(state, np.int64).reshape(-1)
    return int(np.dot(av, np.arange(1, av.size + 1, dtype=np.int64)))


@pytest.mark.parametrize("case", G.MS_CASES)
def test_oracle_reproduces_multiseed_hashes(case):
    g, spec = G.load_ms_case(case)
    steps, check, hid = int(g["steps"]), int(g["check"]), HID[str(g["heuristic"])]
    seeds = g["seeds"] if len(g["seeds"]) <= 16 else g["seeds"][::4]  # the CPU suite replays a quarter of the seeds
    for si, seed in enumerate(seeds):
        row = list(g["seeds"]).index(seed)
        env = OB.OracleEnv(spec, 0, trace=reference_trace(spec, int(seed), steps + 1))
        env.reset(full=True)
        h = np.zeros(1, np.uint64)
        for t in range(steps):
            _, r, d, al = env.step(env.heuristic(hid))
            acc = int(al[0])
            path, core = (int(al[1]) & 0xFFFF, int(al[1]) >> 16) if acc else (-1, 0)
            h = fold(h, [acc], [path], [core], [int(al[2]) if acc else -1], [int(al[3]) if acc else 0], [r], [int(d)])
            if d:
                env.reset(full=False)
            if (t + 1) % check == 0:
                assert int(h[0]) == int(g["hashes"][row][(t + 1) // check - 1]), (case, int(seed), t + 1)
        assert np.array_equal(env.counters()[:8], g["counters"][row]), (case, int(seed))
        assert checksum(env.state()) == int(g["final_checksum"][row]), (case, int(seed))
        env.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", G.MS_CASES)
def test_engine_reproduces_multiseed_hashes(case):
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    g, spec = G.load_ms_case(case)
    steps, check, hid = int(g["steps"]), int(g["check"]), HID[str(g["heuristic"])]
    seeds = [int(s) for s in g["seeds"]]
    n = len(seeds)
    trace = np.stack([reference_trace(spec, s, steps + 1) for s in seeds], axis=1)  # [request, env]
    eng = Engine(spec, n)
    eng.load_trace(trace)
    eng.reset(True)
    dev = eng.device
    act = torch.zeros((n, eng.action_dim), dtype=torch.int32, device=dev)
    rew = torch.zeros(n, dtype=torch.int32, device=dev)
    done = torch.zeros(n, dtype=torch.uint8, device=dev)
    alloc = torch.zeros((n, 4), dtype=torch.int32, device=dev)
    h = torch.zeros(n, dtype=torch.int64, device=dev)  # two's-complement wrap-around = uint64 arithmetic
    M = 1000003
    neg1 = torch.full((n,), -1, dtype=torch.int64, device=dev)
    for t in range(steps):
        eng.step_auto(hid, act, rew, done, None, alloc)
        a = alloc.to(torch.int64)
        acc = a[:, 0]
        ok = acc != 0
        cols = (acc, torch.where(ok, a[:, 1] & 0xFFFF, neg1), torch.where(ok, a[:, 1] >> 16, 0 * neg1),
                torch.where(ok, a[:, 2], neg1), torch.where(ok, a[:, 3], 0 * neg1), rew.to(torch.int64), done.to(torch.int64))
        for c in cols:
            h = h * M + c
        if bool(done[0]):  # every env's episode ends at the same step (one request per step, same episode_length)
            assert bool(done.all())
            eng.reset(False)
        if (t + 1) % check == 0:
            got = h.cpu().numpy().view(np.uint64)
            assert np.array_equal(got, g["hashes"][:, (t + 1) // check - 1]), (case, t + 1, np.flatnonzero(got != g["hashes"][:, (t + 1) // check - 1]))
    assert np.array_equal(eng.counters().cpu().numpy()[:, :8], g["counters"])
    st = eng.export_state().cpu().numpy().astype(np.int64).reshape(n, -1)
    assert np.array_equal(st @ np.arange(1, st.shape[1] + 1, dtype=np.int64), g["final_checksum"])
    assert int(eng.counters()[:, 10].max()) == 0
