"""PhyRMSA-v0: the C oracle replays the golden traces exported from the unmodified reference
(oracle/refharness/gen_phy_golden.py) bit for bit — actions of the reference heuristics,
accept/block, reward, done, counters, episode sums, fragmentation metrics, final channel map."""
import numpy as np
import pytest

import golden_util as G
import oracle_binding as OB

HEUR = {"bmfa_cut": 8, "bmfa": 8, "sapff": 5, "bmff": 7, "sapbm": 6, "faff": 10, "bmfarss": 11, "faffrss": 12}


def heuristic_of(case):
    return next(v for k, v in sorted(HEUR.items(), key=lambda kv: -len(kv[0])) if k in case)


@pytest.mark.parametrize("case", G.PHY_CASES)
def test_oracle_replays_phy_golden(case):
    g, spec = G.load_case(case)
    assert spec.env_kind == 3 and spec.phy is not None
    hid = heuristic_of(case)
    env = OB.OracleEnv(spec, 0, trace=g["requests"])
    env.reset(full=True)
    n = len(g["actions"])
    for t in range(n):
        assert env.request() == g["requests"][t]
        a = env.heuristic(hid)
        assert np.array_equal(a, g["actions"][t]), (case, t, a, g["actions"][t])
        _, r, d, alloc = env.step(g["actions"][t])
        assert alloc[0] == g["accepted"][t] and r == g["reward"][t] and d == bool(g["done"][t]), (case, t)
        st = env.phy_stats()
        assert np.array_equal(st[:6], g["stats"][t][:6]), (case, t, st, g["stats"][t])  # f64 sums bit-exact
        assert st[6] == g["stats"][t][6], (case, t)                                       # number_cuts_total
        assert abs(st[7] - g["stats"][t][7]) <= 1e-12 * abs(g["stats"][t][7]), (case, t)  # rss_total_metric
        if d:
            env.reset(full=False)
        assert np.array_equal(env.counters()[:8], g["counters"][t]), (case, t)
    assert np.array_equal(env.state(), g["final_avail"])
    env.close()


def test_phy_tables_first_match_and_cut_terms():
    g, spec = G.load_case(G.PHY_CASES[0])
    ph, t = spec.phy, spec.tables
    n = t.n_nodes
    assert (ph.table_id.reshape(n, n) == ph.table_id.reshape(n, n).T).all()
    assert (np.diag(ph.table_id.reshape(n, n)) == -1).all() and (ph.table_id.reshape(n, n)[~np.eye(n, dtype=bool)] >= 0).all()
    # end nodes contribute one term per off-path neighbour, interior nodes two (phy_rmsa_env.py:1147-1163)
    deg = np.zeros(n, int)
    for a, b in t.edge_nodes:
        deg[a] += 1
        deg[b] += 1
    pr, ip = 0 * n + 5, 0
    h = t.path_hops[pr, ip]
    nodes = list(t.path_nodes[pr, ip, : h + 1])
    on = set(nodes)
    adj = {i: set() for i in range(n)}
    for a, b in t.edge_nodes:
        adj[a].add(b)
        adj[b].add(a)
    expect = sum((1 if i in (0, h) else 2) * len(adj[v] - on) for i, v in enumerate(nodes))
    assert ph.cut_count[pr * t.k_paths + ip] == expect


def avail_checksum(state):
    """the checksum of available_channels the golden generator stores after every step (gen_phy_golden.py)"""
    av = np.asarray(state, np.int64).ravel()
    return int(np.dot(av, np.arange(1, av.size + 1, dtype=np.int64)))


@pytest.mark.parametrize("case", G.DEFRAG_CASES)
def test_oracle_replays_defragmentation_golden(case):
    """Periodic defragmentation (phy_rmsa_env.py:355-417, 662-764): the unmodified reference ran with defrag_period /
    number_moves / metric; per step its heuristic action, accept / block, counters, the cumulative move counters
    (counted_moves, counted_moves_groom, counted_defrag_cycles) and a checksum of available_channels are pinned, so
    every grooming move and every reallocation has to land on the channel the reference chose."""
    g, spec = G.load_case(case)
    assert spec.phy_defrag_period == int(g["kw_defrag_period"]) and spec.phy_number_moves == int(g["kw_number_moves"])
    hid = heuristic_of(case.replace("defrag_", "").replace("_cut_", "_bmfa_cut_").replace("_rss_", "_bmfarss_"))
    env = OB.OracleEnv(spec, 0, trace=g["requests"])
    env.reset(full=True)
    for t in range(len(g["actions"])):
        a = env.heuristic(hid)
        assert np.array_equal(a, g["actions"][t]), (case, t, a, g["actions"][t])
        _, r, d, alloc = env.step(g["actions"][t])
        assert alloc[0] == g["accepted"][t] and r == g["reward"][t] and d == bool(g["done"][t]), (case, t)
        assert np.array_equal(env.phy_moves(), g["moves"][t]), (case, t, env.phy_moves(), g["moves"][t])
        assert avail_checksum(env.state()) == int(g["avail_sum"][t]), (case, t)
        if d:
            env.reset(full=False)
        assert np.array_equal(env.counters()[:8], g["counters"][t]), (case, t)
    assert np.array_equal(env.state(), g["final_avail"])
    assert g["moves"][:, 0].max() > 0 and g["moves"][:, 2].max() > 0
    env.close()
