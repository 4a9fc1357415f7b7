"""Host-side tables of PhyRMSA-v0 (optical_rl_gym/envs/phy_rmsa_env.py).

The reference looks everything physical up in three arrays it receives as constructor
kwargs (phy_rmsa_env.py:51-53,93-95): ``connections_detail`` (one row per unordered node
pair; columns 0/1 are the node names as numbers), ``modulation_level[row][channel][idp]``
(uint8 level, 100 Gb/s per level per channel, :1398,1424) and ``gsnr[row][channel][idp]``
(f64 dB, :596).  This module flattens them once:

* ``table_id[src_id*N + dst_id]``   first row whose (col0, col1) matches the pair in either
  order — the ``np.where(...)[0][0]`` scan of phy_rmsa_env.py:562-565,1382-1385;
* ``mod_level`` / ``gsnr``          ``[rows][channels][k_paths]`` (the first k path columns);
* ``cut_links`` / ``cut_count``     for every (pair, idp): the neighbour link of every term of
  ``calculate_r_cut(..., modified=True)`` (:1140-1193) — for each path node, each adjacent node
  that is not on the path contributes one term (end nodes) or two (interior nodes).  When the
  channel is free on the whole path the metric is ``T - 2 * sum(bit[neighbour link])``.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .tables import TopologyTables

BIT_RATES = (100, 200, 300, 400, 500, 600)  # phy_rmsa_env.py:38
MAX_CHANNELS_PER_SERVICE = 8  # action = (path, 8 channel numbers); 600 Gb/s at level 1 needs 6


@dataclass
class PhyTables:
    n_tables: int
    num_channels: int
    table_id: np.ndarray  # [N*N] int32, -1 = pair absent
    mod_level: np.ndarray  # [T, C, k] uint8
    gsnr: np.ndarray  # [T, C, k] f64
    cut_links: np.ndarray  # [N*N*k, max_terms] uint8
    cut_count: np.ndarray  # [N*N*k] int32

    @property
    def max_terms(self) -> int:
        return int(self.cut_links.shape[1])


def _scalar(x) -> int:
    return int(np.asarray(x).reshape(-1)[0])


def num_channels(number_spectrum_channels: int = 80, number_spectrum_channels_s_band: int = 108,
                 l_band: bool = True, s_band: bool = True) -> int:
    """topology.graph["num_channel_resources"], optical_network_env.py:78-101."""
    if s_band:
        return 2 * number_spectrum_channels + number_spectrum_channels_s_band
    return 2 * number_spectrum_channels if l_band else number_spectrum_channels


def build_phy_tables(tables: TopologyTables, modulation_level, connections_detail, gsnr,
                     channels: int) -> PhyTables:
    n, k = tables.n_nodes, tables.k_paths
    mod = np.asarray(modulation_level)
    gs = np.asarray(gsnr, np.float64)
    assert mod.ndim == 3 and gs.shape == mod.shape, "modulation_level / gsnr must be [rows, channels, paths]"
    assert mod.shape[1] >= channels and mod.shape[2] >= k, "tables smaller than channels x k_paths"
    rows = mod.shape[0]
    names = []
    for nm in tables.node_names:
        try:
            names.append(int(nm))
        except ValueError:
            names.append(None)
    cd = np.asarray(connections_detail, dtype=object)
    c0 = [_scalar(cd[r, 0]) for r in range(rows)]
    c1 = [_scalar(cd[r, 1]) for r in range(rows)]
    table_id = np.full(n * n, -1, np.int32)
    for s in range(n):
        for d in range(n):
            if s == d or names[s] is None or names[d] is None:
                continue
            for r in range(rows):  # first match in either direction
                if (c0[r] == names[s] and c1[r] == names[d]) or (c0[r] == names[d] and c1[r] == names[s]):
                    table_id[s * n + d] = r
                    break
    # adjacency: node -> [(neighbour, link)]
    adj = [[] for _ in range(n)]
    for l, (a, b) in enumerate(tables.edge_nodes):
        adj[int(a)].append((int(b), l))
        adj[int(b)].append((int(a), l))
    terms = []
    for pr in range(n * n):
        for ip in range(k):
            h = int(tables.path_hops[pr, ip])
            t = []
            if h > 0:
                nl = [int(x) for x in tables.path_nodes[pr, ip, : h + 1]]
                on = set(nl)
                for i, node in enumerate(nl):
                    for nb, link in adj[node]:
                        if nb in on:
                            continue
                        t.extend([link] if i in (0, h) else [link, link])
            terms.append(t)
    mt = max(1, max(len(t) for t in terms))
    cut_links = np.zeros((n * n * k, mt), np.uint8)
    cut_count = np.zeros(n * n * k, np.int32)
    for i, t in enumerate(terms):
        cut_links[i, : len(t)] = t
        cut_count[i] = len(t)
    return PhyTables(
        n_tables=rows, num_channels=int(channels), table_id=table_id,
        mod_level=np.ascontiguousarray(mod[:, :channels, :k].astype(np.uint8)),
        gsnr=np.ascontiguousarray(gs[:, :channels, :k]),
        cut_links=cut_links, cut_count=cut_count,
    )


def load_qot_mat(path: str, key: str = None):
    """The precomputed QoT tables of a MATLAB ``Results_*.mat`` file as the reference's drivers unpack them
    (tests/test_rmsa_us.py:43-49): the struct ``mat[key][0][0]`` holds ``connections_detail`` (field 0),
    ``modulation_level`` (field 1, uint8 ``[pairs][channels][paths]``) and ``gsnr`` (field 2, f64).  ``key``
    defaults to the file's stem (``Results_K3SP_FRP_SLC_CBG_USB14``).  Returns a dict with the three
    constructor kwargs of PhyRMSA-v0 (phy_rmsa_env.py:51-53); the modulation levels are widened to int64
    so that sums and sort keys cannot wrap in uint8, the node columns of
    ``connections_detail`` are unwrapped to plain integers."""
    import os

    from scipy.io import loadmat

    if key is None:
        key = os.path.splitext(os.path.basename(path))[0]
    mat = loadmat(path)
    if key not in mat:
        raise KeyError(f"{path}: no variable {key!r} (have {sorted(k for k in mat if not k.startswith('__'))})")
    d = mat[key][0][0]
    conn = d[0]
    conn2 = np.array([[_scalar(conn[r, c]) for c in (0, 1)] for r in range(conn.shape[0])], np.int32)
    return dict(connections_detail=conn2, modulation_level=np.asarray(d[1]).astype(np.int64),
                gsnr=np.ascontiguousarray(np.asarray(d[2], np.float64)))
