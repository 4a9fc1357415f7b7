"""Flat, device-friendly tables derived from a reference-style topology object.

The reference hands every env a ``networkx.Graph`` whose ``graph`` dict holds
``ksp`` (dict (src, dst) -> list of Path), ``k_paths``, ``modulations``,
``node_indices`` and ``name`` and whose edges carry ``index``/``id``/``length``
(optical_network_env.py:30-31,58-62; rmsa_env.py:71,481,750).  The engine only
needs the integers behind those objects, so this module flattens them once on
the host, by duck typing (no import of networkx or of the reference):

* ``path_links[pair, p, h]``  link ``index`` of hop ``h`` (-1 padded); pair = src_id*N+dst_id
* ``path_hops``, ``path_length`` (f64), ``path_se`` (spectral efficiency of
  ``best_modulation``), ``path_mod`` (row of that modulation in ``modulations``)
* ``path_links_by_id``  same, but through the edge attribute ``id`` that
  ``get_available_slots`` uses (rmsa_env.py:750); equal to
  ``path_links`` for the .txt topologies, ``None`` when ``id`` is not an int.
* modulation rows: maximum_length, spectral_efficiency, minimum_osnr, inband_xt
  (utils.py:14-24)

The tables round-trip through ``.npz`` so fixtures can travel to a machine that
has neither the reference nor networkx.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np


@dataclass
class TopologyTables:
    name: str
    node_names: List[str]
    k_paths: int
    edge_nodes: np.ndarray  # [E,2] int32 node ids (by node_indices order), row = edge "index"
    edge_length: np.ndarray  # [E] f64
    path_links: np.ndarray  # [N*N,k,Hmax] int16, -1 padded
    path_hops: np.ndarray  # [N*N,k] int32 (0 = no such path)
    path_length: np.ndarray  # [N*N,k] f64
    path_se: np.ndarray  # [N*N,k] int32
    path_mod: np.ndarray  # [N*N,k] int32
    path_nodes: np.ndarray  # [N*N,k,Hmax+1] int16 node ids, -1 padded
    mod_names: List[str]
    mod_max_length: np.ndarray  # [M] f64
    mod_se: np.ndarray  # [M] int32
    mod_min_osnr: np.ndarray  # [M] f64 (nan if None)
    mod_inband_xt: np.ndarray  # [M] f64 (nan if None)
    path_links_by_id: Optional[np.ndarray] = field(default=None)

    @property
    def n_nodes(self) -> int:
        return len(self.node_names)

    @property
    def n_edges(self) -> int:
        return int(self.edge_nodes.shape[0])

    @property
    def max_hops(self) -> int:
        return int(self.path_links.shape[2])

    def pair(self, src_id: int, dst_id: int) -> int:
        return src_id * self.n_nodes + dst_id

    # -- persistence -----------------------------------------------------
    def save(self, path: str) -> None:
        d = {k: v for k, v in self.__dict__.items() if isinstance(v, np.ndarray)}
        d["name"] = np.array(self.name)
        d["node_names"] = np.array(self.node_names)
        d["mod_names"] = np.array(self.mod_names)
        d["k_paths"] = np.array(self.k_paths)
        np.savez_compressed(path, **d)

    @staticmethod
    def load(path: str) -> "TopologyTables":
        z = np.load(path, allow_pickle=False)
        kw = {k: z[k] for k in z.files}
        kw["name"] = str(kw["name"])
        kw["node_names"] = [str(x) for x in kw["node_names"]]
        kw["mod_names"] = [str(x) for x in kw["mod_names"]]
        kw["k_paths"] = int(kw["k_paths"])
        kw.setdefault("path_links_by_id", None)
        return TopologyTables(**kw)


def _f(x) -> float:
    return float("nan") if x is None else float(x)


def tables_from_topology(topology) -> TopologyTables:
    """Flatten a reference-style topology (see module docstring)."""
    g = topology.graph
    node_names = [str(x) for x in g["node_indices"]]
    nid = {name: i for i, name in enumerate(g["node_indices"])}
    n = len(node_names)
    k = int(g["k_paths"])
    mods = list(g["modulations"])

    n_edges = topology.number_of_edges()
    edge_nodes = np.full((n_edges, 2), -1, np.int32)
    edge_length = np.zeros(n_edges, np.float64)
    ids_are_int = True
    for u, v, d in topology.edges(data=True):
        e = int(d["index"])
        edge_nodes[e] = (nid[u], nid[v])
        edge_length[e] = float(d["length"])
        if not isinstance(d.get("id"), (int, np.integer)):
            ids_are_int = False

    ksp = g["ksp"]
    hmax = max(len(p.node_list) - 1 for paths in ksp.values() for p in paths)
    shape = (n * n, k)
    path_links = np.full(shape + (hmax,), -1, np.int16)
    path_links_by_id = np.full(shape + (hmax,), -1, np.int16) if ids_are_int else None
    path_nodes = np.full(shape + (hmax + 1,), -1, np.int16)
    path_hops = np.zeros(shape, np.int32)
    path_length = np.zeros(shape, np.float64)
    path_se = np.zeros(shape, np.int32)
    path_mod = np.full(shape, -1, np.int32)

    def mod_row(m):
        for i, x in enumerate(mods):
            if x is m or x == m:
                return i
        return -1

    for (s, t), paths in ksp.items():
        pr = nid[s] * n + nid[t]
        for ip, p in enumerate(paths[:k]):
            nl = list(p.node_list)
            path_hops[pr, ip] = len(nl) - 1
            path_length[pr, ip] = float(p.length)
            for h in range(len(nl) - 1):
                ed = topology[nl[h]][nl[h + 1]]
                path_links[pr, ip, h] = int(ed["index"])
                if ids_are_int:
                    path_links_by_id[pr, ip, h] = int(ed["id"])
            for h, name in enumerate(nl):
                path_nodes[pr, ip, h] = nid[name]
            bm = getattr(p, "best_modulation", None)
            if bm is not None:
                path_se[pr, ip] = int(bm.spectral_efficiency)
                path_mod[pr, ip] = mod_row(bm)

    return TopologyTables(
        name=str(g.get("name", "topology")),
        node_names=node_names,
        k_paths=k,
        edge_nodes=edge_nodes,
        edge_length=edge_length,
        path_links=path_links,
        path_hops=path_hops,
        path_length=path_length,
        path_se=path_se,
        path_mod=path_mod,
        path_nodes=path_nodes,
        mod_names=[str(m.name) for m in mods],
        mod_max_length=np.array([float(m.maximum_length) for m in mods], np.float64),
        mod_se=np.array([int(m.spectral_efficiency) for m in mods], np.int32),
        mod_min_osnr=np.array([_f(m.minimum_osnr) for m in mods], np.float64),
        mod_inband_xt=np.array([_f(m.inband_xt) for m in mods], np.float64),
        path_links_by_id=path_links_by_id,
    )


DATA_DIR = __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.abspath(__file__)), "data")


def builtin_topology(short: str) -> TopologyTables:
    """Tables of the topologies the reference ships (examples/topologies/*.h5), flattened once and kept with the
    package: nsfnet, us14, jpn12_k3, jpn12_k5, spn, germany50."""
    import os

    return TopologyTables.load(os.path.join(DATA_DIR, f"topology_{short}.npz"))


def builtin_qot_tables(name: str = "Results_K3SP_FRP_SLC_CBG_USB14"):
    """The QoT tables of examples/phy_frag_rmsa/inputs/*.mat as numpy arrays: (modulation_level, gsnr,
    connections_detail), as tests/test_rmsa_us.py:43-49 unpacks them."""
    import os

    z = np.load(os.path.join(DATA_DIR, f"phytab_{name}.npz"), allow_pickle=False)
    return z["phy_mod"], z["phy_gsnr"], z["phy_conn"]
