"""Offline topology builder: topology file -> k shortest paths -> reach-based modulation -> flat tables.

What the reference does once per topology with networkx and pickle
(examples/create_topology.py:96-147, examples/graph_utils.py:31-116, optical_rl_gym/utils.py:94-117)
and what its envs then read back from ``graph["ksp"]`` / ``graph["modulations"]`` /
``graph["node_indices"]``.  Here the result goes straight into the ``TopologyTables`` the engine
consumes (``tables.py``), with no networkx and no pickle on the way, so a user can bring a new
topology file to the GPU box.

The k shortest simple paths are the ones ``networkx.shortest_simple_paths(G, s, t, weight="length")``
yields (utils.py:94-98) — third-party arithmetic that is not under /root/reference (networkx,
unpinned by the reference's setup.py; 3.6.1 in the build container).  Its published algorithm is
restated here, tie-breaking included, because the shipped ``.txt`` topologies have integer lengths
and equal-cost paths do occur:

* Yen's algorithm over a candidate heap ordered by ``(cost, push counter)``; a candidate that is
  already waiting in the heap is not pushed twice; for every prefix ("root") of the previous path the
  edges that earlier results take out of the same root are hidden, and so are the root's inner nodes;
* each spur path comes from a *bidirectional* Dijkstra that alternates forward / backward pops,
  orders its fringes by ``(distance, push counter)``, relaxes on strict ``<``, visits neighbours in
  edge-insertion order, and stops at the first node settled from both sides, returning the best
  meeting point *seen so far* (first one wins ties).

Pinned by the tables flattened from the reference's own pickles (``tests/golden/topology_*.npz``):
``tests/test_topology_builder.py`` rebuilds all six from their node / edge lists and compares every
array bit for bit.
"""
from __future__ import annotations

import argparse
import heapq
import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from .tables import TopologyTables

# examples/create_topology.py:45-94 — the modulation set every shipped "*_6-modulations.h5" was built with:
# (name, maximum_length km, spectral_efficiency, minimum_osnr dB, inband_xt dB)
DEFAULT_MODULATIONS: Tuple[Tuple[str, float, int, float, float], ...] = (
    ("BPSK", 100_000, 1, 12.6, -14),
    ("QPSK", 2_000, 2, 12.6, -17),
    ("8QAM", 1_000, 3, 18.6, -20),
    ("16QAM", 500, 4, 22.4, -23),
    ("32QAM", 250, 5, 26.4, -26),
    ("64QAM", 125, 6, 30.4, -29),
)


class TopologyGraph:
    """Undirected graph with the iteration orders the results depend on: nodes in insertion order,
    a node's neighbours in the order its edges were added (what a networkx ``Graph`` gives)."""

    def __init__(self) -> None:
        self.node_names: List[str] = []
        self.node_id: Dict[str, int] = {}
        self.adj: List[Dict[int, int]] = []  # node -> {neighbour: edge index}, insertion ordered
        self.edge_nodes: List[Tuple[int, int]] = []
        self.edge_length: List[float] = []
        self.edge_ids: List[object] = []  # the edge attribute "id" (int for .txt files, str for SNDlib)

    def add_node(self, name: str) -> int:
        name = str(name)
        if name not in self.node_id:
            self.node_id[name] = len(self.node_names)
            self.node_names.append(name)
            self.adj.append({})
        return self.node_id[name]

    def add_edge(self, u: str, v: str, length, edge_id=None) -> int:
        a, b = self.add_node(u), self.add_node(v)
        if b in self.adj[a]:  # a repeated line only updates the attributes (networkx add_edge)
            e = self.adj[a][b]
            self.edge_length[e] = length
            return e
        e = len(self.edge_nodes)
        self.adj[a][b] = e
        self.adj[b][a] = e
        self.edge_nodes.append((a, b))
        self.edge_length.append(length)
        self.edge_ids.append(e if edge_id is None else edge_id)
        return e

    @property
    def n_nodes(self) -> int:
        return len(self.node_names)


# ---- readers (examples/graph_utils.py) ---------------------------------------------------------------
def read_txt_topology(text: str) -> TopologyGraph:
    """graph_utils.py:88-116: lines starting with '#' are comments; first line = number of nodes (named
    "1".."N"), second = number of links, then "a b length" with integer lengths; id = index = line order."""
    g = TopologyGraph()
    lines = [ln for ln in text.splitlines(keepends=True) if not ln.startswith("#")]
    for idx, line in enumerate(lines):
        if idx == 0:
            for i in range(1, int(line) + 1):
                g.add_node(str(i))
        elif idx == 1:
            int(line)  # the link count is parsed and otherwise unused, as in the reference
        elif len(line) > 1:
            info = line.replace("\n", "").split(" ")
            g.add_edge(info[0], info[1], int(info[2]))
    return g


def _geographical_distance(p1: Tuple[float, float], p2: Tuple[float, float]) -> float:
    """graph_utils.py:9-28: haversine on a 6 373 km sphere.  The reference passes ``pos = (x, y)`` and reads
    element 0 as the latitude, i.e. SNDlib's x (a longitude) plays the latitude; reproduced as is."""
    r = 6373.0
    lat1, lon1 = math.radians(p1[0]), math.radians(p1[1])
    lat2, lon2 = math.radians(p2[0]), math.radians(p2[1])
    dlon, dlat = lon2 - lon1, lat2 - lat1
    a = math.sin(dlat / 2) ** 2 + math.cos(lat1) * math.cos(lat2) * math.sin(dlon / 2) ** 2
    return r * (2 * math.atan2(math.sqrt(a), math.sqrt(1 - a)))


def read_sndlib_topology(xml_text: str) -> TopologyGraph:
    """graph_utils.py:31-85: SNDlib native XML; length = great-circle (or Euclidean) distance of the end
    nodes rounded to 3 decimals; edge id = the link's id string, index = document order."""
    import xml.dom.minidom

    doc = xml.dom.minidom.parseString(xml_text).documentElement
    coords = doc.getElementsByTagName("nodes")[0].getAttribute("coordinatesType")
    g = TopologyGraph()
    pos: Dict[str, Tuple[float, float]] = {}
    for node in doc.getElementsByTagName("node"):
        x = node.getElementsByTagName("x")[0].childNodes[0].data
        y = node.getElementsByTagName("y")[0].childNodes[0].data
        g.add_node(node.getAttribute("id"))
        pos[node.getAttribute("id")] = (float(x), float(y))
    for link in doc.getElementsByTagName("link"):
        s = link.getElementsByTagName("source")[0].childNodes[0].data
        t = link.getElementsByTagName("target")[0].childNodes[0].data
        if coords == "geographical":
            length = np.around(_geographical_distance(pos[s], pos[t]), 3)
        else:
            length = np.around(math.sqrt((pos[s][0] - pos[t][0]) ** 2 + (pos[s][1] - pos[t][1]) ** 2), 3)
        g.add_edge(s, t, float(length), edge_id=link.getAttribute("id"))
    return g


def graph_from_tables(t: TopologyTables) -> TopologyGraph:
    """The node / edge lists of existing tables as a graph (edge index order = insertion order)."""
    g = TopologyGraph()
    for name in t.node_names:
        g.add_node(name)
    for (a, b), ln in zip(t.edge_nodes.tolist(), t.edge_length.tolist()):
        e = g.add_edge(t.node_names[a], t.node_names[b], ln)
        if t.path_links_by_id is None:  # the source had non-integer edge ids (SNDlib): keep it that way
            g.edge_ids[e] = f"L{e}"
    return g


# ---- k shortest simple paths (networkx.shortest_simple_paths restated, see module docstring) ---------
def _bidirectional_dijkstra(g: TopologyGraph, source: int, target: int, ignore_nodes: set, ignore_edges: set):
    if source in ignore_nodes or target in ignore_nodes:
        return None
    if source == target:
        return 0, [source]
    dists: List[Dict[int, float]] = [{}, {}]
    paths: List[Dict[int, List[int]]] = [{source: [source]}, {target: [target]}]
    fringe: List[list] = [[], []]
    seen: List[Dict[int, float]] = [{source: 0}, {target: 0}]
    counter = 0
    heapq.heappush(fringe[0], (0, counter, source))
    counter += 1
    heapq.heappush(fringe[1], (0, counter, target))
    counter += 1
    finaldist, finalpath = None, []
    direction = 1
    while fringe[0] and fringe[1]:
        direction = 1 - direction
        dist, _, v = heapq.heappop(fringe[direction])
        if v in dists[direction]:
            continue
        dists[direction][v] = dist
        if v in dists[1 - direction]:
            return finaldist, finalpath
        for w, e in g.adj[v].items():
            if w in ignore_nodes or (v, w) in ignore_edges or (w, v) in ignore_edges:
                continue
            vw = dists[direction][v] + g.edge_length[e]
            if w in dists[direction]:
                continue
            if w not in seen[direction] or vw < seen[direction][w]:
                seen[direction][w] = vw
                heapq.heappush(fringe[direction], (vw, counter, w))
                counter += 1
                paths[direction][w] = paths[direction][v] + [w]
                if w in seen[0] and w in seen[1]:
                    total = seen[0][w] + seen[1][w]
                    if not finalpath or finaldist > total:
                        finaldist = total
                        finalpath = paths[0][w] + paths[1][w][::-1][1:]
    return None


def k_shortest_paths(g: TopologyGraph, source: int, target: int, k: int) -> List[List[int]]:
    """The first ``k`` paths of networkx.shortest_simple_paths(G, source, target, weight="length")."""

    def length_of(path: Sequence[int]):
        tot = 0
        for u, v in zip(path, path[1:]):
            tot = tot + g.edge_length[g.adj[u][v]]
        return tot

    found: List[List[int]] = []
    heap: list = []
    waiting: set = set()
    counter = 0

    def push(cost, path):
        nonlocal counter
        key = tuple(path)
        if key not in waiting:
            heapq.heappush(heap, (cost, counter, path))
            counter += 1
            waiting.add(key)

    prev: Optional[List[int]] = None
    while len(found) < k:
        if prev is None:
            r = _bidirectional_dijkstra(g, source, target, set(), set())
            if r is not None:
                push(r[0], r[1])
        else:
            ignore_nodes: set = set()
            ignore_edges: set = set()
            for i in range(1, len(prev)):
                root = prev[:i]
                root_length = length_of(root)
                for p in found:
                    if p[:i] == root:
                        ignore_edges.add((p[i - 1], p[i]))
                r = _bidirectional_dijkstra(g, root[-1], target, ignore_nodes, ignore_edges)
                if r is not None:
                    push(root_length + r[0], root[:-1] + r[1])
                ignore_nodes.add(root[-1])
        if not heap:
            break
        _, _, path = heapq.heappop(heap)
        waiting.discard(tuple(path))
        found.append(path)
        prev = path
    return found


def best_modulation_row(length: float, modulations: Sequence[Tuple[str, float, int, float, float]]) -> int:
    """utils.py:105-117: the most spectrally efficient format whose maximum_length covers the path
    (stable sort by efficiency, descending); ValueError when none does."""
    order = sorted(range(len(modulations)), key=lambda i: modulations[i][2], reverse=True)
    for i in order:
        if length <= modulations[i][1]:
            return i
    raise ValueError("It was not possible to find a suitable MF for a path with {} km".format(length))


def build_tables(g: TopologyGraph, name: str, k_paths: int = 5,
                 modulations: Sequence[Tuple[str, float, int, float, float]] = DEFAULT_MODULATIONS) -> TopologyTables:
    """examples/create_topology.py:96-147 (``get_topology``) into ``TopologyTables``: for every node pair
    ``idn1 < idn2`` in node order the k shortest paths by length, their length (``np.sum`` of the edge
    lengths, utils.py:101-102) and best modulation; the reverse pair shares the same path objects, i.e.
    the same node order (``k_shortest_paths[n2, n1] = objs``)."""
    n, k = g.n_nodes, int(k_paths)
    per_pair: Dict[Tuple[int, int], List[List[int]]] = {}
    for a in range(n):
        for b in range(a + 1, n):
            per_pair[a, b] = k_shortest_paths(g, a, b, k)
    hmax = max((len(p) - 1 for ps in per_pair.values() for p in ps), default=1)
    shape = (n * n, k)
    path_links = np.full(shape + (hmax,), -1, np.int16)
    ids_are_int = all(isinstance(i, (int, np.integer)) for i in g.edge_ids)
    path_links_by_id = np.full(shape + (hmax,), -1, np.int16) if ids_are_int else None
    path_nodes = np.full(shape + (hmax + 1,), -1, np.int16)
    path_hops = np.zeros(shape, np.int32)
    path_length = np.zeros(shape, np.float64)
    path_se = np.zeros(shape, np.int32)
    path_mod = np.full(shape, -1, np.int32)
    for (a, b), ps in per_pair.items():
        for ip, p in enumerate(ps):
            edges = [g.adj[p[h]][p[h + 1]] for h in range(len(p) - 1)]
            length = np.sum([g.edge_length[e] for e in edges])
            row = best_modulation_row(length, modulations)
            for pr in (a * n + b, b * n + a):
                path_hops[pr, ip] = len(edges)
                path_length[pr, ip] = float(length)
                path_links[pr, ip, :len(edges)] = edges
                if ids_are_int:
                    path_links_by_id[pr, ip, :len(edges)] = [int(g.edge_ids[e]) for e in edges]
                path_nodes[pr, ip, :len(p)] = p
                path_se[pr, ip] = int(modulations[row][2])
                path_mod[pr, ip] = row
    return TopologyTables(
        name=str(name), node_names=list(g.node_names), k_paths=k,
        # orientation as networkx's edges() reports it: the end node that comes first in node order first
        edge_nodes=np.array([(min(a, b), max(a, b)) for a, b in g.edge_nodes], np.int32).reshape(-1, 2),
        edge_length=np.array([float(x) for x in g.edge_length], np.float64),
        path_links=path_links, path_hops=path_hops, path_length=path_length, path_se=path_se, path_mod=path_mod,
        path_nodes=path_nodes, mod_names=[str(m[0]) for m in modulations],
        mod_max_length=np.array([float(m[1]) for m in modulations], np.float64),
        mod_se=np.array([int(m[2]) for m in modulations], np.int32),
        mod_min_osnr=np.array([float("nan") if m[3] is None else float(m[3]) for m in modulations], np.float64),
        mod_inband_xt=np.array([float("nan") if m[4] is None else float(m[4]) for m in modulations], np.float64),
        path_links_by_id=path_links_by_id,
    )


def build_from_file(path: str, name: Optional[str] = None, k_paths: int = 5,
                    modulations: Sequence[Tuple[str, float, int, float, float]] = DEFAULT_MODULATIONS) -> TopologyTables:
    """create_topology.py:98-103: ``.xml`` = SNDlib, ``.txt`` = node/link list; anything else is an error."""
    with open(path, "r") as f:
        text = f.read()
    if path.endswith(".xml"):
        g = read_sndlib_topology(text)
    elif path.endswith(".txt"):
        g = read_txt_topology(text)
    else:
        raise ValueError("Supplied topology is unknown")
    if name is None:
        name = path.replace("\\", "/").split("/")[-1].rsplit(".", 1)[0]
    return build_tables(g, name, k_paths, modulations)


def main(argv=None) -> None:
    ap = argparse.ArgumentParser(description="topology file -> flat k-shortest-path / modulation tables (.npz)")
    ap.add_argument("-t", "--topology", required=True, help=".txt or SNDlib .xml topology file")
    ap.add_argument("-k", "--k_paths", type=int, default=5)
    ap.add_argument("-n", "--name", default=None)
    ap.add_argument("-o", "--output", default=None)
    a = ap.parse_args(argv)
    t = build_from_file(a.topology, a.name, a.k_paths)
    out = a.output or f"{t.name}_{t.k_paths}-paths_{len(t.mod_names)}-modulations.npz"
    t.save(out)
    print(f"{out}: {t.n_nodes} nodes, {t.n_edges} links, k={t.k_paths}, max hops {t.max_hops}")


if __name__ == "__main__":
    main()
