"""Topology builder against the tables flattened from the reference's own pickles.

The golden ``topology_*.npz`` fixtures were made by ``tables_from_topology`` from the shipped ``.h5``
pickles (networkx Yen k-shortest paths + utils.get_best_modulation_format, create_topology.py:96-147);
the builder must reproduce every array from the node / edge lists alone, ties included."""
import glob
import os

import numpy as np
import pytest

from optical_rl_gym_qot_aware_b200 import topology_builder as tb
from optical_rl_gym_qot_aware_b200.tables import TopologyTables

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
REF_TOPOLOGIES = "/root/reference/examples/topologies"


def assert_same_tables(a: TopologyTables, b: TopologyTables):
    for key, v in a.__dict__.items():
        w = getattr(b, key)
        if isinstance(v, np.ndarray):
            assert isinstance(w, np.ndarray) and v.shape == w.shape and v.dtype == w.dtype, key
            assert np.array_equal(v, w, equal_nan=v.dtype.kind == "f"), key
        else:
            assert not isinstance(w, np.ndarray) and v == w, key


@pytest.mark.parametrize("fixture", sorted(glob.glob(os.path.join(GOLDEN, "topology_*.npz"))),
                         ids=lambda p: os.path.basename(p)[9:-4])
def test_builder_reproduces_reference_tables(fixture):
    t = TopologyTables.load(fixture)
    b = tb.build_tables(tb.graph_from_tables(t), t.name, t.k_paths)  # default modulations = create_topology.py:45-94
    assert_same_tables(t, b)


def test_txt_reader_and_ties():
    # two equal-cost routes 1-2-4 / 1-3-4: the order is the bidirectional search's, not alphabetical luck
    g = tb.read_txt_topology("# comment\n4\n5\n1 3 100\n1 2 100\n3 4 100\n2 4 100\n1 4 500\n")
    assert g.node_names == ["1", "2", "3", "4"] and g.edge_nodes == [(0, 2), (0, 1), (2, 3), (1, 3), (0, 3)]
    paths = tb.k_shortest_paths(g, 0, 3, 5)
    assert paths == [[0, 2, 3], [0, 1, 3], [0, 3]]
    t = tb.build_tables(g, "toy", 3)
    assert t.path_hops[t.pair(0, 3)].tolist() == [2, 2, 1] and t.path_hops[t.pair(3, 0)].tolist() == [2, 2, 1]
    assert t.path_length[t.pair(0, 3)].tolist() == [200.0, 200.0, 500.0]
    assert t.path_se[t.pair(0, 3)].tolist() == [5, 5, 4]  # 200 km -> 32QAM (<= 250), 500 km -> 16QAM
    assert t.path_nodes[t.pair(3, 0), 0, :3].tolist() == [0, 2, 3]  # the reverse pair shares the forward node order
    assert t.path_links_by_id is not None and np.array_equal(t.path_links_by_id, t.path_links)
    with pytest.raises(ValueError):
        tb.best_modulation_row(200_000, tb.DEFAULT_MODULATIONS)


def test_sndlib_reader():
    xml = """<?xml version="1.0"?><network><networkStructure><nodes coordinatesType="geographical">
    <node id="A"><coordinates><x>10.0</x><y>50.0</y></coordinates></node>
    <node id="B"><coordinates><x>11.0</x><y>50.0</y></coordinates></node>
    <node id="C"><coordinates><x>10.0</x><y>51.0</y></coordinates></node></nodes>
    <links><link id="L1"><source>A</source><target>B</target></link>
    <link id="L2"><source>B</source><target>C</target></link></links></networkStructure></network>"""
    g = tb.read_sndlib_topology(xml)
    assert g.node_names == ["A", "B", "C"] and g.edge_ids == ["L1", "L2"]
    assert g.edge_length[0] == pytest.approx(111.23, abs=0.01)  # one degree of "latitude" (x) on R = 6373 km and round(g.edge_length[0], 3) == g.edge_length[0]
    t = tb.build_tables(g, "tri", 2)
    assert t.path_links_by_id is None  # string ids: get_available_slots' "id" indexing is unusable
    assert t.path_hops[t.pair(0, 2)].tolist() == [2, 0]


@pytest.mark.skipif(not os.path.isdir(REF_TOPOLOGIES), reason="reference checkout not present")
@pytest.mark.parametrize("fname,fixture,k", [("nsfnet_chen.txt", "nsfnet", 5), ("jpn12.txt", "jpn12_k3", 3),
                                             ("us14.txt", "us14", 3), ("spn.txt", "spn", 3),
                                             ("germany50.xml", "germany50", 5)])
def test_builder_from_reference_files(fname, fixture, k):
    t = TopologyTables.load(os.path.join(GOLDEN, f"topology_{fixture}.npz"))
    b = tb.build_from_file(os.path.join(REF_TOPOLOGIES, fname), t.name, k)
    assert_same_tables(t, b)


REF_INPUTS = "/root/reference/examples/phy_frag_rmsa/inputs"


@pytest.mark.skipif(not os.path.isdir(REF_INPUTS), reason="reference checkout not present")
@pytest.mark.parametrize("key", ["Results_K3SP_FRP_SLC_CBG_USB14", "Results_K3SP_FRP_SLC_CBG_JPN12"])
def test_qot_mat_ingestion_matches_fixture(key):
    from optical_rl_gym_qot_aware_b200.phy import load_qot_mat

    z = np.load(os.path.join(GOLDEN, f"phytab_{key}.npz"))
    q = load_qot_mat(os.path.join(REF_INPUTS, key + ".mat"))
    k = z["phy_mod"].shape[2]
    assert np.array_equal(q["connections_detail"], z["phy_conn"])
    assert np.array_equal(q["modulation_level"][:, :, :k], z["phy_mod"]) and q["modulation_level"].dtype == np.int64
    assert np.array_equal(q["gsnr"][:, :, :k], z["phy_gsnr"])
