"""Helpers shared by the golden-vector tests: load a case and rebuild its EnvSpec."""
import os

import numpy as np

from optical_rl_gym_qot_aware_b200 import spec as S
from optical_rl_gym_qot_aware_b200.tables import TopologyTables

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

TOPO_SHORT = {
    "nsfnet_chen_5-paths_6-modulations.h5": "nsfnet",
    "us14_3-paths_6-modulations.h5": "us14",
    "jpn12_5-paths_6-modulations.h5": "jpn12_k5",
    "jpn12_3-paths_6-modulations.h5": "jpn12_k3",
    "spn_3-paths_6-modulations.h5": "spn",
    "germany50_5-paths_6-modulations.h5": "germany50",
}

CASES = sorted(f[:-4] for f in os.listdir(GOLDEN)
               if f.endswith(".npz") and not f.startswith(("topology_", "gsnr_", "phy_", "phytab_", "defrag_", "qot_", "ms_")))
PHY_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN) if f.startswith("phy_") and f.endswith(".npz"))
QOT_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN) if f.startswith("qot_") and f.endswith(".npz"))
MS_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN) if f.startswith("ms_") and f.endswith(".npz"))
DEFRAG_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN) if f.startswith("defrag_phy_") and f.endswith(".npz"))
GSNR_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN) if f.startswith("gsnr_") and f.endswith(".npz"))
ENV_NAME = {0: "RMSA-v0", 1: "DeepRMSA-v0", 2: "RMCSA-v0", 3: "PhyRMSA-v0"}


def topology(short: str) -> TopologyTables:
    return TopologyTables.load(os.path.join(GOLDEN, f"topology_{short}.npz"))


def load_case(name: str):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    g = {k: z[k] for k in z.files}
    kw = {}
    for k in g["kwargs_keys"]:
        v = g["kw_" + str(k)]
        if v.dtype == object or (v.ndim == 0 and str(v) == "None"):
            kw[str(k)] = None
        elif v.ndim == 0:
            kw[str(k)] = v.item()
        else:
            kw[str(k)] = v
    tables = topology(TOPO_SHORT[str(g["topo_file"])])
    if "tables_key" in g:  # PhyRMSA: the QoT tables the reference run used
        z = np.load(os.path.join(GOLDEN, f"phytab_{str(g['tables_key'])}.npz"), allow_pickle=False)
        kw.update(modulation_level=z["phy_mod"].astype(np.int64), connections_detail=z["phy_conn"],
                  gsnr=z["phy_gsnr"])
        kw.pop("num_spectrum_resources", None)  # irrelevant to PhyRMSA's allocation
        kw.pop("bit_rate_selection", None)
        if "channel_state" in g:  # 0: use_existing_channels stubbed out in the harness, 1: grooming=True, 2: unstubbed
            kw["keep_channel_state"] = int(g["channel_state"]) != 0
    if "adaptive" in g and "gsnr" in g:  # GN-model QoT gate traces (oracle/refharness/gen_qot_golden.py)
        kw.update(qot="gn", qot_adaptive=bool(int(g["adaptive"])))
    spec = S.make_spec(ENV_NAME[int(g["env_kind"])], tables, request_source=S.SRC_REPLAY, **kw)
    return g, spec


def load_ms_case(name: str):
    """Multi-seed hash golden (oracle/refharness/gen_multiseed_golden.py): (fixture dict, EnvSpec for replay)."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    g = {k: z[k] for k in z.files}
    kw = {}
    for k in g["kwargs_keys"]:
        v = g["kw_" + str(k)]
        kw[str(k)] = v.item() if v.ndim == 0 else v
    env_id = str(g["env_id"])
    tables = topology(TOPO_SHORT[str(g["topo_file"])])
    if env_id == "PhyRMSA-v0":  # the C4 bench shape: 320 channels = the shipped 268 + the S-band tail replicated
        t = np.load(os.path.join(GOLDEN, "phytab_Results_K3SP_FRP_SLC_CBG_USB14.npz"), allow_pickle=False)
        cols = np.concatenate([np.arange(268), np.arange(268 - 52, 268)])
        kw.update(modulation_level=t["phy_mod"].astype(np.int64)[:, cols], gsnr=t["phy_gsnr"][:, cols],
                  connections_detail=t["phy_conn"])
        kw.pop("num_spectrum_resources", None)
        kw.pop("bit_rate_selection", None)
    spec = S.make_spec(env_id, tables, request_source=S.SRC_REPLAY, **kw)
    return g, spec
