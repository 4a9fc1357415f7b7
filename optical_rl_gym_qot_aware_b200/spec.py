"""Host-side flattening of the reference's constructor kwargs into one POD spec.

``EnvSpec`` carries exactly what ``orlg_config`` (include/orlg.h) needs: the
flat topology tables plus the scalar env parameters, evaluated on the host with
the reference's own expressions so that every value the device compares against
is bit-identical to what the Python reference would have computed:

* ``mean_iat = 1 / (load / holding)``        optical_network_env.py:127-129
* default bit rates 200..1200 step 50        rmsa_env.py:37-38
* cumulative weights of ``random.choices``   optical_network_env.py:197-206
* RMCSA QoT factors (10**x terms)            rmcsa_env.py:341-384, +4 dB rmcsa_env.py:127-129
"""
from __future__ import annotations

import itertools
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

from .tables import TopologyTables

ENV_RMSA, ENV_DEEPRMSA, ENV_RMCSA, ENV_PHYRMSA = 0, 1, 2, 3
SRC_REPLAY, SRC_PHILOX = 0, 1
OBS_NONE, OBS_F64, OBS_F32, OBS_MATRIX_U8, OBS_PHY_INFO, OBS_PHY_CUTS = 0, 1, 2, 3, 4, 5
H_SP_FF, H_SAP_FF, H_LLP_FF, H_SAP_BM_FC_FF, H_RANDOM = 0, 1, 2, 3, 4
H_PHY_SAP_FF, H_PHY_SAP_BM, H_PHY_BM_FF, H_PHY_BM_FA_CUT = 5, 6, 7, 8  # phy_rmsa_env.py:1676,1254,1317,1375
H_PATH_FF, H_PHY_FA_FF = 9, 10  # rmsa_env.py:974-1008, phy_rmsa_env.py:1508-1569
H_PHY_BM_FA_RSS, H_PHY_FA_FF_RSS = 11, 12  # phy_rmsa_env.py:1441-1505, :1572-1633
H_SAP_GN_FF = 13  # shortest available path, first fit, first candidate that passes the GN-model QoT gate

# SD-FEC SNR thresholds by spectral efficiency 1..6 (target_SNR_dB_vec, examples/create_topology_gn.py:65): the
# minimum_osnr the GN topology builder attaches to BPSK .. 64QAM (:68-110)
SD_FEC_SNR_DB = (3.71925646843142, 6.72955642507124, 10.8453935345953, 13.2406469649752, 16.1608982942870,
                 19.0134649345090)

ENV_IDS = {
    "RMSA-v0": ENV_RMSA,
    "DeepRMSA-v0": ENV_DEEPRMSA,
    "RMCSA-v0": ENV_RMCSA,
    "RMCSA": ENV_RMCSA,
    "PhyRMSA-v0": ENV_PHYRMSA,
}

DEFAULT_BIT_RATES = tuple(range(200, 1201, 50))  # rmsa_env.py:37-38
WORST_XT_BY_CORE = {7: -84.7, 12: -61.9, 19: -54.8}  # rmcsa_env.py:63-67


def _cum(weights) -> np.ndarray:
    # random.choices: cum_weights = list(itertools.accumulate(weights))
    return np.array(list(itertools.accumulate(float(w) for w in weights)), np.float64)


@dataclass
class EnvSpec:
    env_kind: int
    tables: TopologyTables
    num_slots: int = 100
    num_cores: int = 1
    j: int = 1
    episode_length: int = 1000
    allow_rejection: bool = False
    auto_reset: bool = False
    channel_width: float = 12.5
    mean_holding: float = 10800.0
    mean_iat: float = 1080.0
    load: float = 10.0
    bit_rates: Optional[np.ndarray] = None  # None => continuous randint(lo, hi)
    bit_rate_cum: Optional[np.ndarray] = None
    bit_rate_lo: int = 25
    bit_rate_hi: int = 100
    node_probs: np.ndarray = field(default_factory=lambda: np.zeros(0))
    src_cum: np.ndarray = field(default_factory=lambda: np.zeros(0))
    dst_cum: np.ndarray = field(default_factory=lambda: np.zeros(0))
    mod_reach: Optional[np.ndarray] = None
    mod_snr_coeff: Optional[np.ndarray] = None
    worst_xt: Optional[float] = None
    seed: int = 41
    request_source: int = SRC_PHILOX
    obs_kind: int = OBS_NONE
    phy: Optional[object] = None  # phy.PhyTables for PhyRMSA-v0
    # PhyRMSA virtual layer: 0 = channel_state not kept (physical layer only), 1 = grooming=True,
    # 2 = grooming=False but channel_state kept, i.e. what the reference does: sapff/sapbm/bmff/faff/faff_rss
    # try use_existing_channels whatever env.grooming says (phy_rmsa_env.py:1256,1321,1510,1574,1678)
    phy_grooming: int = 0
    # RMSA-v0 / DeepRMSA-v0: keep what the step's info dict needs (rmsa_env.py:276-332): time-weighted link
    # statistics, per-bit-rate histograms, sum of slots x hops of the running services
    info_stats: bool = False
    # GN-model QoT gate (RMSA-v0 / RMCSA-v0; include/orlg.h "QoT gate"): the GSNR of the candidate and of its
    # neighbours (examples/calculate_osnr.py:9-56) must meet the SD-FEC threshold of their modulations
    # PhyRMSA periodic defragmentation (phy_rmsa_env.py:355-417): every phy_defrag_period processed services
    phy_defrag_period: int = 0
    phy_number_moves: int = 0
    phy_metric: int = 0  # 0 "cut" (calculate_r_cut), 1 "rss" (calculate_r_spatial)
    qot_gate: bool = False
    qot_adaptive: bool = False
    qot_thr_db: Optional[np.ndarray] = None   # [16] threshold in dB by spectral efficiency
    qot_se: Optional[np.ndarray] = None       # spectral efficiencies an adaptive RMSA-v0 env tries, descending
    qot_gsnr: Optional[object] = None         # GsnrParams

    @property
    def action_dim(self) -> int:
        return {ENV_RMSA: 2, ENV_DEEPRMSA: 1, ENV_RMCSA: 4, ENV_PHYRMSA: 9}.get(self.env_kind, 2)

    @property
    def obs_dim(self) -> int:
        t = self.tables
        if self.obs_kind in (OBS_F64, OBS_F32):
            return 1 + 2 * t.n_nodes + (2 * self.j + 3) * t.k_paths  # deeprmsa_env.py:35-37
        if self.obs_kind == OBS_MATRIX_U8:
            return 2 * t.n_nodes + self.num_cores * t.n_edges * self.num_slots  # rmsa_env.py:943-946
        if self.obs_kind in (OBS_PHY_INFO, OBS_PHY_CUTS):
            return 2 if self.obs_kind == OBS_PHY_INFO else 1
        return 0

    @property
    def reject_action(self):
        t = self.tables
        if self.env_kind == ENV_RMSA:
            return (t.k_paths, self.num_slots)
        if self.env_kind == ENV_DEEPRMSA:
            return (t.k_paths * self.j,)
        if self.env_kind == ENV_PHYRMSA:
            return (-2,) + (-1,) * 8  # (-2, []) phy_rmsa_env.py:1423
        return (t.k_paths, len(t.mod_names), self.num_cores, self.num_slots)


def rmcsa_qot_factors(tables: TopologyTables, worst_xt: float):
    """Modulation-only factors of ``_crosstalk_is_acceptable`` (rmcsa_env.py:341-384),
    evaluated with the reference's expression order.  ``worst_xt`` is the user
    value; the +4 dB the constructor adds to it and to every ``inband_xt``
    (rmcsa_env.py:127-129) is applied here."""
    nf = 10.0 ** (5.5 / 10.0)
    amp_gain = 10.0 ** (20 / 10.0)
    h = 6.626068e-34
    f_hz = 2.99e8 / (1550 * 1e-9)
    wxt = worst_xt + 4
    coeff, reach = [], []
    for m in range(len(tables.mod_names)):
        snr_min = 10 ** ((float(tables.mod_min_osnr[m]) + 2) / 10)
        coeff.append(snr_min * h * f_hz * amp_gain * nf)
        inband = float(tables.mod_inband_xt[m]) + 4
        reach.append(10 ** ((inband - wxt - 4) / 10))
    return np.array(coeff, np.float64), np.array(reach, np.float64)


def make_spec(
    env_id,
    tables: TopologyTables,
    *,
    episode_length: int = 1000,
    load: float = 10.0,
    mean_service_holding_time: Optional[float] = None,
    mean_service_inter_arrival_time: Optional[float] = None,
    num_spectrum_resources: int = 100,
    num_spatial_resources: int = 7,
    j: int = 1,
    bit_rate_selection: Optional[str] = None,
    bit_rates: Optional[Sequence[int]] = None,
    bit_rate_probabilities=None,
    bit_rate_lower_bound: int = 25,
    bit_rate_higher_bound: int = 100,
    node_request_probabilities=None,
    worst_xt: Optional[float] = None,
    seed: Optional[int] = None,
    allow_rejection: bool = False,
    channel_width: float = 12.5,
    auto_reset: bool = False,
    request_source: int = SRC_PHILOX,
    obs_kind: Optional[int] = None,
    modulation_level=None,
    connections_detail=None,
    gsnr=None,
    number_spectrum_channels: int = 80,
    number_spectrum_channels_s_band: int = 108,
    l_band: bool = True,
    s_band: bool = True,
    grooming: bool = False,
    keep_channel_state: bool = False,
    info_stats: bool = False,
    defrag_period=None,
    number_moves=None,
    metric: str = "cut",
    qot: Optional[str] = None,
    qot_adaptive: bool = False,
    qot_thresholds_db=None,
    qot_gsnr_params=None,
) -> EnvSpec:
    kind = ENV_IDS[env_id] if isinstance(env_id, str) else int(env_id)
    n = tables.n_nodes
    if kind == ENV_DEEPRMSA:
        # deeprmsa_env.py:10-32: load = holding / inter-arrival
        holding = 25.0 if mean_service_holding_time is None else mean_service_holding_time
        iat_in = 0.1 if mean_service_inter_arrival_time is None else mean_service_inter_arrival_time
        load = holding / iat_in
    else:
        holding = 10800.0 if mean_service_holding_time is None else mean_service_holding_time
    mean_iat = 1 / float(load / float(holding))  # optical_network_env.py:127-129

    phy_tables = None
    if kind == ENV_PHYRMSA:
        from . import phy as P

        if defrag_period and (number_moves is None or metric not in ("cut", "rss")):
            raise ValueError("defrag_period needs number_moves and metric in ('cut', 'rss') (phy_rmsa_env.py:54-56)")
        if modulation_level is None or connections_detail is None or gsnr is None:
            raise ValueError("PhyRMSA-v0 needs modulation_level, connections_detail and gsnr")
        channels = P.num_channels(number_spectrum_channels, number_spectrum_channels_s_band, l_band, s_band)
        phy_tables = P.build_phy_tables(tables, modulation_level, connections_detail, gsnr, channels)
        num_spectrum_resources = channels  # channels take the place of slots in the engine
        if bit_rates is None:
            bit_rates = P.BIT_RATES
    if bit_rate_selection is None:
        bit_rate_selection = "continuous" if kind == ENV_RMCSA else "discrete"
    assert bit_rate_selection in ("continuous", "discrete")
    br = cum = None
    if bit_rate_selection == "discrete":
        if bit_rates is None:
            bit_rates = (10, 40, 100) if kind == ENV_RMCSA else DEFAULT_BIT_RATES
        if bit_rate_probabilities is None:
            bit_rate_probabilities = [1.0 / len(bit_rates) for _ in bit_rates]  # rmsa_env.py:104-107
        assert len(bit_rates) == len(bit_rate_probabilities)
        br = np.array(bit_rates, np.int32)
        cum = _cum(bit_rate_probabilities)

    if node_request_probabilities is None:
        p = np.full(n, fill_value=1.0 / n)  # optical_network_env.py:106-109
    else:
        p = np.asarray(node_request_probabilities, np.float64)
        assert len(p) == n
    src_cum = _cum(p)
    dst_cum = np.zeros((n, n), np.float64)
    for s in range(n):  # optical_network_env.py:201-203
        q = np.copy(p)
        q[s] = 0.0
        q = q / np.sum(q)
        dst_cum[s] = _cum(q)

    spec = EnvSpec(
        env_kind=kind,
        tables=tables,
        num_slots=int(num_spectrum_resources),
        num_cores=int(num_spatial_resources) if kind == ENV_RMCSA else 1,
        j=int(j) if kind == ENV_DEEPRMSA else 1,
        episode_length=int(episode_length),
        allow_rejection=bool(allow_rejection),
        auto_reset=bool(auto_reset),
        channel_width=float(channel_width),
        mean_holding=float(holding),
        mean_iat=float(mean_iat),
        load=float(load),
        bit_rates=br,
        bit_rate_cum=cum,
        bit_rate_lo=int(bit_rate_lower_bound),
        bit_rate_hi=int(bit_rate_higher_bound),
        node_probs=p,
        src_cum=src_cum,
        dst_cum=dst_cum,
        seed=41 if seed is None else int(seed),  # optical_network_env.py:266-271
        request_source=request_source,
    )
    if kind == ENV_RMCSA:
        wxt = WORST_XT_BY_CORE.get(spec.num_cores) if worst_xt is None else worst_xt
        if wxt is None:
            raise ValueError("worst_xt must be given for this number of cores")
        spec.worst_xt = float(wxt)
        spec.mod_snr_coeff, spec.mod_reach = rmcsa_qot_factors(tables, float(wxt))
    if obs_kind is None:
        obs_kind = OBS_F64 if kind == ENV_DEEPRMSA else OBS_NONE
    spec.obs_kind = obs_kind
    spec.phy = phy_tables
    spec.info_stats = bool(info_stats) and kind in (ENV_RMSA, ENV_DEEPRMSA)
    if kind == ENV_PHYRMSA:
        spec.phy_grooming = 1 if grooming else (2 if keep_channel_state else 0)
        if defrag_period:
            spec.phy_defrag_period, spec.phy_number_moves = int(defrag_period), int(number_moves)
            spec.phy_metric = 0 if metric == "cut" else 1
            if spec.phy_grooming == 0:  # _groom_defragmentation and the release rules read channel_state
                spec.phy_grooming = 2
    if qot:
        if qot != "gn" or kind not in (ENV_RMSA, ENV_RMCSA):
            raise ValueError("qot='gn' (GN-model QoT gate) is defined for RMSA-v0 and RMCSA-v0")
        spec.qot_gate, spec.qot_adaptive = True, bool(qot_adaptive)
        thr = np.full(16, np.inf, np.float64)  # a spectral efficiency without a threshold can never pass
        thr[0] = -np.inf
        src = dict(enumerate(SD_FEC_SNR_DB, start=1)) if qot_thresholds_db is None else dict(qot_thresholds_db)
        for se, v in src.items():
            thr[int(se)] = float(v)
        spec.qot_thr_db = thr
        ses = sorted({int(x) for x in tables.mod_se}, reverse=True)[:8]
        spec.qot_se = np.array(ses, np.int32)
        spec.qot_gsnr = qot_gsnr_params if qot_gsnr_params is not None else gsnr_params(tables, float(channel_width))
    assert spec.num_slots <= 1024 and tables.n_edges <= 128
    return spec


@dataclass
class GsnrParams:
    """Physical inputs of the GN-model GSNR (examples/calculate_osnr.py:9-56).  The reference
    leaves them undefined; they are fixed here once:

    * spans: ``int(length // max_span_km) + 1`` equal spans per link, 0.2 dB/km, NF 4.5 dB
      (examples/create_topology_gn.py:26-29,122-134);
    * ``attenuation_normalized`` alpha [1/m] = a_dB/km / (20 log10 e) / 1e3, consistent with the
      power loss ``exp(-2 alpha L)`` of calculate_osnr.py:25,51;
    * ``noise_figure_normalized`` = 10**(NF_dB/10);  launch power 0 dBm per channel;
    * bandwidth = slots x 12.5 GHz, centre frequency = f0 + (first_slot + slots/2) x 12.5 GHz with
      f0 = 191.35 THz (lower edge of a 4 THz C band).
    """
    launch_power_w: float
    attenuation_per_m: float
    noise_figure_lin: float
    f0_hz: float
    slot_hz: float
    span_count: np.ndarray  # [E] int32
    span_length_km: np.ndarray  # [E] f64


def gsnr_params(tables: TopologyTables, channel_width: float = 12.5, launch_power_dbm: float = 0.0,
                attenuation_db_km: float = 0.2, noise_figure_db: float = 4.5, max_span_km: float = 80.0,
                f0_hz: float = 191.35e12) -> GsnrParams:
    import math

    count = np.array([int(float(x) // max_span_km) + 1 for x in tables.edge_length], np.int32)
    length = np.array([float(x) / int(c) for x, c in zip(tables.edge_length, count)], np.float64)
    return GsnrParams(
        launch_power_w=1e-3 * 10 ** (launch_power_dbm / 10.0),
        attenuation_per_m=attenuation_db_km / (20.0 * math.log10(math.e)) / 1e3,
        noise_figure_lin=10 ** (noise_figure_db / 10.0),
        f0_hz=float(f0_hz),
        slot_hz=float(channel_width) * 1e9,
        span_count=count,
        span_length_km=length,
    )
