"""ctypes binding of the CPU oracle (oracle/liborlg_oracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package never
imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liborlg_oracle.so")

REQUEST_DTYPE = np.dtype(
    [("arrival", "<f8"), ("holding", "<f8"), ("bit_rate", "<i4"), ("src", "<i2"), ("dst", "<i2")]
)
assert REQUEST_DTYPE.itemsize == 24


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "orlg_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"])
    return _LIB_PATH


class OracleConfig(C.Structure):
    _fields_ = [
        ("env_kind", C.c_int32),
        ("n_nodes", C.c_int32),
        ("n_edges", C.c_int32),
        ("k_paths", C.c_int32),
        ("max_hops", C.c_int32),
        ("path_links", C.c_void_p),
        ("path_links_by_id", C.c_void_p),
        ("path_hops", C.c_void_p),
        ("path_length", C.c_void_p),
        ("path_se", C.c_void_p),
        ("n_mods", C.c_int32),
        ("mod_se", C.c_void_p),
        ("mod_max_length", C.c_void_p),
        ("mod_reach", C.c_void_p),
        ("mod_snr_coeff", C.c_void_p),
        ("num_slots", C.c_int32),
        ("num_cores", C.c_int32),
        ("j", C.c_int32),
        ("episode_length", C.c_int32),
        ("allow_rejection", C.c_int32),
        ("auto_reset", C.c_int32),
        ("channel_width", C.c_double),
        ("mean_iat", C.c_double),
        ("mean_holding", C.c_double),
        ("request_source", C.c_int32),
        ("n_bit_rates", C.c_int32),
        ("seed", C.c_uint64),
        ("bit_rates", C.c_void_p),
        ("bit_rate_cum", C.c_void_p),
        ("bit_rate_lo", C.c_int32),
        ("bit_rate_hi", C.c_int32),
        ("src_cum", C.c_void_p),
        ("dst_cum", C.c_void_p),
        ("phy_n_tables", C.c_int32),
        ("phy_grooming", C.c_int32),
        ("phy_table_id", C.c_void_p),
        ("phy_mod_level", C.c_void_p),
        ("phy_gsnr", C.c_void_p),
        ("path_nodes", C.c_void_p),
        ("edge_nodes", C.c_void_p),
        ("info_stats", C.c_int32),
        ("info_pad", C.c_int32),
        ("qot_gate", C.c_int32),
        ("qot_adaptive", C.c_int32),
        ("qot_n_se", C.c_int32),
        ("qot_pad", C.c_int32),
        ("qot_se", C.c_int32 * 8),
        ("qot_thr_db", C.c_double * 16),
        ("qot_gsnr", C.c_void_p),
        ("phy_defrag_period", C.c_int32),
        ("phy_number_moves", C.c_int32),
        ("phy_metric", C.c_int32),
        ("phy_pad", C.c_int32),
    ]


class OracleGsnrParams(C.Structure):
    _fields_ = [("launch_power_w", C.c_double), ("attenuation_per_m", C.c_double), ("noise_figure_lin", C.c_double),
                ("f0_hz", C.c_double), ("slot_hz", C.c_double), ("span_count", C.c_void_p),
                ("span_length_km", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.oracle_create.restype = C.c_void_p
        L.oracle_create.argtypes = [C.POINTER(OracleConfig), C.c_int64]
        L.oracle_destroy.argtypes = [C.c_void_p]
        L.oracle_set_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.oracle_reset.argtypes = [C.c_void_p, C.c_int]
        L.oracle_action_dim.argtypes = [C.c_void_p]
        L.oracle_obs_dim.argtypes = [C.c_void_p]
        L.oracle_step.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.oracle_observation.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_matrix_observation.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_heuristic.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.oracle_counters.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_state.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_request_now.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_rollout_many.restype = C.c_int64
        L.oracle_rollout_many.argtypes = [
            C.POINTER(OracleConfig), C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int,
            C.c_void_p, C.c_void_p, C.c_void_p,
        ]
        L.oracle_pool_create.restype = C.c_void_p
        L.oracle_pool_create.argtypes = [C.POINTER(OracleConfig), C.c_int64, C.c_int64]
        L.oracle_pool_destroy.argtypes = [C.c_void_p]
        L.oracle_pool_rollout.restype = C.c_int64
        L.oracle_pool_rollout.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.oracle_pool_env.restype = C.c_void_p
        L.oracle_pool_env.argtypes = [C.c_void_p, C.c_int64]
        L.oracle_pool_reward_sum.restype = C.c_int64
        L.oracle_pool_reward_sum.argtypes = [C.c_void_p]
        L.oracle_pool_set_gsnr.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_pool_gsnr_sum.restype = C.c_double
        L.oracle_pool_gsnr_sum.argtypes = [C.c_void_p]
        L.oracle_phy_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_phy_moves.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_phy_channel_state.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_phy_channel_state.restype = C.c_int64
        L.oracle_info.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.oracle_info.restype = C.c_int
        L.oracle_gsnr.argtypes = [C.c_void_p, C.POINTER(OracleGsnrParams), C.c_void_p, C.c_void_p]
        L.oracle_qot_last.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_services.restype = C.c_int64
        L.oracle_services.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_dlog.restype = C.c_double
        L.oracle_dlog.argtypes = [C.c_double]
        L.oracle_philox4x32.argtypes = [C.c_uint32] * 6 + [C.c_void_p]
        L.oracle_philox_request.argtypes = [C.POINTER(OracleConfig), C.c_int64, C.c_int64, C.c_double, C.c_void_p]
        _lib = L
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data


class OracleSpec:
    """Keeps the numpy arrays alive behind an OracleConfig."""

    def __init__(self, spec):
        t = spec.tables
        k = self.keep = {}
        k["path_links"] = np.ascontiguousarray(t.path_links, np.int16)
        k["path_links_by_id"] = (
            None if t.path_links_by_id is None else np.ascontiguousarray(t.path_links_by_id, np.int16)
        )
        k["path_hops"] = np.ascontiguousarray(t.path_hops, np.int32)
        k["path_length"] = np.ascontiguousarray(t.path_length, np.float64)
        k["path_se"] = np.ascontiguousarray(t.path_se, np.int32)
        k["mod_se"] = np.ascontiguousarray(t.mod_se, np.int32)
        k["mod_max_length"] = np.ascontiguousarray(t.mod_max_length, np.float64)
        k["mod_reach"] = None if spec.mod_reach is None else np.ascontiguousarray(spec.mod_reach, np.float64)
        k["mod_snr_coeff"] = (
            None if spec.mod_snr_coeff is None else np.ascontiguousarray(spec.mod_snr_coeff, np.float64)
        )
        k["bit_rates"] = None if spec.bit_rates is None else np.ascontiguousarray(spec.bit_rates, np.int32)
        k["bit_rate_cum"] = (
            None if spec.bit_rate_cum is None else np.ascontiguousarray(spec.bit_rate_cum, np.float64)
        )
        k["src_cum"] = np.ascontiguousarray(spec.src_cum, np.float64)
        k["dst_cum"] = np.ascontiguousarray(spec.dst_cum, np.float64)
        k["path_nodes"] = np.ascontiguousarray(t.path_nodes, np.int16)
        k["edge_nodes"] = np.ascontiguousarray(t.edge_nodes, np.int32)
        ph = getattr(spec, "phy", None)
        k["phy_table_id"] = None if ph is None else np.ascontiguousarray(ph.table_id, np.int32)
        k["phy_mod_level"] = None if ph is None else np.ascontiguousarray(ph.mod_level, np.uint8)
        k["phy_gsnr"] = None if ph is None else np.ascontiguousarray(ph.gsnr, np.float64)
        c = self.cfg = OracleConfig()
        c.env_kind = spec.env_kind
        c.n_nodes, c.n_edges, c.k_paths, c.max_hops = t.n_nodes, t.n_edges, t.k_paths, t.max_hops
        for name in ("path_links", "path_links_by_id", "path_hops", "path_length", "path_se", "mod_se",
                     "mod_max_length", "mod_reach", "mod_snr_coeff", "bit_rates", "bit_rate_cum",
                     "src_cum", "dst_cum", "path_nodes", "edge_nodes", "phy_table_id", "phy_mod_level",
                     "phy_gsnr"):
            setattr(c, name, _ptr(k[name]))
        c.n_mods = len(t.mod_names)
        c.phy_n_tables = 0 if ph is None else ph.n_tables
        c.phy_grooming = int(getattr(spec, "phy_grooming", 0))
        c.info_stats = int(bool(getattr(spec, "info_stats", False)))
        c.num_slots, c.num_cores, c.j = spec.num_slots, spec.num_cores, spec.j
        c.episode_length = spec.episode_length
        c.allow_rejection = int(spec.allow_rejection)
        c.auto_reset = int(spec.auto_reset)
        c.channel_width, c.mean_iat, c.mean_holding = spec.channel_width, spec.mean_iat, spec.mean_holding
        c.request_source = spec.request_source
        c.n_bit_rates = 0 if spec.bit_rates is None else len(spec.bit_rates)
        c.seed = spec.seed & 0xFFFFFFFFFFFFFFFF
        c.bit_rate_lo, c.bit_rate_hi = spec.bit_rate_lo, spec.bit_rate_hi
        c.phy_defrag_period = int(getattr(spec, "phy_defrag_period", 0) or 0)
        c.phy_number_moves = int(getattr(spec, "phy_number_moves", 0) or 0)
        c.phy_metric = int(getattr(spec, "phy_metric", 0) or 0)
        if getattr(spec, "qot_gate", False):
            gp = spec.qot_gsnr
            k["qot_span_count"] = np.ascontiguousarray(gp.span_count, np.int32)
            k["qot_span_km"] = np.ascontiguousarray(gp.span_length_km, np.float64)
            self.qot_gp = OracleGsnrParams(gp.launch_power_w, gp.attenuation_per_m, gp.noise_figure_lin, gp.f0_hz,
                                           gp.slot_hz, k["qot_span_count"].ctypes.data, k["qot_span_km"].ctypes.data)
            c.qot_gate, c.qot_adaptive = 1, int(spec.qot_adaptive)
            c.qot_n_se = len(spec.qot_se)
            for i, v in enumerate(spec.qot_se):
                c.qot_se[i] = int(v)
            for i, v in enumerate(spec.qot_thr_db):
                c.qot_thr_db[i] = float(v)
            c.qot_gsnr = C.cast(C.pointer(self.qot_gp), C.c_void_p)
        self.spec = spec


class OracleEnv:
    """One env of the CPU oracle."""

    def __init__(self, spec, global_env_id: int = 0, trace=None):
        self.L = lib()
        self.ospec = spec if isinstance(spec, OracleSpec) else OracleSpec(spec)
        self.spec = self.ospec.spec
        self.h = self.L.oracle_create(C.byref(self.ospec.cfg), global_env_id)
        self.action_dim = self.L.oracle_action_dim(self.h)
        self.obs_dim = self.L.oracle_obs_dim(self.h)
        self._trace = None
        if trace is not None:
            self.set_trace(trace)

    def set_trace(self, trace):
        self._trace = np.ascontiguousarray(trace, REQUEST_DTYPE)
        self.L.oracle_set_trace(self.h, self._trace.ctypes.data, len(self._trace), 1)

    def reset(self, full=True):
        self.L.oracle_reset(self.h, int(full))

    def step(self, action, want_obs=False):
        a = np.ascontiguousarray(np.atleast_1d(action), np.int32)
        assert a.size == self.action_dim
        r = np.zeros(1, np.int32)
        d = np.zeros(1, np.uint8)
        alloc = np.zeros(4, np.int32)
        obs = np.zeros(self.obs_dim, np.float64) if want_obs else None
        self.L.oracle_step(self.h, a.ctypes.data, r.ctypes.data, d.ctypes.data, _ptr(obs), alloc.ctypes.data)
        return obs, int(r[0]), bool(d[0]), alloc

    def observation(self):
        obs = np.zeros(self.obs_dim, np.float64)
        self.L.oracle_observation(self.h, obs.ctypes.data)
        return obs

    def matrix_observation(self):
        s = self.spec
        n = 2 * s.tables.n_nodes + s.num_cores * s.tables.n_edges * s.num_slots
        obs = np.zeros(n, np.uint8)
        self.L.oracle_matrix_observation(self.h, obs.ctypes.data)
        return obs

    def info(self):
        """info dict values of the last step (see oracle_info in orlg_oracle.h)."""
        o = np.zeros(64, np.float64)
        n = int(self.L.oracle_info(self.h, o.ctypes.data, 64))
        return o[:n]

    def phy_channel_state(self, max_entries=4096):
        o = np.zeros((max_entries, 6), np.int32)
        n = int(self.L.oracle_phy_channel_state(self.h, o.ctypes.data, max_entries))
        return o[: min(n, max_entries)]

    def phy_moves(self):
        """(counted_moves, counted_moves_groom, counted_defrag_cycles) of the current episode."""
        o = np.zeros(3, np.int64)
        self.L.oracle_phy_moves(self.h, o.ctypes.data)
        return o

    def phy_stats(self):
        o = np.zeros(8, np.float64)
        self.L.oracle_phy_stats(self.h, o.ctypes.data)
        return o

    def heuristic(self, hid):
        a = np.zeros(16, np.int32)
        self.L.oracle_heuristic(self.h, hid, a.ctypes.data)
        return a[: self.action_dim].copy()

    def counters(self):
        o = np.zeros(12, np.int64)
        self.L.oracle_counters(self.h, o.ctypes.data)
        return o

    def state(self):
        s = self.spec
        o = np.zeros((s.num_cores, s.tables.n_edges, s.num_slots), np.uint8)
        self.L.oracle_state(self.h, o.ctypes.data)
        return o

    def request(self):
        o = np.zeros(1, REQUEST_DTYPE)
        self.L.oracle_request_now(self.h, o.ctypes.data)
        return o[0]

    def gsnr(self, gp, query):
        """gp: spec.GsnrParams; query: 4 ints.  Returns (candidate dB, min neighbour dB)."""
        cnt = np.ascontiguousarray(gp.span_count, np.int32)
        ln = np.ascontiguousarray(gp.span_length_km, np.float64)
        c = OracleGsnrParams(gp.launch_power_w, gp.attenuation_per_m, gp.noise_figure_lin, gp.f0_hz, gp.slot_hz,
                             cnt.ctypes.data, ln.ctypes.data)
        q = np.ascontiguousarray(query, np.int32)
        out = np.zeros(2, np.float64)
        self.L.oracle_gsnr(self.h, C.byref(c), q.ctypes.data, out.ctypes.data)
        return out

    def qot_last(self):
        """(candidate GSNR dB, lowest neighbour GSNR dB) of the last candidate the QoT gate evaluated."""
        o = np.zeros(2, np.float64)
        self.L.oracle_qot_last(self.h, o.ctypes.data)
        return o

    def services(self):
        """Live services: rows (pair, path, core, initial_slot, number_slots, se)."""
        n = int(self.L.oracle_services(self.h, None, 0))
        out = np.zeros((max(n, 1), 6), np.int32)
        self.L.oracle_services(self.h, out.ctypes.data, n)
        return out[:n]

    def close(self):
        if self.h:
            self.L.oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def rollout_many(spec, env0, n_envs, nsteps, heuristic_id, threads, want_state=False):
    """n_envs independent philox-driven envs; returns (steps, counters, state, reward_sum)."""
    L = lib()
    os_ = spec if isinstance(spec, OracleSpec) else OracleSpec(spec)
    s = os_.spec
    counters = np.zeros((n_envs, 12), np.int64)
    rsum = np.zeros(n_envs, np.int64)
    state = np.zeros((n_envs, s.num_cores, s.tables.n_edges, s.num_slots), np.uint8) if want_state else None
    steps = L.oracle_rollout_many(C.byref(os_.cfg), env0, n_envs, nsteps, heuristic_id, threads,
                                  counters.ctypes.data, _ptr(state), rsum.ctypes.data)
    return steps, counters, state, rsum


class OraclePool:
    """n_envs persistent oracle envs stepped with a heuristic on host threads."""

    def __init__(self, spec, env0: int, n_envs: int):
        self.L = lib()
        self.ospec = spec if isinstance(spec, OracleSpec) else OracleSpec(spec)
        self.n_envs = int(n_envs)
        self.h = self.L.oracle_pool_create(C.byref(self.ospec.cfg), env0, n_envs)

    def rollout(self, nsteps: int, heuristic_id: int, threads: int) -> int:
        return int(self.L.oracle_pool_rollout(self.h, nsteps, heuristic_id, threads))

    def set_gsnr(self, gp):
        """Evaluate calculate_osnr for every decision of the following rollouts (gp: spec.GsnrParams or None)."""
        if gp is None:
            self._gs = None
            self.L.oracle_pool_set_gsnr(self.h, None)
            return
        cnt = np.ascontiguousarray(gp.span_count, np.int32)
        ln = np.ascontiguousarray(gp.span_length_km, np.float64)
        c = OracleGsnrParams(gp.launch_power_w, gp.attenuation_per_m, gp.noise_figure_lin, gp.f0_hz, gp.slot_hz,
                             cnt.ctypes.data, ln.ctypes.data)
        self._gs = (cnt, ln, c)  # keep the buffers alive
        self.L.oracle_pool_set_gsnr(self.h, C.byref(c))

    def gsnr_sum(self) -> float:
        return float(self.L.oracle_pool_gsnr_sum(self.h))

    def counters(self):
        out = np.zeros((self.n_envs, 12), np.int64)
        for i in range(self.n_envs):
            self.L.oracle_counters(self.L.oracle_pool_env(self.h, i), out[i].ctypes.data)
        return out

    def reward_sum(self) -> int:
        return int(self.L.oracle_pool_reward_sum(self.h))

    def close(self):
        if self.h:
            self.L.oracle_pool_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
