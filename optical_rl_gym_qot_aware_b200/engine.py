"""ctypes binding of liborlg.so (include/orlg.h) with torch tensors as device memory.

PyTorch is plumbing here: it owns the caller-side device buffers (actions, obs,
reward, done, traces) and the CUDA stream; every computation happens in the
hand-written kernels behind the C-ABI.  There is no CPU fallback: if the
extension is missing or no CUDA device is present, constructing an Engine fails.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import spec as S

_HERE = os.path.dirname(os.path.abspath(__file__))
# ORLG_LIB selects another build of the same library (A/B runs of kernel variants)
LIB_PATH = os.environ.get("ORLG_LIB") or os.path.join(_HERE, "liborlg.so")

REQUEST_DTYPE = np.dtype(
    [("arrival", "<f8"), ("holding", "<f8"), ("bit_rate", "<i4"), ("src", "<i2"), ("dst", "<i2")]
)
NUM_COUNTERS = 12
COUNTER_NAMES = (
    "services_processed", "services_accepted", "episode_services_processed", "episode_services_accepted",
    "bit_rate_requested", "bit_rate_provisioned", "episode_bit_rate_requested",
    "episode_bit_rate_provisioned", "requests_drawn", "queue_len", "fault", "episodes",
)

EXPORTS = (
    "orlg_abi_version", "orlg_build_info", "orlg_create", "orlg_destroy", "orlg_last_error",
    "orlg_action_dim", "orlg_obs_dim", "orlg_state_words", "orlg_hbm_bytes", "orlg_load_trace",
    "orlg_reset", "orlg_step", "orlg_step_host", "orlg_set_obs_layout", "orlg_observation",
    "orlg_heuristic", "orlg_rollout", "orlg_get_counters", "orlg_export_state", "orlg_get_requests",
    "orlg_launch_count", "orlg_gsnr_configure", "orlg_gsnr", "orlg_gsnr_for_actions", "orlg_phy_get_stats", "orlg_path_first_fit", "orlg_set_next_action_hint", "orlg_step_host_async", "orlg_step_host_wait", "orlg_set_host_wait",
    "orlg_phy_get_channel_state", "orlg_info_dim", "orlg_set_info_output",
    "orlg_step_auto", "orlg_qot_get", "orlg_qot_check", "orlg_step_auto_noalloc",
    "orlg_export_services", "orlg_timed_steps", "orlg_phy_get_moves",
    "orlg_set_counter_snapshot",
)


class OrlgConfig(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("env_kind", C.c_int32), ("device", C.c_int32), ("num_envs", C.c_int32),
        ("global_env_offset", C.c_int64),
        ("n_nodes", C.c_int32), ("n_edges", C.c_int32), ("k_paths", C.c_int32), ("max_hops", C.c_int32),
        ("path_links", C.c_void_p), ("path_links_by_id", C.c_void_p), ("path_hops", C.c_void_p),
        ("path_length", C.c_void_p), ("path_se", C.c_void_p),
        ("n_mods", C.c_int32),
        ("mod_se", C.c_void_p), ("mod_max_length", C.c_void_p), ("mod_reach", C.c_void_p),
        ("mod_snr_coeff", C.c_void_p),
        ("num_slots", C.c_int32), ("num_cores", C.c_int32), ("j", C.c_int32), ("episode_length", C.c_int32),
        ("allow_rejection", C.c_int32), ("auto_reset", C.c_int32),
        ("channel_width", C.c_double), ("mean_iat", C.c_double), ("mean_holding", C.c_double),
        ("request_source", C.c_int32), ("n_bit_rates", C.c_int32), ("seed", C.c_uint64),
        ("bit_rates", C.c_void_p), ("bit_rate_cum", C.c_void_p),
        ("bit_rate_lo", C.c_int32), ("bit_rate_hi", C.c_int32),
        ("src_cum", C.c_void_p), ("dst_cum", C.c_void_p),
        ("heap_capacity", C.c_int32), ("obs_kind", C.c_int32), ("phy_grooming", C.c_int32), ("info_stats", C.c_int32), ("reserved", C.c_int32 * 4),
        ("phy_n_tables", C.c_int32), ("phy_max_terms", C.c_int32),
        ("phy_table_id", C.c_void_p), ("phy_mod_level", C.c_void_p), ("phy_gsnr", C.c_void_p),
        ("phy_cut_links", C.c_void_p), ("phy_cut_count", C.c_void_p),
        ("qot_gate", C.c_int32), ("qot_adaptive", C.c_int32), ("qot_n_se", C.c_int32), ("qot_pad", C.c_int32),
        ("qot_se", C.c_int32 * 8), ("qot_thr_db", C.c_double * 16), ("qot_gsnr", C.c_void_p),
        ("phy_defrag_period", C.c_int32), ("phy_number_moves", C.c_int32), ("phy_metric", C.c_int32), ("phy_pad", C.c_int32),
    ]


class OrlgGsnrParams(C.Structure):
    _fields_ = [("launch_power_w", C.c_double), ("attenuation_per_m", C.c_double), ("noise_figure_lin", C.c_double),
                ("f0_hz", C.c_double), ("slot_hz", C.c_double), ("span_count", C.c_void_p),
                ("span_length_km", C.c_void_p)]


_lib = None


def load_library():
    """Load liborlg.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`"
            )
        L = C.CDLL(LIB_PATH)
        vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
        L.orlg_abi_version.restype = i32
        L.orlg_build_info.restype = C.c_char_p
        L.orlg_create.argtypes = [C.POINTER(OrlgConfig), C.POINTER(vp)]
        L.orlg_destroy.argtypes = [vp]
        L.orlg_last_error.restype = C.c_char_p
        L.orlg_last_error.argtypes = [vp]
        for name in ("orlg_action_dim", "orlg_obs_dim"):
            getattr(L, name).argtypes = [vp]
        for name in ("orlg_state_words", "orlg_hbm_bytes", "orlg_launch_count"):
            getattr(L, name).argtypes = [vp]
            getattr(L, name).restype = i64
        L.orlg_load_trace.argtypes = [vp, vp, i64]
        L.orlg_reset.argtypes = [vp, i32, vp]
        L.orlg_step.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.orlg_step_host.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.orlg_step_host_async.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.orlg_step_host_wait.argtypes = [vp, vp]
        L.orlg_set_host_wait.argtypes = [vp, i32]
        L.orlg_set_obs_layout.argtypes = [vp, i64, i64]
        L.orlg_observation.argtypes = [vp, i32, vp, vp]
        L.orlg_heuristic.argtypes = [vp, i32, vp, vp]
        L.orlg_rollout.argtypes = [vp, i32, i32, vp, vp]
        L.orlg_get_counters.argtypes = [vp, vp, vp]
        L.orlg_export_state.argtypes = [vp, vp, vp]
        L.orlg_get_requests.argtypes = [vp, vp, vp]
        L.orlg_gsnr_configure.argtypes = [vp, C.POINTER(OrlgGsnrParams)]
        L.orlg_gsnr.argtypes = [vp, vp, vp, vp]
        L.orlg_gsnr_for_actions.argtypes = [vp, vp, vp, vp]
        L.orlg_phy_get_stats.argtypes = [vp, vp, vp]
        L.orlg_phy_get_channel_state.argtypes = [vp, vp, i32, vp, vp]
        L.orlg_info_dim.argtypes = [vp]
        L.orlg_set_info_output.argtypes = [vp, vp]
        L.orlg_path_first_fit.argtypes = [vp, vp, vp, vp]
        L.orlg_set_next_action_hint.argtypes = [vp, i32, vp]
        L.orlg_step_auto.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp]
        L.orlg_step_auto_noalloc.argtypes = [vp, i32, vp, vp, vp, vp, vp]
        L.orlg_export_services.argtypes = [vp, vp, i32, vp, vp]
        L.orlg_timed_steps.argtypes = [vp, i32, i32, vp, vp, vp, vp, i32, vp, C.c_size_t, i32, vp, vp, vp]
        L.orlg_phy_get_moves.argtypes = [vp, vp, vp]
        L.orlg_set_counter_snapshot.argtypes = [vp, vp]
        L.orlg_qot_get.argtypes = [vp, vp, vp]
        L.orlg_qot_check.argtypes = [vp, vp, vp]
        _lib = L
    return _lib


def _np_ptr(a):
    return None if a is None else a.ctypes.data


class Engine:
    """One engine = one GPU's shard of envs (device-resident state)."""

    def __init__(self, spec: S.EnvSpec, num_envs: int, device: int = 0, global_env_offset: int = 0,
                 heap_capacity: int = 0, obs_layout: str = "feature_major"):
        import torch

        if not torch.cuda.is_available():
            raise RuntimeError("optical_rl_gym_qot_aware_b200 needs a CUDA device (no CPU fallback)")
        self.torch = torch
        self.L = load_library()
        self.spec = spec
        self.num_envs = int(num_envs)
        self.device = torch.device("cuda", device)
        t = spec.tables
        k = self._keep = {}
        k["path_links"] = np.ascontiguousarray(t.path_links, np.int16)
        by_id = t.path_links_by_id
        k["path_links_by_id"] = None if by_id is None else np.ascontiguousarray(by_id, np.int16)
        k["path_hops"] = np.ascontiguousarray(t.path_hops, np.int32)
        k["path_length"] = np.ascontiguousarray(t.path_length, np.float64)
        k["path_se"] = np.ascontiguousarray(t.path_se, np.int32)
        k["mod_se"] = np.ascontiguousarray(t.mod_se, np.int32)
        k["mod_max_length"] = np.ascontiguousarray(t.mod_max_length, np.float64)
        for name in ("mod_reach", "mod_snr_coeff"):
            v = getattr(spec, name)
            k[name] = None if v is None else np.ascontiguousarray(v, np.float64)
        k["bit_rates"] = None if spec.bit_rates is None else np.ascontiguousarray(spec.bit_rates, np.int32)
        k["bit_rate_cum"] = None if spec.bit_rate_cum is None else np.ascontiguousarray(spec.bit_rate_cum, np.float64)
        k["src_cum"] = np.ascontiguousarray(spec.src_cum, np.float64)
        k["dst_cum"] = np.ascontiguousarray(spec.dst_cum, np.float64)
        ph = getattr(spec, "phy", None)
        if ph is not None:
            k["phy_table_id"] = np.ascontiguousarray(ph.table_id, np.int32)
            k["phy_mod_level"] = np.ascontiguousarray(ph.mod_level, np.uint8)
            k["phy_gsnr"] = np.ascontiguousarray(ph.gsnr, np.float64)
            k["phy_cut_links"] = np.ascontiguousarray(ph.cut_links, np.uint8)
            k["phy_cut_count"] = np.ascontiguousarray(ph.cut_count, np.int32)
        c = OrlgConfig()
        if ph is not None:
            c.phy_n_tables, c.phy_max_terms = ph.n_tables, ph.max_terms
            c.phy_grooming = int(getattr(spec, "phy_grooming", 0))
            c.phy_defrag_period = int(getattr(spec, "phy_defrag_period", 0) or 0)
            c.phy_number_moves = int(getattr(spec, "phy_number_moves", 0) or 0)
            c.phy_metric = int(getattr(spec, "phy_metric", 0) or 0)
        c.abi_version = self.L.orlg_abi_version()
        c.env_kind, c.device, c.num_envs = spec.env_kind, device, self.num_envs
        c.global_env_offset = global_env_offset
        c.n_nodes, c.n_edges, c.k_paths, c.max_hops = t.n_nodes, t.n_edges, t.k_paths, t.max_hops
        for name in k:
            setattr(c, name, _np_ptr(k[name]))
        c.n_mods = len(t.mod_names)
        c.num_slots, c.num_cores, c.j = spec.num_slots, spec.num_cores, spec.j
        c.episode_length = spec.episode_length
        c.allow_rejection, c.auto_reset = int(spec.allow_rejection), int(spec.auto_reset)
        c.channel_width, c.mean_iat, c.mean_holding = spec.channel_width, spec.mean_iat, spec.mean_holding
        c.request_source = spec.request_source
        c.n_bit_rates = 0 if spec.bit_rates is None else len(spec.bit_rates)
        c.seed = spec.seed & 0xFFFFFFFFFFFFFFFF
        c.bit_rate_lo, c.bit_rate_hi = spec.bit_rate_lo, spec.bit_rate_hi
        c.heap_capacity = heap_capacity
        c.obs_kind = spec.obs_kind
        c.info_stats = int(bool(getattr(spec, "info_stats", False)))
        if getattr(spec, "qot_gate", False):  # GN-model QoT gate (include/orlg.h "QoT gate")
            gp = spec.qot_gsnr
            k["qot_span_count"] = np.ascontiguousarray(gp.span_count, np.int32)
            k["qot_span_km"] = np.ascontiguousarray(gp.span_length_km, np.float64)
            self._qot_gp = OrlgGsnrParams(gp.launch_power_w, gp.attenuation_per_m, gp.noise_figure_lin, gp.f0_hz,
                                          gp.slot_hz, k["qot_span_count"].ctypes.data, k["qot_span_km"].ctypes.data)
            c.qot_gate, c.qot_adaptive, c.qot_n_se = 1, int(spec.qot_adaptive), len(spec.qot_se)
            for i, v in enumerate(spec.qot_se):
                c.qot_se[i] = int(v)
            for i, v in enumerate(spec.qot_thr_db):
                c.qot_thr_db[i] = float(v)
            c.qot_gsnr = C.cast(C.pointer(self._qot_gp), C.c_void_p)
            self.gsnr_params = gp
        h = C.c_void_p()
        rc = self.L.orlg_create(C.byref(c), C.byref(h))
        if rc != 0:
            raise RuntimeError(f"orlg_create failed ({rc}): {self.L.orlg_last_error(None).decode()}")
        self.h = h
        self.action_dim = self.L.orlg_action_dim(h)
        self.obs_dim = self.L.orlg_obs_dim(h)
        self.obs_layout = obs_layout
        if obs_layout == "feature_major":
            self._ck(self.L.orlg_set_obs_layout(h, 1, self.num_envs))
        self._trace = None
        self._obs_dtype = {S.OBS_F64: torch.float64, S.OBS_F32: torch.float32, S.OBS_MATRIX_U8: torch.uint8,
                           S.OBS_PHY_INFO: torch.float64, S.OBS_PHY_CUTS: torch.float64}.get(spec.obs_kind)

    # -- helpers -----------------------------------------------------------
    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(f"orlg call failed ({rc}): {self.L.orlg_last_error(self.h).decode()}")

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    @staticmethod
    def _p(t):
        return None if t is None else C.c_void_p(t.data_ptr())

    def new_obs(self, kind=None):
        """Device buffer for observations, returned as an [num_envs, obs_dim] view."""
        torch = self.torch
        dt = self._obs_dtype if kind is None else {S.OBS_F64: torch.float64, S.OBS_F32: torch.float32,
                                                   S.OBS_MATRIX_U8: torch.uint8, S.OBS_PHY_INFO: torch.float64,
                                                   S.OBS_PHY_CUTS: torch.float64}[kind]
        dim = self.obs_dim if kind is None else self._dim_of(kind)
        if self.obs_layout == "feature_major":
            return torch.empty((dim, self.num_envs), dtype=dt, device=self.device).t()
        return torch.empty((self.num_envs, dim), dtype=dt, device=self.device)

    def _dim_of(self, kind):
        s, t = self.spec, self.spec.tables
        if kind == S.OBS_MATRIX_U8:
            return 2 * t.n_nodes + s.num_cores * t.n_edges * s.num_slots
        if kind in (S.OBS_PHY_INFO, S.OBS_PHY_CUTS):
            return 2 if kind == S.OBS_PHY_INFO else 1
        return 1 + 2 * t.n_nodes + (2 * s.j + 3) * t.k_paths

    def _obs_base(self, obs):
        # the buffer start is the same for both layouts (the view is a transpose)
        return None if obs is None else C.c_void_p(obs.data_ptr())

    # -- ABI calls -----------------------------------------------------------
    def load_trace(self, trace):
        """trace: numpy structured array [n_requests, num_envs] of REQUEST_DTYPE."""
        torch = self.torch
        tr = np.ascontiguousarray(trace, REQUEST_DTYPE)
        assert tr.ndim == 2 and tr.shape[1] == self.num_envs
        self._trace = torch.from_numpy(tr.view(np.uint8).reshape(tr.shape[0], -1)).to(self.device)
        self._ck(self.L.orlg_load_trace(self.h, self._p(self._trace), tr.shape[0]))

    def reset(self, full=True):
        self._ck(self.L.orlg_reset(self.h, int(full), self._stream()))

    def step(self, actions, reward=None, done=None, obs=None, alloc=None):
        assert actions.dtype == self.torch.int32 and actions.is_cuda and actions.is_contiguous()
        self._ck(self.L.orlg_step(self.h, self._p(actions), self._p(reward), self._p(done),
                                  self._obs_base(obs), self._p(alloc), self._stream()))

    def step_host(self, actions, reward=None, done=None, obs=None, alloc=None):
        """Host (pinned) tensors in and out; copies and sync happen inside the C call."""
        self._ck(self.L.orlg_step_host(self.h, self._p(actions), self._p(reward), self._p(done),
                                       self._p(obs), self._p(alloc), self._stream()))

    def step_host_async(self, actions, reward=None, done=None, obs=None, alloc=None, stream=None):
        """Enqueue H2D actions, one step, D2H outputs on ``stream`` (a torch.cuda.Stream; default: the
        current stream) without waiting; pair with :meth:`step_host_wait`."""
        st = self._stream() if stream is None else C.c_void_p(stream.cuda_stream)
        self._ck(self.L.orlg_step_host_async(self.h, self._p(actions), self._p(reward), self._p(done),
                                             self._p(obs), self._p(alloc), st))

    def step_host_wait(self, stream=None):
        st = self._stream() if stream is None else C.c_void_p(stream.cuda_stream)
        self._ck(self.L.orlg_step_host_wait(self.h, st))

    def observation(self, obs, kind=None):
        kind = self.spec.obs_kind if kind is None else kind
        self._ck(self.L.orlg_observation(self.h, kind, self._obs_base(obs), self._stream()))

    def heuristic(self, heuristic_id, actions):
        self._ck(self.L.orlg_heuristic(self.h, heuristic_id, self._p(actions), self._stream()))

    def set_next_action_hint(self, heuristic_id, actions):
        """Every later step() also writes the heuristic's action for the new pending request into
        ``actions`` (int32 [num_envs, action_dim] device tensor, kept alive here); None turns it off."""
        self._hint = actions
        self._ck(self.L.orlg_set_next_action_hint(self.h, int(heuristic_id), self._p(actions)))

    def path_first_fit(self, paths, actions):
        """PathOnlyFirstFitAction (rmsa_env.py:974-1008): int32 [num_envs] paths -> [num_envs, 2] actions."""
        assert paths.dtype == self.torch.int32 and paths.is_cuda and paths.is_contiguous()
        self._ck(self.L.orlg_path_first_fit(self.h, self._p(paths), self._p(actions), self._stream()))

    def rollout(self, heuristic_id, nsteps, reward_sum=None):
        self._ck(self.L.orlg_rollout(self.h, heuristic_id, int(nsteps), self._p(reward_sum), self._stream()))

    def step_auto(self, heuristic_id, actions_out=None, reward=None, done=None, obs=None, alloc=None):
        """One step with the device heuristic as the policy (orlg_step_auto)."""
        self._ck(self.L.orlg_step_auto(self.h, int(heuristic_id), self._p(actions_out), self._p(reward), self._p(done),
                                       self._obs_base(obs), self._p(alloc), self._stream()))

    def step_auto_noalloc(self, heuristic_id, reward=None, done=None, obs=None, alloc=None):
        """orlg_step_auto_noalloc: as step_auto(heuristic_id, None, ...), with the action scratch in a static device
        array of the library (num_envs * action_dim <= 262144); the call allocates nothing."""
        self._ck(self.L.orlg_step_auto_noalloc(self.h, int(heuristic_id), self._p(reward), self._p(done),
                                               self._obs_base(obs), self._p(alloc), self._stream()))

    def services(self, max_rows: int = 512):
        """Running services of every env: (int32 [num_envs, max_rows, 6], int32 [num_envs] counts); rows are
        (src_id * n_nodes + dst_id, path, core, first slot, slots, spectral efficiency) in release-queue order."""
        torch = self.torch
        out = torch.zeros((self.num_envs, max_rows, 6), dtype=torch.int32, device=self.device)
        cnt = torch.zeros(self.num_envs, dtype=torch.int32, device=self.device)
        self._ck(self.L.orlg_export_services(self.h, self._p(out), max_rows, self._p(cnt), self._stream()))
        return out, cnt

    def timed_steps(self, nsteps, actions=None, reward=None, done=None, obs=None, heuristic_id=None, flush=None,
                    use_graph=True):
        """orlg_timed_steps: (ms per step [nsteps], ms of the step's first launch [nsteps]) as numpy f32.
        heuristic_id None: orlg_step(actions) as configured; else orlg_step_auto(heuristic_id, actions)."""
        ms = np.zeros(nsteps, np.float32)
        first = np.zeros(nsteps, np.float32)
        mode = 0 if heuristic_id is None else 1
        fb = 0 if flush is None else flush.numel() * flush.element_size()
        self._ck(self.L.orlg_timed_steps(self.h, mode, int(heuristic_id or 0), self._p(actions), self._p(reward),
                                         self._p(done), self._obs_base(obs), int(nsteps), self._p(flush), fb,
                                         int(bool(use_graph)), ms.ctypes.data, first.ctypes.data, self._stream()))
        return ms, first

    def qot_last(self, out=None):
        """QoT gate: f64 [num_envs, 2] (candidate GSNR dB, lowest neighbour GSNR dB) of the last candidates evaluated."""
        if out is None:
            out = self.torch.empty((self.num_envs, 2), dtype=self.torch.float64, device=self.device)
        self._ck(self.L.orlg_qot_get(self.h, self._p(out), self._stream()))
        return out

    def qot_check(self):
        """QoT gate self-check: int64 [num_envs] rows whose running sums differ from a from-scratch evaluation."""
        out = self.torch.empty(self.num_envs, dtype=self.torch.int64, device=self.device)
        self._ck(self.L.orlg_qot_check(self.h, self._p(out), self._stream()))
        return out

    def counters(self):
        out = self.torch.empty((self.num_envs, NUM_COUNTERS), dtype=self.torch.int64, device=self.device)
        self._ck(self.L.orlg_get_counters(self.h, self._p(out), self._stream()))
        return out

    def export_state(self):
        s, t = self.spec, self.spec.tables
        out = self.torch.empty((self.num_envs, s.num_cores, t.n_edges, s.num_slots), dtype=self.torch.uint8,
                               device=self.device)
        self._ck(self.L.orlg_export_state(self.h, self._p(out), self._stream()))
        return out

    def requests(self):
        out = self.torch.empty((self.num_envs, REQUEST_DTYPE.itemsize), dtype=self.torch.uint8, device=self.device)
        self._ck(self.L.orlg_get_requests(self.h, self._p(out), self._stream()))
        return out.cpu().numpy().view(REQUEST_DTYPE).reshape(self.num_envs)

    INFO_KEYS = ("service_blocking_rate", "episode_service_blocking_rate", "bit_rate_blocking_rate",
                 "episode_bit_rate_blocking_rate", "network_compactness", "network_compactness_difference",
                 "avg_link_compactness", "avg_link_utilization", "fairness")

    def set_info_output(self, enable: bool = True):
        """The info dict of RMSAEnv.step (rmsa_env.py:293-332) written by every following step; returns the
        f64 tensor [num_envs, 9 + n_bit_rates] it lands in (columns: INFO_KEYS, then bit_rate_blocking_<b>)."""
        if not enable:
            self._ck(self.L.orlg_set_info_output(self.h, None))
            self._info = None
            return None
        n = int(self.L.orlg_info_dim(self.h))
        if n <= 0:
            raise RuntimeError("engine was created without info_stats")
        self._info = self.torch.zeros((self.num_envs, n), dtype=self.torch.float64, device=self.device)
        self._ck(self.L.orlg_set_info_output(self.h, self._p(self._info)))
        return self._info

    def phy_channel_state(self, max_entries: int = 256):
        """PhyRMSA-v0 channel_state: (int32 [num_envs, max_entries, 6], int32 [num_envs] counts); rows are
        (src_id * n_nodes + dst_id, idp, channel, used, free, capacity) in append order."""
        torch = self.torch
        out = torch.zeros((self.num_envs, max_entries, 6), dtype=torch.int32, device=self.device)
        cnt = torch.zeros(self.num_envs, dtype=torch.int32, device=self.device)
        self._ck(self.L.orlg_phy_get_channel_state(self.h, self._p(out), max_entries, self._p(cnt), self._stream()))
        return out, cnt

    def set_counter_snapshot(self, enable: bool = True):
        """Every following step also writes the first eight counters at the info dict's point in the step (before
        the next request is counted and before an episode reset); returns the int64 [num_envs, 8] tensor."""
        if not enable:
            self._snap = None
            self._ck(self.L.orlg_set_counter_snapshot(self.h, None))
            return None
        self._snap = self.torch.zeros((self.num_envs, 8), dtype=self.torch.int64, device=self.device)
        self._ck(self.L.orlg_set_counter_snapshot(self.h, self._p(self._snap)))
        return self._snap

    def phy_moves(self):
        """PhyRMSA-v0 defragmentation: int64 [num_envs, 3] counted_moves, counted_moves_groom, counted_defrag_cycles."""
        out = self.torch.empty((self.num_envs, 3), dtype=self.torch.int64, device=self.device)
        self._ck(self.L.orlg_phy_get_moves(self.h, self._p(out), self._stream()))
        return out

    def phy_stats(self):
        """PhyRMSA-v0: f64 [num_envs, 8] (see orlg_phy_get_stats in include/orlg.h)."""
        out = self.torch.empty((self.num_envs, 8), dtype=self.torch.float64, device=self.device)
        self._ck(self.L.orlg_phy_get_stats(self.h, self._p(out), self._stream()))
        return out

    def gsnr_configure(self, gp=None):
        """gp: spec.GsnrParams (default: spec.gsnr_params(tables, channel_width))."""
        if gp is None:
            gp = S.gsnr_params(self.spec.tables, self.spec.channel_width)
        cnt = np.ascontiguousarray(gp.span_count, np.int32)
        ln = np.ascontiguousarray(gp.span_length_km, np.float64)
        c = OrlgGsnrParams(gp.launch_power_w, gp.attenuation_per_m, gp.noise_figure_lin, gp.f0_hz, gp.slot_hz,
                           cnt.ctypes.data, ln.ctypes.data)
        self._ck(self.L.orlg_gsnr_configure(self.h, C.byref(c)))
        self.gsnr_params = gp

    def gsnr(self, query, out=None):
        """query: int32 [num_envs, 4] device tensor (path | core << 16, first_slot, n_slots, se or 0).
        Returns f64 [num_envs, 2]: candidate GSNR dB, lowest neighbour GSNR dB."""
        torch = self.torch
        assert query.dtype == torch.int32 and query.is_cuda and query.is_contiguous()
        if out is None:
            out = torch.empty((self.num_envs, 2), dtype=torch.float64, device=self.device)
        self._ck(self.L.orlg_gsnr(self.h, self._p(query), self._p(out), self._stream()))
        return out

    def set_host_wait(self, blocking: bool):
        """step_host_wait sleeps on an event (True) instead of spinning on the stream (False, default)."""
        self._ck(self.L.orlg_set_host_wait(self.h, int(bool(blocking))))

    def gsnr_for_actions(self, actions, out=None):
        """actions: int32 device tensor in the env's own action format (RMSA-v0 [num_envs, 2], RMCSA-v0
        [num_envs, 4]) for the pending requests.  Returns f64 [num_envs, 2] like :meth:`gsnr`."""
        torch = self.torch
        assert actions.dtype == torch.int32 and actions.is_cuda and actions.is_contiguous()
        assert actions.numel() == self.num_envs * self.action_dim
        if out is None:
            out = torch.empty((self.num_envs, 2), dtype=torch.float64, device=self.device)
        self._ck(self.L.orlg_gsnr_for_actions(self.h, self._p(actions), self._p(out), self._stream()))
        return out

    @property
    def launch_count(self):
        return int(self.L.orlg_launch_count(self.h))

    @property
    def hbm_bytes(self):
        return int(self.L.orlg_hbm_bytes(self.h))

    def close(self):
        if getattr(self, "h", None):
            self.L.orlg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
