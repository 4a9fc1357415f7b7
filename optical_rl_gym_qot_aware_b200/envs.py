"""Gym-facing surface of the engine: same ids, kwargs and return shapes as the reference.

The reference exposes its envs through ``gym.make(id, **kwargs)`` and the gym ``Env``
protocol (optical_rl_gym/__init__.py:3-31; rmsa_env.py:222-341,343-457;
deeprmsa_env.py:48-124; rmcsa_env.py:209-339).  This module keeps that surface and puts
the C-ABI engine underneath:

* :class:`BatchedEnv`   ``num_envs`` envs on one GPU; torch CUDA tensors in and out
  (``reset() -> obs``, ``step(actions) -> (obs, reward, done, info)``).
* :class:`OrlgVecEnv`   Stable-Baselines3 ``VecEnv`` protocol over host numpy arrays
  (``reset``, ``step_async``/``step_wait``, auto-reset = the reference's counters-only
  reset), built on the host-buffer call ``orlg_step_host``.
* :class:`GymEnv`       ``num_envs == 1`` compatibility wrapper returning the scalar
  4-tuple of the reference, with the attributes its heuristics read
  (``current_service``, ``topology.graph["available_slots"]``, ``k_paths``, ``j`` ...).
* :func:`make`          ``make("RMSA-v0", topology=..., num_envs=..., **kwargs)``.

All stepping happens in the CUDA kernels; nothing here computes an env transition.
"""
from __future__ import annotations

from dataclasses import dataclass
from types import SimpleNamespace
from typing import Any, Dict, Optional

import numpy as np

from . import spec as S
from .tables import TopologyTables, tables_from_topology

HEURISTICS = {
    "shortest_path_first_fit": S.H_SP_FF, "sp_ff": S.H_SP_FF,
    "shortest_available_path_first_fit": S.H_SAP_FF, "sap_ff": S.H_SAP_FF, "ksp_ff": S.H_SAP_FF,
    "least_loaded_path_first_fit": S.H_LLP_FF, "llp_ff": S.H_LLP_FF,
    "shortest_available_path_best_modulation_first_core_first_fit": S.H_SAP_BM_FC_FF,
    "sap_bm_fc_ff": S.H_SAP_BM_FC_FF,
    # PhyRMSA-v0 (phy_rmsa_env.py), physical layer
    "sapff_rmsa": S.H_PHY_SAP_FF, "phy_aware_sapbm_rmsa": S.H_PHY_SAP_BM, "phy_aware_bmff_rmsa": S.H_PHY_BM_FF,
    "phy_aware_bmfa_rmsa": S.H_PHY_BM_FA_CUT, "bm_fa_cut": S.H_PHY_BM_FA_CUT,
    "phy_aware_faff_rmsa": S.H_PHY_FA_FF, "fa_ff": S.H_PHY_FA_FF,
    "phy_aware_bmfa_rss_rmsa": S.H_PHY_BM_FA_RSS, "bm_fa_rss": S.H_PHY_BM_FA_RSS,
    "phy_aware_faff_rss_rmsa": S.H_PHY_FA_FF_RSS, "fa_ff_rss": S.H_PHY_FA_FF_RSS,
}


def _as_tables(topology, k_paths: Optional[int] = None) -> TopologyTables:
    if isinstance(topology, TopologyTables):
        return topology
    if isinstance(topology, str):
        if topology.endswith((".txt", ".xml")):  # a raw topology file: build the tables here (k via k_paths=)
            from .topology_builder import build_from_file

            return build_from_file(topology, k_paths=5 if k_paths is None else int(k_paths))
        return TopologyTables.load(topology)
    return tables_from_topology(topology)


def _heuristic_id(h) -> int:
    return HEURISTICS[h] if isinstance(h, str) else int(h)


# -- minimal space descriptions (gym itself is API-only for the reference) -----------------
@dataclass
class Discrete:
    n: int
    shape = ()
    dtype = np.int64


@dataclass
class MultiDiscrete:
    nvec: tuple

    @property
    def shape(self):
        return (len(self.nvec),)

    dtype = np.int64


@dataclass
class Box:
    low: float
    high: float
    shape: tuple
    dtype: Any


class BatchedEnv:
    """``num_envs`` independent envs of one kind on one GPU."""

    def __init__(self, env_id: str, topology, num_envs: int = 1, device: int = 0, global_env_offset: int = 0,
                 request_source: str = "philox", obs_dtype: Optional[str] = None, auto_reset: bool = False,
                 observation: Optional[str] = None, **kwargs):
        import torch

        from .engine import Engine

        self.torch = torch
        self.env_id = env_id
        tables = _as_tables(topology, kwargs.pop("k_paths", None))  # k_paths only matters for raw topology files
        kind = S.ENV_IDS[env_id]
        obs_kind = None
        if kind == S.ENV_PHYRMSA and observation in ("info", "cuts"):
            # the step's number_cuts_total (and rss_total_metric): what UseInfoReward feeds back
            obs_kind = S.OBS_PHY_INFO if observation == "info" else S.OBS_PHY_CUTS
        elif observation == "matrix":  # SimpleMatrixObservation rmsa_env.py:940-971
            obs_kind = S.OBS_MATRIX_U8
        elif kind == S.ENV_DEEPRMSA:
            obs_kind = S.OBS_F32 if obs_dtype in ("float32", "f32") else S.OBS_F64
        src = {"philox": S.SRC_PHILOX, "replay": S.SRC_REPLAY}[request_source]
        self.spec = S.make_spec(env_id, tables, request_source=src, auto_reset=auto_reset, obs_kind=obs_kind,
                                **kwargs)
        self.engine = Engine(self.spec, num_envs, device=device, global_env_offset=global_env_offset)
        self.num_envs = int(num_envs)
        self.device = self.engine.device
        # info_stats=True: the step's info dict of rmsa_env.py:293-332 comes from the step kernel itself
        self._info = self.engine.set_info_output() if self.spec.info_stats else None
        # the counters behind the four blocking rates, as they are where the reference builds its info dict (after the
        # decision, before the next request is counted and before an episode reset)
        self._snap = self.engine.set_counter_snapshot()
        self._stepped = False
        t = tables
        # attributes the reference's drivers and heuristics read
        self.k_paths = t.k_paths
        self.j = self.spec.j
        self.num_spectrum_resources = self.spec.num_slots
        self.num_spatial_resources = self.spec.num_cores
        self.episode_length = self.spec.episode_length
        self.allow_rejection = self.spec.allow_rejection
        self.load = self.spec.load
        self.mean_service_holding_time = self.spec.mean_holding
        self.mean_service_inter_arrival_time = self.spec.mean_iat
        rej = 1 if self.spec.allow_rejection else 0
        if kind == S.ENV_RMSA:  # rmsa_env.py:197-208
            self.action_space = MultiDiscrete((t.k_paths + rej, self.spec.num_slots + rej))
        elif kind == S.ENV_DEEPRMSA:  # deeprmsa_env.py:35-43
            self.action_space = Discrete(t.k_paths * self.j + rej)
        elif kind == S.ENV_PHYRMSA:  # the reference declares (k, S) (phy_rmsa_env.py:251-256) but takes (path, channels)
            self.action_space = MultiDiscrete((t.k_paths + rej,) + (self.spec.num_slots,) * 8)
        else:  # rmcsa_env.py:181-194
            self.action_space = MultiDiscrete((t.k_paths + rej, len(t.mod_names) + rej,
                                               self.spec.num_cores + rej, self.spec.num_slots + rej))
        if self.spec.obs_kind in (S.OBS_F64, S.OBS_F32):
            self.observation_space = Box(0, 1, (self.engine.obs_dim,),
                                         np.float64 if self.spec.obs_kind == S.OBS_F64 else np.float32)
        elif self.spec.obs_kind == S.OBS_MATRIX_U8:
            self.observation_space = Box(0, 1, (self.engine.obs_dim,), np.uint8)
        elif self.spec.obs_kind in (S.OBS_PHY_INFO, S.OBS_PHY_CUTS):
            self.observation_space = Box(0, np.inf, (self.engine.obs_dim,), np.float64)
        else:
            self.observation_space = None
        n, dev = self.num_envs, self.device
        self._reward = torch.zeros(n, dtype=torch.int32, device=dev)
        self._done = torch.zeros(n, dtype=torch.uint8, device=dev)
        self._alloc = torch.zeros((n, 4), dtype=torch.int32, device=dev)
        self._obs = self.engine.new_obs() if self.engine.obs_dim else None
        self._actions = torch.zeros((n, self.engine.action_dim), dtype=torch.int32, device=dev)
        self._started = False
        if src == S.SRC_PHILOX:
            self.reset(only_episode_counters=False)  # the constructor draws the first request (rmsa_env.py:219-220)

    # -- request sources ---------------------------------------------------------------
    def load_trace(self, trace) -> None:
        """Replay source: ``trace[r, i]`` is request r of local env i (engine.REQUEST_DTYPE);
        performs the constructor-time reset, consuming request 0."""
        self.engine.load_trace(trace)
        self.reset(only_episode_counters=False)

    # -- gym protocol --------------------------------------------------------------------
    def reset(self, only_episode_counters: bool = True):
        """rmsa_env.py:343-457: by default only the per-episode counters are cleared and the
        network keeps its state."""
        self.engine.reset(full=not only_episode_counters or not self._started)
        self._started = True
        return self.observation()

    def observation(self):
        if self._obs is None:
            return None
        self.engine.observation(self._obs)
        return self._obs

    def step(self, actions):
        """actions: int tensor/array [num_envs, action_dim] (or [num_envs] for DeepRMSA).
        Returns (obs, reward, done, info); info holds the decision taken per env."""
        torch = self.torch
        a = actions
        if not torch.is_tensor(a):
            a = torch.as_tensor(np.asarray(a), device=self.device)
        a = a.to(device=self.device, dtype=torch.int32).reshape(self.num_envs, self.engine.action_dim).contiguous()
        self.engine.step(a, self._reward, self._done, self._obs, self._alloc)
        self._stepped = True
        info = {"accepted": self._alloc[:, 0], "path": self._alloc[:, 1] & 0xFFFF, "core": self._alloc[:, 1] >> 16,
                "initial_slot": self._alloc[:, 2], "number_slots": self._alloc[:, 3]}
        return self._obs, self._reward, self._done, info

    def reward(self):
        return self._reward

    def render(self, mode="human"):  # rmsa_env.py:459-460
        return None

    def close(self):
        self.engine.close()

    # -- device heuristics and fused rollouts -----------------------------------------------
    def heuristic(self, name):
        """Action the named reference heuristic would take, for every env (device tensor)."""
        self.engine.heuristic(_heuristic_id(name), self._actions)
        return self._actions

    def rollout(self, name, nsteps: int, reward_sum=None):
        """``nsteps`` x (heuristic -> step) in one launch (utils.evaluate_heuristic, utils.py:124-162)."""
        self.engine.rollout(_heuristic_id(name), nsteps, reward_sum)

    # -- views ------------------------------------------------------------------------------------
    def counters(self):
        return self.engine.counters()

    def phy_info(self) -> Dict[str, Any]:
        """PhyRMSA-v0: the remaining entries of the reference's info dict (phy_rmsa_env.py:319-348),
        per env, with the fragmentation metrics of the current channel map."""
        st = self.engine.phy_stats()
        return {
            "number_cuts_total": st[:, 6], "rss_total_metric": st[:, 7],
            "total_path_length": st[:, 0] / (st[:, 5] + 1), "avrage_gsnr": st[:, 1] / (st[:, 4] + 1),
            "average_mod_level": st[:, 3] / (st[:, 4] + 1), "average_path_index": st[:, 2] / (st[:, 5] + 1),
            "path_index": st[:, 2], "physical_paths": st[:, 5],
            **self._phy_moves(),
        }

    def _phy_moves(self):
        mv = self.engine.phy_moves().to(self.torch.float64)  # phy_rmsa_env.py:340-342
        return {"num_moves": mv[:, 0] / 2 + mv[:, 1], "num_moves_groom": mv[:, 1], "num_defrag_cycle": mv[:, 2]}

    def available_channels(self):
        """PhyRMSA-v0: topology.graph["available_channels"] of every env, uint8 [num_envs, E, C]."""
        return self.engine.export_state()[:, 0]

    def available_slots(self):
        """topology.graph["available_slots"] of every env: uint8 [num_envs, cores, E, S]."""
        return self.engine.export_state()

    def current_requests(self):
        return self.engine.requests()

    def info(self) -> Dict[str, Any]:
        """The reference's info dict (rmsa_env.py:293-332), per env.  With ``info_stats=True`` every entry
        (compactness, link statistics, fairness, per-bit-rate blocking) is the value the last step computed at
        the reference's point in the step; otherwise only the counter-derived rates, from the current counters."""
        if self._info is not None:
            out = {k: self._info[:, i] for i, k in enumerate(self.engine.INFO_KEYS)}
            for i, b in enumerate(np.asarray(self.spec.bit_rates).tolist()):
                out[f"bit_rate_blocking_{b}"] = self._info[:, 9 + i]
            return out
        # after a step: the snapshot the step kernel took at the reference's point (rmsa_env.py:293-307); before the
        # first step there is no info dict in the reference either - the current counters stand in
        c = (self._snap if self._stepped else self.counters()[:, :8]).to(self.torch.float64)
        safe = lambda x: self.torch.clamp(x, min=1.0)  # noqa: E731
        return {
            "service_blocking_rate": (c[:, 0] - c[:, 1]) / safe(c[:, 0]),
            "episode_service_blocking_rate": (c[:, 2] - c[:, 3]) / safe(c[:, 2]),
            "bit_rate_blocking_rate": (c[:, 4] - c[:, 5]) / safe(c[:, 4]),
            "episode_bit_rate_blocking_rate": (c[:, 6] - c[:, 7]) / safe(c[:, 6]),
        }

    def episode_stats(self, reward_sum=None, group=None):
        """Statistics summed over every rank's shard (the path's only collective)."""
        from .sharding import reduce_episode_stats

        return reduce_episode_stats(self.counters(), reward_sum, group)


class OrlgVecEnv:
    """Stable-Baselines3 ``VecEnv`` protocol (not shipped by the reference) over
    host numpy arrays.  ``step_wait`` makes one ``orlg_step_host`` call: actions H2D from pinned
    memory, one step of every env, obs/reward/done D2H, stream synchronised.  Envs whose episode
    ended are reset in-kernel with the reference's counters-only reset (rmsa_env.py:344-389)."""

    def __init__(self, env_id: str, topology, num_envs: int, **kwargs):
        import torch

        kwargs.setdefault("auto_reset", True)
        self.env = BatchedEnv(env_id, topology, num_envs=num_envs, **kwargs)
        self.num_envs = self.env.num_envs
        self.action_space = self.env.action_space
        self.observation_space = self.env.observation_space
        e = self.env.engine
        odt = {S.OBS_F64: torch.float64, S.OBS_F32: torch.float32, S.OBS_MATRIX_U8: torch.uint8}.get(
            self.env.spec.obs_kind)
        n = self.num_envs
        self._h_act = torch.zeros((n, e.action_dim), dtype=torch.int32).pin_memory()
        self._h_rew = torch.zeros(n, dtype=torch.int32).pin_memory()
        self._h_done = torch.zeros(n, dtype=torch.uint8).pin_memory()
        self._h_alloc = torch.zeros((n, 4), dtype=torch.int32).pin_memory()
        # the device buffer is feature-major; the host copy is exposed as its [num_envs, obs_dim] view
        self._h_obs = torch.zeros((e.obs_dim, n), dtype=odt).pin_memory() if e.obs_dim else None
        self._pending = False
        self._stream = torch.cuda.Stream(device=self.env.device)  # own stream: several VecEnvs overlap

    def _obs_view(self):
        return None if self._h_obs is None else self._h_obs.numpy().T

    def reset(self):
        obs = self.env.reset(only_episode_counters=True)
        if obs is not None:
            self._h_obs.copy_(obs.t())
            self.env.torch.cuda.synchronize(self.env.device)
        return self._obs_view()

    def step_async(self, actions) -> None:
        a = np.asarray(actions, dtype=np.int32).reshape(self.num_envs, -1)
        self._h_act.numpy()[:] = a
        self._pending = True
        self.env._stepped = True  # info() now reads the step kernel's counter snapshot
        self._stream.wait_stream(self.env.torch.cuda.current_stream(self.env.device))
        self.env.engine.step_host_async(self._h_act, self._h_rew, self._h_done, self._h_obs, self._h_alloc,
                                        stream=self._stream)

    def step_wait(self):
        assert self._pending, "step_wait without step_async"
        self._pending = False
        self.env.engine.step_host_wait(stream=self._stream)
        rew = self._h_rew.numpy().astype(np.float32)
        done = self._h_done.numpy().astype(bool)
        al = self._h_alloc.numpy()
        # SB3 indexes infos[i]: always one dict per env; the same columns as arrays are kept in `last_info_arrays`
        cols = {"accepted": al[:, 0].astype(bool), "path": al[:, 1] & 0xFFFF, "core": al[:, 1] >> 16,
                "initial_slot": al[:, 2], "number_slots": al[:, 3]}
        self.last_info_arrays = cols
        lists = {k: v.tolist() for k, v in cols.items()}
        infos = [{k: lists[k][i] for k in lists} for i in range(self.num_envs)]
        return self._obs_view(), rew, done, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def info(self):
        """The info dict of the last step, per env (arrays): what Monitor / VecMonitor log at an episode's end."""
        return self.env.info()

    def seed(self, seed=None):
        return [None] * self.num_envs

    def get_attr(self, name, indices=None):
        return [getattr(self.env, name)] * self.num_envs

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False] * self.num_envs

    def close(self):
        self.env.close()


class GymEnv:
    """Scalar gym ``Env`` look-alike over a one-env engine: ``step`` returns the reference's
    4-tuple ``(observation, reward, done, info)`` (rmsa_env.py:336-341) with Python scalars."""

    metadata = {"metrics": ["service_blocking_rate", "episode_service_blocking_rate", "bit_rate_blocking_rate",
                            "episode_bit_rate_blocking_rate"]}

    def __init__(self, env_id: str, topology, **kwargs):
        self._phy = S.ENV_IDS[env_id] == S.ENV_PHYRMSA
        if self._phy:
            kwargs.setdefault("observation", "info")  # number_cuts_total / rss_total_metric at the reference's point
        self._b = BatchedEnv(env_id, topology, num_envs=1, **kwargs)
        for name in ("k_paths", "j", "num_spectrum_resources", "num_spatial_resources", "episode_length",
                     "allow_rejection", "load", "mean_service_holding_time", "mean_service_inter_arrival_time",
                     "action_space", "observation_space", "spec"):
            setattr(self, name, getattr(self._b, name))
        self.tables = self._b.spec.tables
        self.topology = _TopologyView(self)
        self.k_shortest_paths = _KspView(self.tables)
        self.channel_width = self._b.spec.channel_width
        self.grooming = self._b.spec.phy_grooming == 1
        self._state_cache = None
        if self._phy:  # phy_rmsa_env.py:51-53,93-95
            ph = self._b.spec.phy
            self.modulation_level, self.gsnr = ph.mod_level.astype(np.int64), ph.gsnr
            self.connections_detail = np.asarray(kwargs.get("connections_detail"))

    # -- read-only views of the network state for code written against the reference env (its module-level
    # heuristics rmsa_env.py:854-937, deeprmsa_env.py:135-155 read these); every call works on a host copy of the
    # occupancy fetched once per step --
    def _state(self):
        if self._state_cache is None:
            self._state_cache = self._b.available_slots()[0].cpu().numpy().astype(np.int64)  # [cores, E, S]
        return self._state_cache

    def _links(self, path):
        return [self.topology[path.node_list[i]][path.node_list[i + 1]]["index"] for i in range(len(path.node_list) - 1)]

    def get_number_slots(self, path, modulation_format=None):
        """rmsa_env.py:708-719 / rmcsa_env.py:753-765"""
        import math

        se = (modulation_format or path.best_modulation).spectral_efficiency
        return math.ceil(self.current_service.bit_rate / (se * self.channel_width)) + 1

    def is_path_free(self, path, *args):
        """rmsa_env.py:721-734 (path, initial_slot, number_slots) / rmcsa_env.py:767-794 (path, core, initial_slot,
        number_slots)"""
        core, slot, n = (args if len(args) == 3 else (0,) + tuple(args))
        if slot + n > self.num_spectrum_resources:
            return False
        st = self._state()
        return all(not np.any(st[core, l, slot:slot + n] == 0) for l in self._links(path))

    def get_available_slots(self, path):
        """rmsa_env.py:745-756: AND of the path's link rows"""
        st = self._state()
        out = np.ones(self.num_spectrum_resources, np.int64)
        for l in self._links(path):
            out = out * st[0, l]
        return out

    @staticmethod
    def rle(inarray):
        """rmsa_env.py:758-772: (initial indices, values, lengths)"""
        ia = np.asarray(inarray)
        n = len(ia)
        if n == 0:
            return None, None, None
        y = np.array(ia[1:] != ia[:-1])
        i = np.append(np.where(y), n - 1)
        z = np.diff(np.append(-1, i))
        return np.cumsum(np.append(0, z))[:-1], ia[i], z

    def get_available_blocks(self, path_index):
        """rmsa_env.py:774-804: the first j blocks of >= the required slots on path `path_index`"""
        cs = self.current_service
        path = self.k_shortest_paths[cs.source, cs.destination][path_index]
        slots = self.get_number_slots(path)
        starts, values, lengths = self.rle(self.get_available_slots(path))
        ok = [i for i in range(len(values)) if values[i] == 1 and lengths[i] >= slots][: self.j]
        return starts[ok], lengths[ok]

    def is_channel_free(self, path, channel_number):
        """phy_rmsa_env.py:1029-1035"""
        st = self._state()
        return all(st[0, l, channel_number] != 0 for l in self._links(path))

    @property
    def channel_state(self):
        """phy_rmsa_env.py:158-166: {(src_id, dst_id, idp): [(channel, used, free, capacity), ...]} in append order"""
        n, k = self.tables.n_nodes, self.k_paths
        out = {(a, b, i): [] for a in range(n) for b in range(n) for i in range(k)}
        if self._b.spec.phy_grooming:
            rows, cnt = self._b.engine.phy_channel_state(4096)
            for r in rows[0, : int(cnt[0])].cpu().numpy():
                out[(int(r[0]) // n, int(r[0]) % n, int(r[1]))].append((int(r[2]), int(r[3]), int(r[4]), int(r[5])))
        return out

    # -- the attributes reference heuristics read (rmsa_env.py:854-937, deeprmsa_env.py:135-155) --
    @property
    def current_service(self):
        r = self._b.current_requests()[0]
        c = self._b.counters()[0].cpu().numpy()
        names = self.tables.node_names
        return SimpleNamespace(
            service_id=int(c[2]) - 1 if self._b.spec.env_kind != S.ENV_RMCSA else int(c[2]),  # rmsa_env.py:661-674
            source=names[int(r["src"])], source_id=int(r["src"]), destination=names[int(r["dst"])],
            destination_id=int(r["dst"]), arrival_time=float(r["arrival"]), holding_time=float(r["holding"]),
            bit_rate=int(r["bit_rate"]))

    def _counter(self, i):
        return int(self._b.counters()[0, i])

    services_processed = property(lambda s: s._counter(0))
    services_accepted = property(lambda s: s._counter(1))
    episode_services_processed = property(lambda s: s._counter(2))
    episode_services_accepted = property(lambda s: s._counter(3))
    bit_rate_requested = property(lambda s: s._counter(4))
    bit_rate_provisioned = property(lambda s: s._counter(5))
    episode_bit_rate_requested = property(lambda s: s._counter(6))
    episode_bit_rate_provisioned = property(lambda s: s._counter(7))

    def _obs_np(self, obs):
        if obs is None or self._phy:  # rmsa_env.py:697-701 / phy_rmsa_env.py: {"topology", "current_service"}
            return {"topology": self.topology, "current_service": self.current_service}
        return obs[0].cpu().numpy().copy()

    def _phy_action(self, action):
        """(path, [(channel, used, free, capacity, virtual), ...]) of the reference heuristics
        (phy_rmsa_env.py:1429-1432) or a flat (path, ch0, ch1, ...) -> the engine's 9 ints."""
        a = np.full(9, -1, np.int64)
        if len(action) == 2 and isinstance(action[1], (list, tuple)):
            a[0] = action[0]
            for i, c in enumerate(action[1][:8]):
                a[1 + i] = c[0] if isinstance(c, (list, tuple)) else c
        else:
            flat = np.asarray(action, dtype=np.int64).reshape(-1)
            a[: len(flat)] = flat
        return a

    def reset(self, only_episode_counters: bool = True):
        self._state_cache = None
        return self._obs_np(self._b.reset(only_episode_counters))

    def observation(self):
        return self._obs_np(self._b.observation())

    def step(self, action):
        """4-tuple like rmsa_env.py:336-341; PhyRMSA-v0 returns the reference's 5-tuple
        (obs, reward, done, False, info) of phy_rmsa_env.py:418-424."""
        if self._phy:
            action = self._phy_action(action)
        self._state_cache = None
        obs, rew, done, info = self._b.step(np.atleast_1d(np.asarray(action, dtype=np.int64))[None])
        i = {k: (float(v[0]) if v.dtype.is_floating_point else int(v[0])) for k, v in self._b.info().items()}
        i.update({k: int(v[0]) for k, v in info.items()})
        if self._phy:
            i.update({k: float(v[0]) for k, v in self._b.phy_info().items()})
            if obs is not None:  # the step's own metrics, taken before the releases (phy_rmsa_env.py:313-314)
                m = obs[0].cpu().numpy()
                i["number_cuts_total"] = float(m[0])
                if m.size > 1:
                    i["rss_total_metric"] = float(m[1])
            return self._obs_np(obs), int(rew[0]), bool(done[0]), False, i
        return self._obs_np(obs), int(rew[0]), bool(done[0]), i

    def heuristic(self, name):
        a = self._b.heuristic(name)[0].cpu().numpy()
        if self._phy:  # rebuild the reference's channel tuples (channel, used, free, capacity, False)
            if a[0] < 0:
                return (-2, [])
            ph, t, cs = self.spec.phy, self.tables, self._b.current_requests()[0]
            table = ph.table_id[int(cs["src"]) * t.n_nodes + int(cs["dst"])]
            un, out = int(cs["bit_rate"]), []
            if a[0] >= 20:  # virtual layer: (idp + 20, [(channel, share, left over, capacity, True)]) :1380,1664-1670
                rows, cnt = self._b.engine.phy_channel_state(4096)
                rows = rows[0, : int(cnt[0])].cpu().numpy()
                key = int(cs["src"]) * t.n_nodes + int(cs["dst"])
                free = {int(r[2]): (int(r[4]), int(r[5])) for r in rows if r[0] == key and r[1] == a[0] - 20}
                for ch in [int(c) for c in a[1:] if c >= 0]:
                    fr, cap = free[ch]
                    un -= fr * 100
                    out.append((ch, fr + un / 100, un / -100, cap, True) if un <= 0 else (ch, fr, 0, cap, True))
                return (int(a[0]), out)
            for ch in [int(c) for c in a[1:] if c >= 0]:
                lvl = int(ph.mod_level[table, ch, int(a[0])])
                un -= lvl * 100
                out.append((ch, lvl + un / 100, un / -100, lvl, False) if un <= 0 else (ch, lvl, 0, lvl, False))
            return (int(a[0]), out)
        return int(a[0]) if a.size == 1 else tuple(int(x) for x in a)

    def load_trace(self, trace):
        self._state_cache = None
        self._b.load_trace(np.asarray(trace).reshape(-1, 1))

    def render(self, mode="human"):
        return None

    def close(self):
        self._b.close()


class _KspView:
    """``env.k_shortest_paths[src, dst]`` -> list of path objects with the fields of utils.Path (utils.py:27-36)."""

    def __init__(self, t: TopologyTables):
        self._t, self._cache = t, {}
        self._nid = {name: i for i, name in enumerate(t.node_names)}

    def __getitem__(self, key):
        a, b = key
        if key not in self._cache:
            t = self._t
            pr = self._nid[a] * t.n_nodes + self._nid[b]
            out = []
            for ip in range(t.k_paths):
                h = int(t.path_hops[pr, ip])
                if h <= 0:
                    continue
                m = int(t.path_mod[pr, ip])
                mod = SimpleNamespace(name=t.mod_names[m] if m >= 0 else None, spectral_efficiency=int(t.path_se[pr, ip]),
                                      maximum_length=float(t.mod_max_length[m]) if m >= 0 else None)
                out.append(SimpleNamespace(path_id=pr * t.k_paths + ip, idp=ip, hops=h, length=float(t.path_length[pr, ip]),
                                           node_list=tuple(t.node_names[int(x)] for x in t.path_nodes[pr, ip, : h + 1]),
                                           best_modulation=mod, current_modulation=None))
            self._cache[key] = out
        return self._cache[key]


class _TopologyView:
    """``env.topology``: ``.graph`` (see _GraphView) and ``topology[n1][n2]["index" | "id" | "length"]``."""

    def __init__(self, env: "GymEnv"):
        self.graph = _GraphView(env)
        t = env.tables
        self._adj: Dict[str, Dict[str, Dict[str, Any]]] = {n: {} for n in t.node_names}
        for e, (a, b) in enumerate(t.edge_nodes):
            d = {"index": e, "id": e, "length": float(t.edge_length[e])}
            self._adj[t.node_names[int(a)]][t.node_names[int(b)]] = d
            self._adj[t.node_names[int(b)]][t.node_names[int(a)]] = d

    def __getitem__(self, node):
        return self._adj[node]

    def nodes(self):
        return list(self._adj)

    def number_of_edges(self):
        return sum(len(v) for v in self._adj.values()) // 2


class _GraphView(dict):
    """``env.topology.graph`` with a lazily fetched ``available_slots`` (rmsa_env.py:396-398)."""

    def __init__(self, env: GymEnv):
        super().__init__(k_paths=env.k_paths, num_spectrum_resources=env.num_spectrum_resources,
                         name=env.tables.name, node_indices=list(env.tables.node_names))
        self._env = env

    def __getitem__(self, key):
        if key in ("available_slots", "available_channels"):
            a = self._env._b.available_slots()[0].cpu().numpy().astype(np.int64)
            return a if self._env._b.spec.env_kind == S.ENV_RMCSA else a[0]
        if key == "num_channel_resources":
            return self._env.num_spectrum_resources
        return super().__getitem__(key)


def make(env_id: str, topology=None, num_envs: Optional[int] = None, vec_env: bool = False, **kwargs):
    """Counterpart of ``gym.make(id, **kwargs)`` for the ids in optical_rl_gym/__init__.py:3-31.
    ``num_envs=None`` -> scalar :class:`GymEnv`; an int -> :class:`BatchedEnv` (device tensors) or,
    with ``vec_env=True``, :class:`OrlgVecEnv` (host numpy, SB3 protocol)."""
    if env_id not in S.ENV_IDS:
        raise KeyError(f"unknown env id {env_id!r}; known: {sorted(S.ENV_IDS)}")
    if topology is None:
        raise TypeError("make() needs topology=")
    if num_envs is None:
        return GymEnv(env_id, topology, **kwargs)
    if vec_env:
        return OrlgVecEnv(env_id, topology, num_envs, **kwargs)
    return BatchedEnv(env_id, topology, num_envs=num_envs, **kwargs)


class PathOnlyFirstFitAction:
    """rmsa_env.py:974-1008 / rmcsa_env.py:950-982: the agent picks only the path; the first fitting slot is found on
    the device (``orlg_path_first_fit``).  Wraps a :class:`BatchedEnv` of RMSA-v0 or RMCSA-v0."""

    def __init__(self, env: BatchedEnv):
        # RMCSA-v0 (rmcsa_env.py:950-982): the path's best modulation by reach, first core with room, first fit
        assert env.spec.env_kind in (S.ENV_RMSA, S.ENV_RMCSA)
        self.env = env
        self.action_space = Discrete(env.k_paths + (1 if env.allow_rejection else 0))
        self.observation_space = env.observation_space

    def __getattr__(self, name):
        return getattr(self.env, name)

    def action(self, paths):
        torch = self.env.torch
        p = torch.as_tensor(paths, device=self.env.device).to(torch.int32).reshape(self.env.num_envs).contiguous()
        self.env.engine.path_first_fit(p, self.env._actions)
        return self.env._actions

    def step(self, paths):
        return self.env.step(self.action(paths))


class UseInfoReward:
    """wrappers.py:4-16: replace the reward by ``info[key]``.  With PhyRMSA-v0 and
    ``key in ("number_cuts_total", "rss_total_metric")`` this is the physical-layer
    fragmentation-aware reward of the phy-frag configuration."""

    def __init__(self, env, key: str):
        self.env, self.key = env, key

    def __getattr__(self, name):
        return getattr(self.env, name)

    def step(self, action):
        out = self.env.step(action)
        info = out[-1]
        return (out[0], info[self.key]) + tuple(out[2:])
