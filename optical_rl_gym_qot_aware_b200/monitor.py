"""Episode logging and the evaluate-heuristic loop of the reference's drivers.

The reference's experiments wrap an env in Stable-Baselines3's ``Monitor`` and call
``utils.evaluate_heuristic`` (tests/test_rmsa_threads_us.py:62-128, optical_rl_gym/utils.py:124-162);
the plotting scripts then read the ``*.monitor.csv`` files.  SB3 is a third-party dependency the
reference does not vendor (``stable-baselines3[extra]<2.0.0``, examples/stable_baselines3/requirements.txt:1)
and it is absent here, so its published file format is restated:

    #{"t_start": <time.time() at construction>, "env_id": "<id>"}
    r,l,t,<reset keywords...>,<info keywords...>
    <episode reward rounded to 6>,<episode length>,<seconds since t_start rounded to 6>,<info[key] at done...>

``Monitor`` does this for a scalar ``GymEnv`` (4-tuple or PhyRMSA 5-tuple steps); ``VecMonitor`` for the
batched surfaces (``BatchedEnv`` / ``OrlgVecEnv``): one row per finished episode of any env, in env order
within a step, as SB3's ``VecMonitor`` writes them.
"""
from __future__ import annotations

import csv
import json
import os
import time
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

EXT = "monitor.csv"


def _env_id_of(spec) -> Optional[str]:
    """The registered id (optical_rl_gym/__init__.py:3-31) of an ``EnvSpec``."""
    kind = getattr(spec, "env_kind", None)
    if kind is None:
        return None
    from .spec import ENV_IDS

    return next((name for name, k in ENV_IDS.items() if k == kind), None)


class ResultsWriter:
    """SB3 ``ResultsWriter``: header comment line + ``csv.DictWriter`` rows, flushed per row."""

    def __init__(self, filename: str = "", header: Optional[Dict[str, Any]] = None, extra_keys: Sequence[str] = ()):
        header = header or {}
        if not filename.endswith(EXT):
            filename = os.path.join(filename, EXT) if os.path.isdir(filename) else filename + "." + EXT
        self.filename = filename
        d = os.path.dirname(filename)
        if d:
            os.makedirs(d, exist_ok=True)
        self._fh = open(filename, "wt", newline="\n")
        self._fh.write("#%s\n" % json.dumps(header))
        self._w = csv.DictWriter(self._fh, fieldnames=("r", "l", "t") + tuple(extra_keys))
        self._w.writeheader()
        self._fh.flush()

    def write_row(self, row: Dict[str, Any]) -> None:
        self._w.writerow(row)
        self._fh.flush()

    def close(self) -> None:
        self._fh.close()


class Monitor:
    """Wrap a scalar env; on every ``done`` append ``{"r", "l", "t", *info_keywords}`` to the csv and put
    it into ``info["episode"]``.  Attribute access falls through to the wrapped env, so the reference's
    heuristics (which read ``env.env...`` or ``env...`` attributes) keep working."""

    def __init__(self, env, filename: Optional[str] = None, allow_early_resets: bool = True,
                 reset_keywords: Sequence[str] = (), info_keywords: Sequence[str] = (), env_id: Optional[str] = None):
        self.env = env
        self.t_start = time.time()
        self.reset_keywords, self.info_keywords = tuple(reset_keywords), tuple(info_keywords)
        self.allow_early_resets = allow_early_resets
        if env_id is None:
            spec = getattr(env, "spec", None)
            env_id = getattr(spec, "env_id", None) or getattr(spec, "id", None) or _env_id_of(spec)
        self.results_writer = None
        if filename is not None:
            self.results_writer = ResultsWriter(filename, {"t_start": self.t_start, "env_id": env_id},
                                                self.reset_keywords + self.info_keywords)
        self.rewards: List[float] = []
        self.needs_reset = True
        self.episode_returns: List[float] = []
        self.episode_lengths: List[int] = []
        self.episode_times: List[float] = []
        self.total_steps = 0
        self.current_reset_info: Dict[str, Any] = {}

    def __getattr__(self, name):
        return getattr(self.__dict__["env"], name)

    def reset(self, **kwargs):
        if not self.allow_early_resets and not self.needs_reset:
            raise RuntimeError("Tried to reset an environment before done. If you want to allow early resets, "
                               "wrap your env with Monitor(env, path, allow_early_resets=True)")
        self.rewards = []
        self.needs_reset = False
        for key in self.reset_keywords:
            if key not in kwargs:
                raise ValueError(f"Expected you to pass keyword argument {key} into reset")
            self.current_reset_info[key] = kwargs.pop(key)
        return self.env.reset(**kwargs)

    def step(self, action):
        if self.needs_reset:
            raise RuntimeError("Tried to step environment that needs reset")
        out = self.env.step(action)
        reward, done, info = out[1], out[2], out[-1]
        self.rewards.append(reward)
        if done:
            self.needs_reset = True
            ep_rew, ep_len = sum(self.rewards), len(self.rewards)
            ep_info = {"r": round(ep_rew, 6), "l": ep_len, "t": round(time.time() - self.t_start, 6)}
            for key in self.info_keywords:
                ep_info[key] = info[key]
            self.episode_returns.append(ep_rew)
            self.episode_lengths.append(ep_len)
            self.episode_times.append(time.time() - self.t_start)
            ep_info.update(self.current_reset_info)
            if self.results_writer:
                self.results_writer.write_row(ep_info)
            info["episode"] = ep_info
        self.total_steps += 1
        return out

    def close(self) -> None:
        if self.results_writer is not None:
            self.results_writer.close()
        if hasattr(self.env, "close"):
            self.env.close()

    def get_total_steps(self) -> int:
        return self.total_steps

    def get_episode_rewards(self) -> List[float]:
        return self.episode_returns

    def get_episode_lengths(self) -> List[int]:
        return self.episode_lengths

    def get_episode_times(self) -> List[float]:
        return self.episode_times


class VecMonitor:
    """The batched counterpart: wraps anything whose ``step(actions)`` returns
    ``(obs, rewards[n], dones[n], info)``.  Per-env return and length are accumulated on the host from the
    step's reward / done vectors; ``info_keywords`` are read, for the envs that finished, from the step's
    info (a dict of per-env arrays — ``BatchedEnv`` — or a list of per-env dicts — ``OrlgVecEnv``) and, for
    keys it does not hold, from ``env.info()`` (the counter-derived rates of the reference's info dict)."""

    def __init__(self, venv, filename: Optional[str] = None, info_keywords: Sequence[str] = (), env_id: Optional[str] = None):
        self.venv = venv
        self.num_envs = int(venv.num_envs)
        self.t_start = time.time()
        self.info_keywords = tuple(info_keywords)
        if env_id is None:
            spec = getattr(venv, "spec", None)
            env_id = getattr(spec, "env_id", None) or _env_id_of(spec)
        self.results_writer = None
        if filename is not None:
            self.results_writer = ResultsWriter(filename, {"t_start": self.t_start, "env_id": env_id}, self.info_keywords)
        self.episode_returns = np.zeros(self.num_envs, np.float64)
        self.episode_lengths = np.zeros(self.num_envs, np.int64)
        self.episode_count = 0

    def __getattr__(self, name):
        return getattr(self.__dict__["venv"], name)

    def reset(self, *a, **kw):
        obs = self.venv.reset(*a, **kw)
        self.episode_returns[:] = 0
        self.episode_lengths[:] = 0
        return obs

    @staticmethod
    def _host(x):
        return x.detach().cpu().numpy() if hasattr(x, "detach") else np.asarray(x)

    def _lookup(self, info, key: str, i: int, late: Dict[str, Any]):
        if isinstance(info, dict) and key in info:
            return self._host(info[key])[i].item()
        if isinstance(info, (list, tuple)) and key in info[i]:
            return info[i][key]
        if key not in late:  # fetched once per step, only when an episode ended
            late.update({k: self._host(v) for k, v in self.venv.info().items()} if hasattr(self.venv, "info") else {})
        return late[key][i].item()

    def step(self, actions):
        out = self.venv.step(actions)
        rew, done, info = self._host(out[1]), self._host(out[2]).astype(bool), out[3]
        self.episode_returns += rew
        self.episode_lengths += 1
        if done.any():
            late: Dict[str, Any] = {}
            rows = []
            for i in np.flatnonzero(done):
                row = {"r": round(float(self.episode_returns[i]), 6), "l": int(self.episode_lengths[i]),
                       "t": round(time.time() - self.t_start, 6)}
                for key in self.info_keywords:
                    row[key] = self._lookup(info, key, int(i), late)
                rows.append((int(i), row))
                if self.results_writer:
                    self.results_writer.write_row(row)
                if isinstance(info, (list, tuple)):
                    info[i]["episode"] = row
            if isinstance(info, dict):
                info["episode"] = rows
            self.episode_count += len(rows)
            self.episode_returns[done] = 0
            self.episode_lengths[done] = 0
        return out

    def close(self) -> None:
        if self.results_writer is not None:
            self.results_writer.close()
        self.venv.close()


def load_results(path: str) -> Tuple[Dict[str, Any], Dict[str, np.ndarray]]:
    """Read one ``*.monitor.csv`` (or the ``monitor.csv`` of a directory): ``(header, columns)`` with every
    column as a float array — what the reference's plotting scripts get from ``pd.read_csv(f, skiprows=1)``."""
    if os.path.isdir(path):
        path = os.path.join(path, EXT)
    with open(path, "rt") as fh:
        first = fh.readline()
        if not first.startswith("#"):
            raise ValueError(f"{path}: not a monitor file (no #-header)")
        header = json.loads(first[1:])
        rows = list(csv.DictReader(fh))
    cols = {k: np.array([float(r[k]) for r in rows], np.float64) for k in (rows[0].keys() if rows else ())}
    return header, cols


def _run_episode(env, policy: Callable, on_step: Optional[Callable], render: bool) -> Tuple[float, int]:
    """One episode: counters-only ``reset()``, then policy -> step until the env reports done.
    Works for the 4-tuple envs and for PhyRMSA-v0's 5-tuple alike (reward and done are fields 1 and 2)."""
    env.reset()
    total, steps = 0.0, 0
    while True:
        result = env.step(policy(env))
        total += result[1]
        steps += 1
        if on_step is not None:
            on_step({"env": env, "reward": result[1], "done": result[2], "episode_reward": total,
                     "episode_length": steps, "info": result[-1]}, globals())
        if render:
            env.render()
        if result[2]:
            return total, steps


def evaluate_heuristic(env, heuristic, n_eval_episodes: int = 10, render: bool = False,
                       callback: Optional[Callable] = None, reward_threshold: Optional[float] = None,
                       return_episode_rewards: bool = False):
    """Same contract as utils.py:124-162 (mean / std of the episode rewards over ``n_eval_episodes`` episodes,
    optional threshold assertion, optional raw lists) for both step arities — the reference's own version
    unpacks five values and therefore only runs with PhyRMSA-v0.  ``heuristic`` is a callable
    taking the env, or the name of one of the engine's device heuristics (``env.heuristic(name)``)."""
    policy = (lambda e: e.heuristic(heuristic)) if isinstance(heuristic, str) else heuristic
    episodes = [_run_episode(env, policy, callback, render) for _ in range(n_eval_episodes)]
    rewards, lengths = [r for r, _ in episodes], [n for _, n in episodes]
    if return_episode_rewards and reward_threshold is None:
        return rewards, lengths
    mean_reward, std_reward = float(np.mean(rewards)), float(np.std(rewards))
    if reward_threshold is not None and not mean_reward > reward_threshold:
        raise AssertionError(f"Mean reward below threshold: {mean_reward:.2f} < {reward_threshold:.2f}")
    return (rewards, lengths) if return_episode_rewards else (mean_reward, std_reward)
