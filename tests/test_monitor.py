"""Monitor / VecMonitor csv files and the evaluate-heuristic loop.  The CPU half runs on
stand-in envs (host logic only); the GPU half drives the real scalar and batched envs."""
import os

import numpy as np
import pytest

import golden_util as G
from optical_rl_gym_qot_aware_b200.monitor import Monitor, VecMonitor, evaluate_heuristic, load_results


class _Scalar:
    """Episode of `n` steps, reward 1 on even steps; `arity` 4 (RMSA) or 5 (PhyRMSA) step tuples."""

    def __init__(self, n=5, arity=4):
        self.n, self.arity, self.t, self.resets = n, arity, 0, 0

    def reset(self):
        self.resets += 1
        self.t = 0
        return "obs"

    def step(self, a):
        self.t += 1
        info = {"episode_service_blocking_rate": 1.0 - 1.0 / self.t, "x": a}
        out = ("obs", 1 if self.t % 2 == 0 else 0, self.t == self.n)
        return out + ((False, info) if self.arity == 5 else (info,))

    def heuristic(self, name):
        return 7


@pytest.mark.parametrize("arity", [4, 5])
def test_monitor_csv_format_and_evaluate_heuristic(tmp_path, arity):
    env = Monitor(_Scalar(5, arity), str(tmp_path / "SAP-FF"), info_keywords=("episode_service_blocking_rate",),
                  env_id="RMSA-v0")
    with pytest.raises(RuntimeError):
        env.step(0)  # needs reset first
    seen = []
    mean, std = evaluate_heuristic(env, "shortest_available_path_first_fit", n_eval_episodes=3,
                                   callback=lambda l, g: seen.append(l["episode_length"]))
    assert (mean, std) == (2.0, 0.0) and seen == [1, 2, 3, 4, 5] * 3 and env.resets == 3
    rewards, lengths = evaluate_heuristic(env, lambda e: 1, n_eval_episodes=2, return_episode_rewards=True)
    assert rewards == [2.0, 2.0] and lengths == [5, 5]
    with pytest.raises(AssertionError):
        evaluate_heuristic(env, lambda e: 1, n_eval_episodes=1, reward_threshold=10)
    env.close()
    lines = open(tmp_path / "SAP-FF.monitor.csv").read().splitlines()
    assert lines[0].startswith('#{"t_start": ') and lines[0].endswith('"env_id": "RMSA-v0"}')
    assert lines[1] == "r,l,t,episode_service_blocking_rate"
    assert len(lines) == 2 + 6 and all(ln.startswith("2,5,") and ln.endswith(",0.8") for ln in lines[2:])
    header, cols = load_results(str(tmp_path / "SAP-FF.monitor.csv"))
    assert header["env_id"] == "RMSA-v0" and cols["l"].tolist() == [5.0] * 6 and np.all(np.diff(cols["t"]) >= 0)
    assert env.get_episode_lengths() == [5] * 6 and env.get_total_steps() == 30


class _Batched:
    num_envs = 3

    def __init__(self):
        self.t = 0

    def reset(self):
        return None

    def step(self, actions):
        self.t += 1
        done = np.array([self.t % 2 == 0, self.t % 3 == 0, False])
        return None, np.array([1, 0, 1]), done, {"accepted": np.array([1, 0, 1])}

    def info(self):
        return {"episode_service_blocking_rate": np.array([0.0, 1.0, 0.0])}

    def close(self):
        pass


def test_vecmonitor_rows(tmp_path):
    v = VecMonitor(_Batched(), str(tmp_path), info_keywords=("episode_service_blocking_rate", "accepted"), env_id="X")
    v.reset()
    for _ in range(6):
        v.step(None)
    v.close()
    _, cols = load_results(str(tmp_path))
    # env 0 finishes at t=2,4,6 (2 steps, return 2), env 1 at t=3,6 (3 steps, return 0); at t=6 env 0 is written first
    assert cols["l"].tolist() == [2, 3, 2, 2, 3] and cols["r"].tolist() == [2, 0, 2, 2, 0]
    assert cols["episode_service_blocking_rate"].tolist() == [0, 1, 0, 0, 1] and cols["accepted"].tolist() == [1, 0, 1, 1, 0]
    assert v.episode_count == 5


@pytest.mark.gpu
def test_monitor_on_the_real_envs(tmp_path):
    """Scalar RMSA-v0 under Monitor + evaluate_heuristic, and a 64-env BatchedEnv under VecMonitor: an episode
    is episode_length - 1 steps and the logged reward equals the accepted count."""
    import optical_rl_gym_qot_aware_b200 as pkg

    t = pkg.TopologyTables.load(os.path.join(G.GOLDEN, "topology_nsfnet.npz"))
    kw = dict(topology=t, load=50, num_spectrum_resources=100, episode_length=40, seed=3)
    env = Monitor(pkg.make("RMSA-v0", **kw), str(tmp_path / "sapff"),
                  info_keywords=("episode_service_blocking_rate", "episode_bit_rate_blocking_rate"))
    rewards, lengths = evaluate_heuristic(env, "shortest_available_path_first_fit", n_eval_episodes=4,
                                          return_episode_rewards=True)
    env.close()
    header, cols = load_results(str(tmp_path / "sapff.monitor.csv"))
    assert header["env_id"] == "RMSA-v0" and lengths == [39] * 4 and cols["l"].tolist() == [39.0] * 4
    assert cols["r"].tolist() == rewards
    # blocking rate over the 40 services the episode counters saw (39 served + the pending one re-counted, F7)
    assert np.all((cols["episode_service_blocking_rate"] >= 0) & (cols["episode_service_blocking_rate"] <= 1))

    b = VecMonitor(pkg.make("RMSA-v0", num_envs=64, auto_reset=True, **kw), str(tmp_path / "vec"),
                   info_keywords=("accepted",))
    b.reset()
    for _ in range(39 * 2):
        b.step(b.heuristic("shortest_available_path_first_fit"))
    b.close()
    _, cols = load_results(str(tmp_path / "vec.monitor.csv"))
    assert cols["l"].tolist() == [39.0] * 128 and b.episode_count == 128 and cols["r"].max() <= 39
