"""Generate golden vectors by running the UNMODIFIED reference (TEST INFRASTRUCTURE).

Run in the build container (needs /root/reference):

    python oracle/refharness/gen_golden.py            # writes tests/golden/*.npz

For every case it drives a reference env (gym.make under the stubs in
oracle/refharness/stubs) with a reference heuristic or seeded random actions and
records, per step: the request that was pending (raw f64 arrival/holding, node
ids, bit rate), the action, accepted/reward/done, the allocation (path, slot,
slots, core), for DeepRMSA the f64 observation; and at the end the
available_slots array and the counters.  Episodes are reset with the
reference's default reset() (counters only).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)

import refenv  # noqa: E402

gym = refenv.install()

from optical_rl_gym.envs import deeprmsa_env, rmcsa_env, rmsa_env  # noqa: E402

from optical_rl_gym_qot_aware_b200.tables import tables_from_topology  # noqa: E402

GOLDEN = os.path.join(REPO, "tests", "golden")

REQ_DTYPE = np.dtype(
    [("arrival", "<f8"), ("holding", "<f8"), ("bit_rate", "<i4"), ("src", "<i2"), ("dst", "<i2")]
)

# node probabilities of tests/test_deeprmsa.py:30-47 (NSFNET, 14 nodes)
DEEPRMSA_NODE_PROBS = np.array(
    [0.01801802, 0.04004004, 0.05305305, 0.01901902, 0.04504505, 0.02402402, 0.06706707,
     0.08908909, 0.13813814, 0.12212212, 0.07607608, 0.12012012, 0.01901902, 0.16916917]
)


def req_of(env):
    s = env.current_service
    return (float(s.arrival_time), float(s.holding_time), int(s.bit_rate), int(s.source_id),
            int(s.destination_id))


def counters_of(env):
    return np.array(
        [env.services_processed, env.services_accepted, env.episode_services_processed,
         env.episode_services_accepted, env.bit_rate_requested, env.bit_rate_provisioned,
         env.episode_bit_rate_requested, env.episode_bit_rate_provisioned], np.int64)


INFO_KEYS = ("service_blocking_rate", "episode_service_blocking_rate", "bit_rate_blocking_rate",
             "episode_bit_rate_blocking_rate", "network_compactness", "network_compactness_difference",
             "avg_link_compactness", "avg_link_utilization")
ONLY = set(sys.argv[1:])  # optional: names of the cases to (re)generate


def run_case(name, env_id, topo_file, kwargs, policy, n_steps, action_seed=0, want_info=False):
    if ONLY and name not in ONLY:
        return
    topo = refenv.load_topology(topo_file)
    env = gym.make(env_id, topology=topo, **kwargs)
    kind = {"RMSA-v0": 0, "DeepRMSA-v0": 1, "RMCSA-v0": 2}[env_id]
    adim = {0: 2, 1: 1, 2: 4}[kind]
    rng = np.random.RandomState(action_seed)
    reqs = np.zeros(n_steps + 1, REQ_DTYPE)
    actions = np.zeros((n_steps, adim), np.int32)
    accepted = np.zeros(n_steps, np.uint8)
    reward = np.zeros(n_steps, np.int32)
    done = np.zeros(n_steps, np.uint8)
    alloc = np.full((n_steps, 4), -1, np.int32)  # path, slot, nslots, core
    counters = np.zeros((n_steps, 8), np.int64)
    obs = None
    obs0 = None
    info_main = np.zeros((n_steps, len(INFO_KEYS)), np.float64)
    info_fair = np.zeros(n_steps, np.float64)
    info_br = np.zeros((n_steps, len(getattr(env, "bit_rates", []))), np.float64)
    if kind == 1:
        obs0 = np.asarray(env.observation(), np.float64)
        obs = np.zeros((n_steps, obs0.size), np.float64)
    reqs[0] = req_of(env)
    S, k = env.num_spectrum_resources, env.k_paths
    for t in range(n_steps):
        if policy == "random":
            if kind == 0:
                a = [rng.randint(k + 1), rng.randint(S + 1)]
            elif kind == 1:
                a = [rng.randint(k * env.j + 1)]
            else:
                a = [rng.randint(k + 1), rng.randint(len(env.modulation_formats) + 1),
                     rng.randint(env.num_spatial_resources + 1), rng.randint(S + 1)]
        else:
            a = policy(env)
            a = [a] if np.isscalar(a) else list(a)
            if kind == 2 and len(a) == 3:  # F16: 3-tuple on "no fit" -> full reject action
                a = [k, len(env.modulation_formats), env.num_spatial_resources, S]
        actions[t] = a
        svc = env.current_service
        out = env.step(a[0] if kind == 1 else a)
        o, r, d = out[0], out[1], out[2]
        if want_info:  # the info dict of rmsa_env.py:293-332, f64 as computed
            info_main[t] = [out[3][key] for key in INFO_KEYS]
            info_fair[t] = out[3]["fairness"]
            info_br[t] = [out[3][f"bit_rate_blocking_{b}"] for b in env.bit_rates]
        accepted[t] = 1 if svc.accepted else 0
        reward[t] = r
        done[t] = 1 if d else 0
        if svc.accepted:
            core = svc.core if kind == 2 else 0
            pid = [p is svc.path or p == svc.path for p in env.k_shortest_paths[svc.source, svc.destination]].index(True)
            alloc[t] = (pid, svc.initial_slot, svc.number_slots, core)
        if kind == 1:
            obs[t] = o
        counters[t] = counters_of(env)
        reqs[t + 1] = req_of(env)
        if d:
            env.reset()
            counters[t] = counters_of(env)  # value after the counters-only reset
    avail = np.asarray(env.topology.graph["available_slots"], np.uint8)
    if avail.ndim == 2:
        avail = avail[None]
    out = dict(
        env_kind=kind, topo_file=topo_file, requests=reqs, actions=actions, accepted=accepted,
        reward=reward, done=done, alloc=alloc, counters=counters, final_avail=avail,
        kwargs_keys=np.array(sorted(k for k, v in kwargs.items() if v is not None)),
    )
    for key, v in kwargs.items():
        if key != "topology" and v is not None:
            out["kw_" + key] = np.asarray(v)
    if want_info:
        out.update(info_main=info_main, info_fairness=info_fair, info_bit_rate_blocking=info_br)
    if obs is not None:
        out["obs"] = obs
        out["obs0"] = obs0
    if kind == 2:
        out["worst_xt_post"] = np.float64(env.worst_xt)
        out["inband_xt_post"] = np.array([m.inband_xt for m in env.modulation_formats], np.float64)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **out)
    blk = 1.0 - accepted.mean()
    print(f"{name}: {n_steps} steps, blocking {blk:.3f}, episodes {int(done.sum())}, "
          f"{os.path.getsize(path) / 1024:.0f} KiB")


TOPOLOGIES = {
    "nsfnet": "nsfnet_chen_5-paths_6-modulations.h5",
    "us14": "us14_3-paths_6-modulations.h5",
    "jpn12_k5": "jpn12_5-paths_6-modulations.h5",
    "jpn12_k3": "jpn12_3-paths_6-modulations.h5",
    "spn": "spn_3-paths_6-modulations.h5",
    "germany50": "germany50_5-paths_6-modulations.h5",
}


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    for short, f in TOPOLOGIES.items():
        if not ONLY:
            tables_from_topology(refenv.load_topology(f)).save(os.path.join(GOLDEN, f"topology_{short}.npz"))
    nsf, spn, jpn = TOPOLOGIES["nsfnet"], TOPOLOGIES["spn"], TOPOLOGIES["jpn12_k5"]
    c1 = dict(seed=10, load=50, num_spectrum_resources=100, episode_length=1000)
    run_case("rmsa_nsfnet_sapff", "RMSA-v0", nsf, c1, rmsa_env.shortest_available_path_first_fit, 2500)
    run_case("rmsa_nsfnet_spff", "RMSA-v0", nsf, dict(c1, seed=11, episode_length=200),
             rmsa_env.shortest_path_first_fit, 1000)
    run_case("rmsa_nsfnet_llpff", "RMSA-v0", nsf, dict(c1, seed=12, episode_length=300),
             rmsa_env.least_loaded_path_first_fit, 1000)
    run_case("rmsa_nsfnet_random", "RMSA-v0", nsf,
             dict(c1, seed=13, load=20, episode_length=100, allow_rejection=True), "random", 1500, 7)
    c2 = dict(seed=10, j=1, mean_service_holding_time=7.5, mean_service_inter_arrival_time=1 / 12.0,
              num_spectrum_resources=100, episode_length=50, node_request_probabilities=DEEPRMSA_NODE_PROBS)
    run_case("deeprmsa_nsfnet_kspff", "DeepRMSA-v0", nsf, c2,
             deeprmsa_env.shortest_available_path_first_fit, 1200)
    run_case("deeprmsa_nsfnet_j3_random", "DeepRMSA-v0", nsf,
             dict(c2, seed=21, j=3, allow_rejection=True, episode_length=70), "random", 1000, 3)
    run_case("deeprmsa_nsfnet_spff", "DeepRMSA-v0", nsf, dict(c2, seed=22, allow_rejection=True),
             deeprmsa_env.shortest_path_first_fit, 600)
    # the info dict: compactness, time-weighted link statistics, per-bit-rate blocking
    run_case("info_rmsa_nsfnet_sapff", "RMSA-v0", nsf, dict(c1, seed=14, load=45, mean_service_holding_time=25,
                                                             episode_length=150),
             rmsa_env.shortest_available_path_first_fit, 700, want_info=True)
    run_case("info_rmsa_nsfnet_random", "RMSA-v0", nsf,
             dict(c1, seed=15, load=30, mean_service_holding_time=25, episode_length=90, allow_rejection=True),
             "random", 600, 9, want_info=True)
    run_case("info_deeprmsa_nsfnet_kspff", "DeepRMSA-v0", nsf, dict(c2, seed=23),
             deeprmsa_env.shortest_available_path_first_fit, 600, want_info=True)
    c5 = dict(seed=30, load=100, mean_service_holding_time=25, num_spectrum_resources=100, episode_length=400)
    run_case("rmsa_spn_sapff", "RMSA-v0", spn, c5, rmsa_env.shortest_available_path_first_fit, 1500)
    run_case("rmsa_spn320_sapff", "RMSA-v0", spn, dict(c5, seed=31, load=400, num_spectrum_resources=320),
             rmsa_env.shortest_available_path_first_fit, 1500)
    c3 = dict(seed=40, load=250, mean_service_holding_time=25, num_spectrum_resources=320,
              num_spatial_resources=7, worst_xt=-84.7, episode_length=500)
    run_case("rmcsa_jpn12_heur", "RMCSA-v0", jpn, c3,
             rmcsa_env.shortest_available_path_best_modulation_first_core_first_fit, 1500)
    run_case("rmcsa_nsfnet_heur", "RMCSA-v0", nsf,
             dict(seed=41, load=250, mean_service_holding_time=25, num_spectrum_resources=64,
                  num_spatial_resources=7, worst_xt=-84.7, episode_length=300),
             rmcsa_env.shortest_available_path_best_modulation_first_core_first_fit, 1200)
    run_case("rmcsa_jpn12_random", "RMCSA-v0", jpn,
             dict(c3, seed=42, load=60, num_spectrum_resources=64, num_spatial_resources=12, worst_xt=None,
                  allow_rejection=True, episode_length=200), "random", 1500, 5)


if __name__ == "__main__":
    main()
