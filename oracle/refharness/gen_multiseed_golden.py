"""Multi-seed hash goldens from the UNMODIFIED reference (TEST INFRASTRUCTURE).

    python oracle/refharness/gen_multiseed_golden.py [case ...]     # writes tests/golden/ms_*.npz

SURVEY.md 7.3 / 8c ask for >= 64 seeds x >= 10 000 steps per env id.  Storing such traces is out of the question
(15 MB per config), and it is not needed: the reference's request stream is a pure function of its seed (CPython's
random.Random; tests/ref_trace.py regenerates it and is pinned by the stored traces), so a fixture holds, per seed,
only a rolling 64-bit hash of everything a step decides - (accepted, path, core, first slot, slots, reward, done) -
sampled every CHECK steps, plus the final counters and a checksum of the final spectrum.  The reference env is driven
by its own heuristic, one multiprocessing.Process per group of seeds.
"""
import multiprocessing as mp
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

DEEPRMSA_NODE_PROBS = [0.01801802, 0.04004004, 0.05305305, 0.01901902, 0.04504505, 0.02402402, 0.06706707,
                       0.08908909, 0.13813814, 0.12212212, 0.07607608, 0.12012012, 0.01901902, 0.16916917]
NSF, US, JPN = ("nsfnet_chen_5-paths_6-modulations.h5", "us14_3-paths_6-modulations.h5", "jpn12_5-paths_6-modulations.h5")
CHECK = 500
# name -> env id, topology, kwargs (without seed), heuristic, seeds, steps
CASES = {
    "ms_rmsa_nsfnet_sapff": ("RMSA-v0", NSF, dict(load=50, num_spectrum_resources=100, episode_length=1000), "sapff", 64, 10000),
    "ms_rmsa_nsfnet_llpff": ("RMSA-v0", NSF, dict(load=80, num_spectrum_resources=100, episode_length=300), "llpff", 64, 10000),
    "ms_deeprmsa_nsfnet_kspff": ("DeepRMSA-v0", NSF, dict(j=1, mean_service_holding_time=7.5,
                                                          mean_service_inter_arrival_time=1 / 12.0, num_spectrum_resources=100,
                                                          episode_length=50, node_request_probabilities=DEEPRMSA_NODE_PROBS),
                                 "kspff", 64, 10000),
    "ms_rmcsa_jpn12_heur": ("RMCSA-v0", JPN, dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                                  num_spatial_resources=7, worst_xt=-84.7, episode_length=1000), "rmcsa", 64, 10000),
    "ms_rmsa_us14_320_full": ("RMSA-v0", US, dict(load=400, mean_service_holding_time=25, num_spectrum_resources=320,
                                                  episode_length=700), "sapff", 64, 10000),  # load >> capacity: full network
    "ms_phy_us14_bmfa_cut_320ch": ("PhyRMSA-v0", US, dict(load=1400, mean_service_holding_time=25, episode_length=200,
                                                          number_spectrum_channels=106, number_spectrum_channels_s_band=108,
                                                          num_spectrum_resources=64, bit_rate_selection="discrete",
                                                          grooming=False), "bmfa", 16, 4000),  # the C4 bench shape; steady state (and blocking) from ~2 000 steps on
}


def fold(h, *vals):
    for v in vals:
        h = (h * 1000003 + (int(v) & 0xFFFFFFFFFFFFFFFF)) & 0xFFFFFFFFFFFFFFFF
    return h


def worker(name, seeds, out):
    import refenv

    gym = refenv.install()
    from optical_rl_gym.envs import deeprmsa_env, phy_rmsa_env, rmcsa_env, rmsa_env

    env_id, topo_file, kw, pol, _, steps = CASES[name]
    res = {}
    for seed in seeds:
        topo = refenv.load_topology(topo_file)
        kws = dict(kw, seed=seed)
        if env_id == "PhyRMSA-v0":
            from scipy.io import loadmat

            d = loadmat(os.path.join(refenv.REFERENCE_ROOT, "examples", "phy_frag_rmsa", "inputs",
                                     "Results_K3SP_FRP_SLC_CBG_USB14.mat"))["Results_K3SP_FRP_SLC_CBG_USB14"][0][0]
            cols = np.concatenate([np.arange(268), np.arange(268 - 52, 268)])  # 320 channels, as bench.py builds them
            kws.update(connections_detail=d[0], modulation_level=d[1].astype(np.int64)[:, cols], gsnr=d[2][:, cols])
        env = gym.make(env_id, topology=topo, **kws)
        heur = {"sapff": rmsa_env.shortest_available_path_first_fit, "llpff": rmsa_env.least_loaded_path_first_fit,
                "kspff": deeprmsa_env.shortest_available_path_first_fit,
                "rmcsa": rmcsa_env.shortest_available_path_best_modulation_first_core_first_fit,
                "bmfa": phy_rmsa_env.phy_aware_bmfa_rmsa}[pol]
        h, checks = 0, []
        for t in range(steps):
            a = heur(env)
            svc = env.current_service
            if pol == "rmcsa" and len(a) == 3:  # F16
                a = [env.k_paths, len(env.modulation_formats), env.num_spatial_resources, env.num_spectrum_resources]
            o = env.step(a)
            acc = int(bool(svc.accepted))
            if env_id == "PhyRMSA-v0":
                path = svc.path.idp if acc else -1  # physical layer only (grooming=False, BM-FA-Cut)
                first = a[1][0][0] if acc else -1
                n = len(a[1]) if acc else 0
                h = fold(h, acc, path, 0, first, n, int(o[1]), int(bool(o[2])))
            else:
                if acc:
                    pid = [p is svc.path or p == svc.path for p in env.k_shortest_paths[svc.source, svc.destination]].index(True)
                    core = svc.core if env_id == "RMCSA-v0" else 0
                    h = fold(h, 1, pid, core, svc.initial_slot, svc.number_slots, int(o[1]), int(bool(o[2])))
                else:
                    h = fold(h, 0, -1, 0, -1, 0, int(o[1]), int(bool(o[2])))
            if o[2]:
                env.reset()
            if (t + 1) % CHECK == 0:
                checks.append(h)
        key = "available_channels" if env_id == "PhyRMSA-v0" else "available_slots"
        av = np.asarray(env.topology.graph[key], np.int64).ravel()
        csum = int(np.dot(av, np.arange(1, av.size + 1, dtype=np.int64)))
        cnt = [env.services_processed, env.services_accepted, env.episode_services_processed, env.episode_services_accepted,
               env.bit_rate_requested, env.bit_rate_provisioned, env.episode_bit_rate_requested, env.episode_bit_rate_provisioned]
        res[seed] = (checks, cnt, csum)
    out.put(res)


def run(name):
    env_id, topo_file, kw, pol, n_seeds, steps = CASES[name]
    seeds = [1000 + i for i in range(n_seeds)]
    procs = min(os.cpu_count() or 1, n_seeds)
    q = mp.Queue()
    ps = [mp.Process(target=worker, args=(name, seeds[i::procs], q)) for i in range(procs)]
    for p in ps:
        p.start()
    res = {}
    for _ in ps:
        res.update(q.get())
    for p in ps:
        p.join()
    out = dict(env_id=np.array(env_id), topo_file=np.array(topo_file), heuristic=np.array(pol), seeds=np.array(seeds, np.int64),
               steps=np.int64(steps), check=np.int64(CHECK),
               hashes=np.array([res[s][0] for s in seeds], np.uint64),
               counters=np.array([res[s][1] for s in seeds], np.int64), final_checksum=np.array([res[s][2] for s in seeds], np.int64),
               kwargs_keys=np.array(sorted(k for k, v in kw.items() if v is not None)))
    for k, v in kw.items():
        if v is not None:
            out["kw_" + k] = np.asarray(v)
    np.savez_compressed(os.path.join(REPO, "tests", "golden", name + ".npz"), **out)
    acc = out["counters"][:, 1].sum() / out["counters"][:, 0].sum()
    print(f"{name}: {n_seeds} seeds x {steps} steps, accepted {acc:.3f}", flush=True)


if __name__ == "__main__":
    for nm in (sys.argv[1:] or list(CASES)):
        run(nm)
