"""Times the UNMODIFIED Python reference on bench.py's workloads (TEST / MEASUREMENT INFRASTRUCTURE).

    python oracle/refharness/time_reference.py [seconds per leg] [workload ...]   # writes profiles/python_reference.json

The reference cannot travel to the GPU box (no vendoring), so this runs in the build container: for every bench
workload the reference env (gym.make under oracle/refharness/stubs) is driven by the reference's own heuristic -
the pattern of tests/test_rmsa_threads_us.py:130-256: one multiprocessing.Process per env, no communication -
(a) in one process and (b) in os.cpu_count() processes, each for `seconds` of wall time after a warm-up; the result is
steps / wall second.  The QoT-gate workloads call the reference's own calculate_osnr for every GSNR through the
harness of gen_qot_golden.py (the reference ships no env that calls it).
bench.py attaches the numbers to its lines as cpu_baseline.python_reference.
"""
import json
import multiprocessing as mp
import os
import platform
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)

DEEPRMSA_NODE_PROBS = [0.01801802, 0.04004004, 0.05305305, 0.01901902, 0.04504505, 0.02402402, 0.06706707,
                       0.08908909, 0.13813814, 0.12212212, 0.07607608, 0.12012012, 0.01901902, 0.16916917]
NSF, US, JPN, SPN = ("nsfnet_chen_5-paths_6-modulations.h5", "us14_3-paths_6-modulations.h5",
                     "jpn12_5-paths_6-modulations.h5", "spn_3-paths_6-modulations.h5")
# the kwargs of bench.py's WORKLOADS, as the reference's constructors take them
CONFIGS = {
    "c1_rmsa_nsfnet_100": ("RMSA-v0", NSF, dict(load=50, num_spectrum_resources=100, episode_length=1000), "sapff"),
    "c2_deeprmsa_nsfnet_100": ("DeepRMSA-v0", NSF, dict(j=1, mean_service_holding_time=7.5,
                                                         mean_service_inter_arrival_time=1 / 12.0, num_spectrum_resources=100,
                                                         episode_length=50, node_request_probabilities=DEEPRMSA_NODE_PROBS), "kspff"),
    "c3_rmcsa_jpn12_320": ("RMCSA-v0", JPN, dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                                 num_spatial_resources=7, worst_xt=-84.7, episode_length=1000), "rmcsa"),
    "c3_rmcsa_jpn12_320_gn": ("RMCSA-v0", JPN, dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                                    num_spatial_resources=7, worst_xt=-84.7, episode_length=1000), "gn"),
    "c4_phyrmsa_us14_320": ("PhyRMSA-v0", US, dict(load=1400, mean_service_holding_time=25, episode_length=200,
                                                   number_spectrum_channels=106, number_spectrum_channels_s_band=108, num_spectrum_resources=64,
                                                   bit_rate_selection="discrete", grooming=False), "bmfa"),
    "c5_rmsa_spn_100": ("RMSA-v0", SPN, dict(load=100, mean_service_holding_time=25, num_spectrum_resources=100,
                                             episode_length=1000), "sapff"),
    "c5_rmsa_spn_320": ("RMSA-v0", SPN, dict(load=400, mean_service_holding_time=25, num_spectrum_resources=320,
                                             episode_length=1000), "sapff"),
    "c6_rmsa_us14_320_gn": ("RMSA-v0", US, dict(load=100, mean_service_holding_time=25, num_spectrum_resources=320,
                                                episode_length=1000), "gn"),
}


def worker(name, seed, seconds, warm_steps, out):
    try:
        _worker(name, seed, seconds, warm_steps, out)
    except BaseException as ex:  # a dead worker must not hang the parent
        out.put((0, 1.0))
        raise ex


def _worker(name, seed, seconds, warm_steps, out):
    import numpy as np

    env_id, topo_file, kw, pol = CONFIGS[name]
    kw = dict(kw, seed=seed)
    if pol == "gn":
        import gen_qot_golden as Q  # installs the stubs, injects Span / Link, imports calculate_osnr

        GG, gym, refenv = Q.GG, Q.gym, Q.refenv
    else:
        import refenv

        gym = refenv.install()
    from optical_rl_gym.envs import deeprmsa_env, phy_rmsa_env, rmcsa_env, rmsa_env

    topo = refenv.load_topology(topo_file)
    if env_id == "PhyRMSA-v0":
        from scipy.io import loadmat

        d = loadmat(os.path.join(refenv.REFERENCE_ROOT, "examples", "phy_frag_rmsa", "inputs",
                                 "Results_K3SP_FRP_SLC_CBG_USB14.mat"))["Results_K3SP_FRP_SLC_CBG_USB14"][0][0]
        cols = np.concatenate([np.arange(268), np.arange(268 - 52, 268)])  # 320 channels, as bench.py builds them
        kw.update(connections_detail=d[0], modulation_level=d[1].astype(np.int64)[:, cols], gsnr=d[2][:, cols])
    env = gym.make(env_id, topology=topo, **kw)
    if pol == "gn":
        from optical_rl_gym_qot_aware_b200 import spec as S
        from optical_rl_gym_qot_aware_b200.tables import tables_from_topology

        kind = 0 if env_id == "RMSA-v0" else 2
        gate = Q.Gate(env, kind, S.gsnr_params(tables_from_topology(topo)))
        gate.scene = Q.AdaptiveScene(env, gate.scene.gp, kind)

        def one():
            if kind == 0:
                a = Q.rmsa_heuristic(env, gate, True)
                return Q.rmsa_gated_step(env, gate, True, a)[0]
            a = Q.rmcsa_heuristic(env, gate, True)
            return Q.rmcsa_gated_step(env, gate, a)
    else:
        heur = {"sapff": rmsa_env.shortest_available_path_first_fit,
                "kspff": deeprmsa_env.shortest_available_path_first_fit,
                "rmcsa": rmcsa_env.shortest_available_path_best_modulation_first_core_first_fit,
                "bmfa": phy_rmsa_env.phy_aware_bmfa_rmsa}[pol]

        def one():
            a = heur(env)
            if pol == "rmcsa" and len(a) == 3:  # F16
                a = [env.k_paths, len(env.modulation_formats), env.num_spatial_resources, env.num_spectrum_resources]
            return env.step(a)

    for _ in range(warm_steps):
        if one()[2]:
            env.reset()
    steps, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        if one()[2]:
            env.reset()
        steps += 1
    out.put((steps, time.perf_counter() - t0))


def run(name, procs, seconds, warm_steps):
    q = mp.Queue()
    ps = [mp.Process(target=worker, args=(name, 10 + i, seconds, warm_steps, q)) for i in range(procs)]
    t0 = time.perf_counter()
    for p in ps:
        p.start()
    res = [q.get() for _ in ps]
    for p in ps:
        p.join()
    return sum(s / t for s, t in res), sum(s for s, _ in res), time.perf_counter() - t0


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 30.0
    names = sys.argv[2:] or list(CONFIGS)
    ncpu = os.cpu_count() or 1
    cpu = platform.processor() or "unknown"
    try:
        cpu = [ln.split(":", 1)[1].strip() for ln in open("/proc/cpuinfo") if ln.startswith("model name")][0]
    except Exception:
        pass
    path = os.path.join(REPO, "profiles", "python_reference.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    for name in names:
        warm = {"gn": 300, "bmfa": 200}.get(CONFIGS[name][3], 600)
        v1, s1, _ = run(name, 1, seconds, warm)
        vn, sn, _ = run(name, ncpu, seconds, warm)
        out[name] = {"single_process": v1, "all_cores": vn, "cores": ncpu, "cpu_model": cpu, "unit": "env-steps/s",
                     "seconds_per_leg": seconds, "steps_timed": [s1, sn], "where": "build container",
                     "what": "unmodified reference under the gym stubs, its own heuristic, one multiprocessing.Process per env"
                             + (" ; every GSNR of the gate through the reference's calculate_osnr"
                                if CONFIGS[name][3] == "gn" else "")}
        print(name, out[name], flush=True)
        json.dump(out, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
