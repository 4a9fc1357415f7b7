"""PhyRMSA-v0 golden traces from the UNMODIFIED reference (TEST INFRASTRUCTURE).

    python oracle/refharness/gen_phy_golden.py      # writes tests/golden/phy_*.npz

Drives reference PhyRMSAEnv instances (grooming=False, no defragmentation) on the shipped
topologies and .mat QoT tables (modulation_level cast to int64 first) with
the reference's own heuristics and records per step: the pending request, the action
(path, channel numbers), accepted/reward/done, counters, and the info entries
number_cuts_total / rss_total_metric / total_path_length / avrage_gsnr / average_mod_level /
path_index / physical_paths; at the end available_channels.

BM-FA-Cut (phy_aware_bmfa_rmsa) runs exactly as shipped.  SAP-FF, SAP-BM and BM-FF always
call use_existing_channels (grooming, a later scope row); for them the harness rebinds that
module attribute to a stub returning (-3, []) so that only their physical-layer logic runs —
the reference's source files stay untouched.  The fixture carries the [rows, channels, k]
slices of the QoT tables it used, so the tests need neither scipy nor the reference.
"""
import os
import sys

import numpy as np
from scipy.io import loadmat

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)

import refenv  # noqa: E402

gym = refenv.install()

from optical_rl_gym.envs import phy_rmsa_env as ref  # noqa: E402

GOLDEN = os.path.join(REPO, "tests", "golden")
REQ_DTYPE = np.dtype([("arrival", "<f8"), ("holding", "<f8"), ("bit_rate", "<i4"), ("src", "<i2"), ("dst", "<i2")])
INPUTS = os.path.join(refenv.REFERENCE_ROOT, "examples", "phy_frag_rmsa", "inputs")
ADIM = 9


def load_tables(key):
    d = loadmat(os.path.join(INPUTS, f"{key}.mat"))[key][0][0]
    return d[0], d[1].astype(np.int64), d[2]  # connections, modulation (int64: F15), gsnr


def req_of(env):
    s = env.current_service
    return (float(s.arrival_time), float(s.holding_time), int(s.bit_rate), int(s.source_id), int(s.destination_id))


def counters_of(env):
    return np.array([env.services_processed, env.services_accepted, env.episode_services_processed,
                     env.episode_services_accepted, env.bit_rate_requested, env.bit_rate_provisioned,
                     env.episode_bit_rate_requested, env.episode_bit_rate_provisioned], np.int64)


ONLY = set(sys.argv[1:])  # optional: names of the cases to (re)generate


def run(name, topo_file, mat_key, heuristic, no_groom_stub, n_steps, **kw):
    if ONLY and name not in ONLY:
        return
    topo = refenv.load_topology(topo_file)
    conn, mod, gs = load_tables(mat_key)
    kwargs = dict(num_spectrum_resources=64, bit_rate_selection="discrete", grooming=False,
                  mean_service_holding_time=25)
    kwargs.update(number_spectrum_channels=80, number_spectrum_channels_s_band=108)
    kwargs.update(kw)
    env = gym.make("PhyRMSA-v0", topology=topo, modulation_level=mod, connections_detail=conn, gsnr=gs, **kwargs)
    saved = ref.use_existing_channels
    if no_groom_stub:
        ref.use_existing_channels = lambda env_: (-3, [])
    try:
        k = env.k_paths
        reqs = np.zeros(n_steps + 1, REQ_DTYPE)
        actions = np.full((n_steps, ADIM), -1, np.int32)
        accepted = np.zeros(n_steps, np.uint8)
        virtual = np.zeros(n_steps, np.uint8)
        reward = np.zeros(n_steps, np.int32)
        done = np.zeros(n_steps, np.uint8)
        counters = np.zeros((n_steps, 8), np.int64)
        stats = np.zeros((n_steps, 8), np.float64)
        # periodic defragmentation (phy_rmsa_env.py:355-417): cumulative move counters and a checksum of
        # available_channels after every step, so that each reallocation is pinned, not only the final map
        moves = np.zeros((n_steps, 3), np.int64)
        avail_sum = np.zeros(n_steps, np.int64)
        weights = None
        reqs[0] = req_of(env)
        for t in range(n_steps):
            path, chans = heuristic(env)
            assert path <= 10 or not no_groom_stub, "virtual-layer action although use_existing_channels is stubbed"
            virtual[t] = int(path > 10)
            actions[t, 0] = path
            for i, c in enumerate(chans):
                actions[t, 1 + i] = c[0]
            svc = env.current_service
            _, r, d, _, info = env.step((path, chans))
            accepted[t], reward[t], done[t] = int(svc.accepted), r, int(d)
            # episode sums behind the info entries (phy_rmsa_env.py:319-348), read before any reset
            stats[t] = (env.total_path_length_episode, env.total_gsnr_episode, env.total_path_index_episode,
                        env.total_modulation_level_episode, env.channels_accepted_episode,
                        env.physical_services_accepted_episode, info["number_cuts_total"], info["rss_total_metric"])
            reqs[t + 1] = req_of(env)
            moves[t] = (env.counted_moves, env.counted_moves_groom, env.counted_defrag_cycles)
            av = np.asarray(env.topology.graph["available_channels"], np.int64).ravel()
            if weights is None:
                weights = np.arange(1, av.size + 1, dtype=np.int64)
            avail_sum[t] = int(np.dot(av, weights))
            if d:
                env.reset()
            counters[t] = counters_of(env)
        out = dict(env_kind=3, topo_file=topo_file, requests=reqs, actions=actions, accepted=accepted, reward=reward,
                   done=done, counters=counters, stats=stats, virtual=virtual, moves=moves, avail_sum=avail_sum,
                   channel_state=np.int32(0 if no_groom_stub else (1 if kwargs.get("grooming") else 2)),
                   final_avail=np.asarray(env.topology.graph["available_channels"], np.uint8)[None],
                   kwargs_keys=np.array(sorted(kwargs)), tables_key=np.array(mat_key))
        tab = os.path.join(GOLDEN, f"phytab_{mat_key}.npz")
        if not os.path.exists(tab):  # the [rows, channels, k] slices of the shipped QoT tables, once per topology
            np.savez_compressed(
                tab, phy_mod=mod[:, :, :k].astype(np.uint8), phy_gsnr=np.ascontiguousarray(gs[:, :, :k]),
                phy_conn=np.array([[int(np.asarray(conn[r, c]).reshape(-1)[0]) for c in (0, 1)]
                                   for r in range(conn.shape[0])], np.int32))
        for key, v in kwargs.items():
            out["kw_" + key] = np.asarray(v)
        path_ = os.path.join(GOLDEN, name + ".npz")
        np.savez_compressed(path_, **out)
        print(f"{name}: {n_steps} steps, blocking {1 - accepted.mean():.3f}, virtual {virtual.mean():.3f}, episodes {int(done.sum())}, "
              f"cuts {stats[-1, 6]:.3f} rss {stats[-1, 7]:.4f}, {os.path.getsize(path_) / 1024:.0f} KiB")
    finally:
        ref.use_existing_channels = saved


def main():
    us, jpn = "us14_3-paths_6-modulations.h5", "jpn12_3-paths_6-modulations.h5"
    US, JPN = "Results_K3SP_FRP_SLC_CBG_USB14", "Results_K3SP_FRP_SLC_CBG_JPN12"
    small = dict(number_spectrum_channels=30, number_spectrum_channels_s_band=40)   # 100 channels
    tiny = dict(number_spectrum_channels=20, number_spectrum_channels_s_band=24)    # 64 channels
    run("phy_us14_bmfa_cut", us, US, ref.phy_aware_bmfa_rmsa, False, 600, seed=10, load=1400, episode_length=200)
    run("phy_us14_bmfa_cut_100ch", us, US, ref.phy_aware_bmfa_rmsa, False, 1500, seed=15, load=1400,
        episode_length=200, **small)
    run("phy_jpn12_bmfa_cut_64ch", jpn, JPN, ref.phy_aware_bmfa_rmsa, False, 1500, seed=11, load=1120,
        episode_length=205, **tiny)
    run("phy_us14_sapff_100ch", us, US, ref.sapff_rmsa, True, 1500, seed=12, load=1500, episode_length=150, **small)
    run("phy_us14_bmff_100ch", us, US, ref.phy_aware_bmff_rmsa, True, 1500, seed=13, load=1600, episode_length=300,
        **small)
    run("phy_us14_faff_100ch", us, US, ref.phy_aware_faff_rmsa, True, 1500, seed=16, load=1500, episode_length=250,
        **small)
    run("phy_us14_bmfarss_100ch", us, US, ref.phy_aware_bmfa_rss_rmsa, False, 1200, seed=17, load=1450,
        episode_length=220, **small)
    run("phy_jpn12_faffrss_64ch", jpn, JPN, ref.phy_aware_faff_rss_rmsa, True, 1200, seed=18, load=1150,
        episode_length=180, **tiny)
    # virtual layer: env.grooming=True, and the reference's grooming=False behaviour of the heuristics that call
    # use_existing_channels unconditionally (no stub)
    run("phy_us14_groom_bmfa_100ch", us, US, ref.phy_aware_bmfa_rmsa, False, 1500, seed=19, load=1900,
        episode_length=230, grooming=True, **small)
    run("phy_jpn12_groom_faffrss_64ch", jpn, JPN, ref.phy_aware_faff_rss_rmsa, False, 1200, seed=20, load=1500,
        episode_length=170, grooming=True, **tiny)
    run("phy_us14_cstate_sapff_100ch", us, US, ref.sapff_rmsa, False, 1500, seed=21, load=1900, episode_length=210,
        **small)
    # the heuristics of the stubbed fixtures above once more as shipped (no stub: they look at the virtual layer first)
    run("phy_us14_cstate_bmff_100ch", us, US, ref.phy_aware_bmff_rmsa, False, 1500, seed=25, load=1800, episode_length=260,
        **small)
    run("phy_us14_cstate_faff_100ch", us, US, ref.phy_aware_faff_rmsa, False, 1500, seed=26, load=1800, episode_length=270,
        **small)
    run("phy_jpn12_cstate_sapbm_64ch", jpn, JPN, ref.phy_aware_sapbm_rmsa, False, 1500, seed=27, load=1400,
        episode_length=120, **tiny)
    # periodic defragmentation: targets for the round-2 kernel; no engine path consumes them yet
    run("defrag_phy_us14_defrag_cut_100ch", us, US, ref.phy_aware_bmfa_rmsa, False, 1200, seed=22, load=1700,
        episode_length=240, defrag_period=10, number_moves=10, metric="cut", **small)
    run("defrag_phy_us14_defrag_groom_cut_100ch", us, US, ref.phy_aware_bmfa_rmsa, False, 1200, seed=23, load=1900,
        episode_length=240, grooming=True, defrag_period=10, number_moves=10, metric="cut", **small)
    run("defrag_phy_jpn12_defrag_rss_64ch", jpn, JPN, ref.phy_aware_bmfa_rss_rmsa, False, 900, seed=24, load=1300,
        episode_length=190, defrag_period=20, number_moves=5, metric="rss", **tiny)
    run("phy_jpn12_sapbm_64ch", jpn, JPN, ref.phy_aware_sapbm_rmsa, True, 1500, seed=14, load=1300,
        episode_length=100, **tiny)


if __name__ == "__main__":
    main()
