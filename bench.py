#!/usr/bin/env python
"""bench.py — env-steps/sec of the batched stepping engine on B200 (driver contract).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA engine
    python bench.py --impl reference --gpus N --steps K ...   # CPU arm (oracle port on all host threads)

One "step" = one pass of the hot path over one batch of envs: every env takes one gym step with the workload's
policy evaluated on the device.  The default (headline) workload is the configuration BASELINE.json's metric and
target are quoted on: RMSA-v0 on the US topology, 320 slots, GN-model QoT — the GSNR of every candidate lightpath
and of its neighbours decides accept / block and the modulation format (include/orlg.h "QoT gate") — 65 536 envs per
GPU (weak scaling: envs are sharded with no data-path collective; NCCL only all-reduces the episode statistics).
At N = 1 the line also carries a `workloads` block: one record per BASELINE.json config (C1 .. C5).

`value`    device-resident throughput: actions, state and outputs stay in HBM.  The K steps are issued by one C
           call (orlg_timed_steps) as ONE CUDA graph - (L2-evicting memset, event, step launches, event) x K - so no
           host work sits between an event record and the launch it times; value = envs x K / sum of the K
           event intervals, max over ranks.
`e2e`      the same metric through the host-buffer C-ABI call (orlg_step_host): the policy's actions travel
           device -> pinned host -> device, reward / done / allocation / GSNR come back D2H every step; copies and
           host work are inside the timed region.
`roofline` the step's dominant kernel: algorithmic bytes per env-step x envs / its CUDA-event
           duration, against MEASURED_PEAKS.json hbm_gbs.
`cpu_baseline` the C oracle port (oracle/, test infrastructure) stepping a bounded sample of the same workload
           on all host threads; `python_reference` (from profiles/) is the unmodified Python reference timed in the
           build container.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
for _p in (REPO, os.path.join(REPO, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

from optical_rl_gym_qot_aware_b200 import spec as S  # noqa: E402
from optical_rl_gym_qot_aware_b200.tables import builtin_qot_tables, builtin_topology  # noqa: E402

METRIC = "env-steps/sec"

# node probabilities of the reference's DeepRMSA driver (tests/test_deeprmsa.py:30-47)
DEEPRMSA_NODE_PROBS = [0.01801802, 0.04004004, 0.05305305, 0.01901902, 0.04504505, 0.02402402, 0.06706707,
                       0.08908909, 0.13813814, 0.12212212, 0.07607608, 0.12012012, 0.01901902, 0.16916917]

# name -> env id, topology, constructor kwargs, device policy, observation, step form ("hint": orlg_step with the fused
# next-action hint, 1 launch; "auto": orlg_step_auto, gate + step = 2 launches), steps run before any timing,
# SURVEY.md 8(d) bytes per env-step (upper bound for the first-fit workloads), envs in the CPU arm's sample
WORKLOADS = {
    "c6_rmsa_us14_320_gn": dict(
        env="RMSA-v0", topo="us14", hid=S.H_SAP_GN_FF, obs=S.OBS_NONE, step="auto", preroll=600, bytes=None, cpu_envs=512,
        envs=262144,  # the target names no batch size; the gate kernel is still filling up at 65 536 envs (profiles/r02b_env_sweep.md)
        kw=dict(load=100, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=1000, qot="gn",
                qot_adaptive=True),
        what="BASELINE target config: 320-slot RMSA on the US topology, GN-model GSNR of the candidate and its "
             "neighbours feeding accept/block and the modulation format"),
    "c1_rmsa_nsfnet_100": dict(
        env="RMSA-v0", topo="nsfnet", hid=S.H_SAP_FF, obs=S.OBS_NONE, step="hint", preroll=600, bytes=609, cpu_envs=8192,
        kw=dict(load=50, num_spectrum_resources=100, episode_length=1000), what="BASELINE configs[0]"),
    "c2_deeprmsa_nsfnet_100": dict(
        env="DeepRMSA-v0", topo="nsfnet", hid=S.H_SAP_FF, obs=S.OBS_F32, step="hint", preroll=600, bytes=825, cpu_envs=8192,
        kw=dict(j=1, mean_service_holding_time=7.5, mean_service_inter_arrival_time=1 / 12.0, num_spectrum_resources=100,
                episode_length=50, node_request_probabilities=DEEPRMSA_NODE_PROBS), what="BASELINE configs[1]"),
    "c3_rmcsa_jpn12_320_gn": dict(
        env="RMCSA-v0", topo="jpn12_k5", hid=S.H_SAP_GN_FF, obs=S.OBS_NONE, step="auto", preroll=1500, bytes=None,
        cpu_envs=256,
        kw=dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320, num_spatial_resources=7,
                worst_xt=-84.7, episode_length=1000, qot="gn", qot_adaptive=True),
        what="BASELINE configs[2]: QoT-aware RMCSA with GN-model GSNR on JPN12, 320 slots, modulation-adaptive"),
    "c3_rmcsa_jpn12_320": dict(
        env="RMCSA-v0", topo="jpn12_k5", hid=S.H_SAP_BM_FC_FF, obs=S.OBS_NONE, step="hint", preroll=1500, bytes=6454,
        cpu_envs=4096,
        kw=dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320, num_spatial_resources=7,
                worst_xt=-84.7, episode_length=1000), what="configs[2] without the GN model (closed-form QoT gate only)"),
    "c4_phyrmsa_us14_320": dict(
        env="PhyRMSA-v0", topo="us14", hid=S.H_PHY_BM_FA_CUT, obs=S.OBS_PHY_CUTS, step="hint", preroll=6000, bytes=1426,
        cpu_envs=256,
        kw=dict(load=1400, mean_service_holding_time=25, episode_length=200, number_spectrum_channels=106, number_spectrum_channels_s_band=108),
        what="BASELINE configs[3]: phy-frag RMSA on US14, 320 channels, BM-FA-Cut, cuts reward"),
    "c5_rmsa_spn_100": dict(
        env="RMSA-v0", topo="spn", hid=S.H_SAP_FF, obs=S.OBS_NONE, step="hint", preroll=600, bytes=517, cpu_envs=8192,
        kw=dict(load=100, mean_service_holding_time=25, num_spectrum_resources=100, episode_length=1000),
        what="BASELINE configs[4] (SPN, 100 slots)"),
    "c5_rmsa_spn_320": dict(
        env="RMSA-v0", topo="spn", hid=S.H_SAP_FF, obs=S.OBS_NONE, step="hint", preroll=600, bytes=1039, cpu_envs=4096,
        kw=dict(load=400, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=1000),
        what="BASELINE configs[4] (SPN, 320 slots)"),
}
WORKLOADS["c6_rmsa_us14_320_gn_64k"] = dict(WORKLOADS["c6_rmsa_us14_320_gn"], envs=65536, ref="c6_rmsa_us14_320_gn",
                                             what="the headline workload at 65 536 envs per GPU (the batch of BASELINE configs[1])")
HEADLINE = "c6_rmsa_us14_320_gn"
DEFAULT_ENVS = 65536  # BASELINE configs[1]; a workload may name its own (`envs`)
BLOCK = ["c6_rmsa_us14_320_gn_64k", "c1_rmsa_nsfnet_100", "c2_deeprmsa_nsfnet_100", "c3_rmcsa_jpn12_320_gn", "c3_rmcsa_jpn12_320",
         "c4_phyrmsa_us14_320", "c5_rmsa_spn_100", "c5_rmsa_spn_320"]
DATA = "synthetic (Philox request streams, seed 0x5eed, keyed by global env id)"
OBS_NAME = {0: "none", 1: "f64", 2: "f32", 3: "matrix_u8", 4: "phy cuts+rss f64", 5: "phy cuts f64"}


def make_workload(name, seed=0x5EED):
    w = WORKLOADS[name]
    kw = dict(w["kw"])
    if w["env"] == "PhyRMSA-v0":
        # the shipped QoT tables have 268 channels; 320 = them + the S-band tail replicated
        mod, gsnr, conn = builtin_qot_tables()
        cols = np.concatenate([np.arange(268), np.arange(268 - 52, 268)])
        kw.update(modulation_level=mod[:, cols], gsnr=gsnr[:, cols], connections_detail=conn)
    spec = S.make_spec(w["env"], builtin_topology(w["topo"]), seed=seed, request_source=S.SRC_PHILOX, auto_reset=True,
                       obs_kind=w["obs"], **kw)
    return spec, w


def measured_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock and throttle reasons sampled every 5 ms during the timed region (NVML; the nvidia-smi -lms loop of
    the recipe starts too slowly for a sub-second region)."""

    def __init__(self, index):
        self.index, self.sm, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self.thread = None

    def start(self):
        try:
            import pynvml as N

            N.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")  # NVML enumerates every GPU of the box
            idx = self.index
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                idx = int(vis.split(",")[self.index])
            h = N.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM))
            bits = {"hw_slowdown": N.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": N.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": N.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": N.nvmlClocksThrottleReasonSwPowerCap}

            def pump():
                while not self._stop.is_set():
                    try:
                        self.sm.append(float(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)))
                        r = N.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        for nm, bit in bits.items():
                            if r & bit:
                                self.reasons.add(nm)
                    except Exception:
                        pass
                    time.sleep(0.005)

            self.thread = threading.Thread(target=pump, daemon=True)
            self.thread.start()
        except Exception as ex:  # no NVML: report it instead of inventing clocks
            self.reasons.add(f"nvml unavailable: {type(ex).__name__}")

    def stop(self):
        self._stop.set()
        if self.thread is not None:
            self.thread.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


_KSPFF_LUT = {}
_KSPFF_WS = {}


def host_policy_kspff(obs_fm, n_nodes, k, j, out=None):
    """K-SP-FF (deeprmsa_env.py:146-155) evaluated on the host from the observation: path p has a fitting block iff
    its first block-length feature is not -1.  obs_fm: [obs_dim, n_env] (feature-major host buffer).  The k
    availability bits of an env are packed into one byte (Horner) and mapped through a table."""
    assert k <= 8
    rowlen = 2 * j + 3
    base = 1 + 2 * n_nodes
    lut = _KSPFF_LUT.get((k, j))
    if lut is None:
        lut = np.array([next((p * j for p in range(k) if c >> p & 1), k * j) for c in range(1 << k)], np.int32)
        _KSPFF_LUT[(k, j)] = lut
    n = obs_fm.shape[1]
    ws = _KSPFF_WS.get(n)
    if ws is None:  # per batch size: bit scratch, packed code, index scratch (no allocation in the loop)
        ws = _KSPFF_WS[n] = (np.empty(n, np.bool_), np.empty(n, np.uint8), np.empty(n, np.intp))
    bit, code, idx = ws
    np.not_equal(obs_fm[base + (k - 1) * rowlen + 1], -1.0, out=code.view(np.bool_))
    for p in range(k - 2, -1, -1):
        np.not_equal(obs_fm[base + p * rowlen + 1], -1.0, out=bit)
        np.add(code, code, out=code)
        np.bitwise_or(code, bit.view(np.uint8), out=code)
    np.copyto(idx, code)
    if out is None:
        return lut.take(idx).reshape(-1, 1)
    np.take(lut, idx, out=out.reshape(-1), mode="clip")
    return out


def cpu_arm(spec, w, n_sample, threads, seconds=None, steps=None, warmup=0):
    """The oracle port on host threads over a persistent pool of n_sample envs.  Timed either for `seconds` (50-step
    batches) or for exactly `steps` steps of every env in ONE pool call (threads are spawned once per call, so a
    per-step call would time pthread_create as much as the envs).  Returns (env-steps/s, ms per step, steps, blocking)."""
    import oracle_binding as OB

    pool = OB.OraclePool(spec, 0, n_sample)
    pool.rollout(w["preroll"], w["hid"], threads)
    if warmup:
        pool.rollout(warmup, w["hid"], threads)
    done_steps, passes = 0, 0
    t0 = time.perf_counter()
    if steps is not None:
        done_steps += pool.rollout(steps, w["hid"], threads)
        passes = steps
    else:
        while time.perf_counter() - t0 < seconds:
            done_steps += pool.rollout(50, w["hid"], threads)
            passes += 50
    dt = time.perf_counter() - t0
    cnt = pool.counters()
    blocking = 1.0 - cnt[:, 1].sum() / max(1, cnt[:, 0].sum())
    pool.close()
    return done_steps / dt, dt * 1e3 / max(1, passes), passes, float(blocking)


def envs_of(name, args):
    return int(args.envs or WORKLOADS[name].get("envs", DEFAULT_ENVS))


def python_reference(name):
    """The unmodified Python reference timed in the build container (oracle/refharness/time_reference.py writes
    profiles/python_reference.json; it cannot travel to the GPU box)."""
    p = os.path.join(REPO, "profiles", "python_reference.json")
    try:
        return json.load(open(p)).get(WORKLOADS.get(name, {}).get("ref", name))
    except Exception:
        return None


def first_fit_bytes(spec, eng, act, upper):
    """Algorithmic bytes per env-step of the first-fit workloads, from the decisions actually taken: shortest-
    available-path first-fit stops at the first path (RMCSA: first (path, core)) that fits, so the rows it has to
    read follow from the action; SURVEY 8(d)'s figures assume all k paths (x all cores)."""
    n_env = eng.num_envs
    a = act.cpu().numpy()
    rq, t = eng.requests(), spec.tables
    k, cores = t.k_paths, spec.num_cores
    hops = t.path_hops[rq["src"].astype(np.int64) * t.n_nodes + rq["dst"]].astype(np.int64)  # [n, k]
    cum = np.concatenate([np.zeros((n_env, 1), np.int64), np.cumsum(hops, axis=1)], axis=1)
    idx = np.arange(n_env)
    ok = (a[:, 0] >= 0) & (a[:, 0] < k)
    pth = np.where(ok, a[:, 0], 0)
    core = a[:, 2] if spec.env_kind == S.ENV_RMCSA else np.zeros(n_env, a.dtype)
    rows_read = np.where(ok, cum[idx, pth] * cores + hops[idx, pth] * (core + 1), cum[:, k] * cores)
    cnt = eng.counters()
    accept = float(cnt[:, 1].sum()) / max(1.0, float(cnt[:, 0].sum()))
    rows_rw = accept * 3.0 * float(np.where(ok, hops[idx, pth], 0).mean())  # write the chosen rows, read+write a release
    words = (spec.num_slots + 31) // 32
    b = int(round(4 * words * (float(rows_read.mean()) + rows_rw) + 168))
    note = (f"rows of the paths{' x cores' if cores > 1 else ''} first-fit examined for the decisions taken "
            f"({float(rows_read.mean()):.2f} rows read per step on average), chosen rows written, one release per accepted "
            f"service, 168 B fixed (SURVEY.md 8(d) upper bound: {upper} B)")
    return b, note


def qot_gate_bytes(spec, eng, act, sample=4096):
    """Algorithmic bytes per env-step of the QoT gate kernel, from the state of a sample of
    envs: the payload of every running service (8 B each: membership test), the occupancy rows of the k candidate
    paths, ONE gate evaluation - the running sums of every neighbour (16 B per hop) - and, per accepted request, its
    own row written plus a read-modify-write of the neighbours' sums on the shared links (32 B each), the same again
    when it is released; 96 B fixed (request, verdict, GSNR pair, free-row stack, counters)."""
    t = spec.tables
    n = min(sample, eng.num_envs)
    svc, cnt = eng.services(max_rows=512)
    svc, cnt = svc[:n].cpu().numpy().astype(np.int64), cnt[:n].cpu().numpy()
    a = act[:n].cpu().numpy()
    rq = eng.requests()[:n]
    k = t.k_paths
    masks = np.zeros(t.n_nodes * t.n_nodes * k, np.uint64)
    for i in range(masks.size):
        for h in range(int(t.path_hops.reshape(-1)[i])):
            masks[i] |= np.uint64(1) << np.uint64(int(t.path_links.reshape(-1, t.max_hops)[i, h]))
    hops_flat = t.path_hops.reshape(-1).astype(np.int64)
    pair = rq["src"].astype(np.int64) * t.n_nodes + rq["dst"]
    ok = (a[:, 0] >= 0) & (a[:, 0] < k)
    cidx = pair * k + np.where(ok, a[:, 0], 0)
    ccore = a[:, 2] if spec.env_kind == S.ENV_RMCSA else np.zeros(n, np.int64)
    sidx = svc[:, :, 0] * k + svc[:, :, 1]
    valid = np.arange(svc.shape[1])[None, :] < np.minimum(cnt, svc.shape[1])[:, None]
    common = masks[sidx] & masks[cidx][:, None]
    nb = valid & (common != 0) & (svc[:, :, 2] == ccore[:, None]) & ok[:, None]
    shared = np.zeros(common.shape, np.int64)
    c = common.copy()
    while c.any():
        shared += (c & np.uint64(1)).astype(np.int64)
        c >>= np.uint64(1)
    nb_rows = float((nb * hops_flat[sidx]).sum(1).mean())       # neighbour hops read per gate evaluation
    nb_shared = float((nb * shared).sum(1).mean())               # shared (neighbour, link) pairs
    cn = eng.counters()
    accept = float(cn[:, 1].sum()) / max(1.0, float(cn[:, 0].sum()))
    hk = float(t.path_hops[pair].sum(1).mean())
    words = (spec.num_slots + 31) // 32
    q = float(cnt.mean())
    chops = float(np.where(ok, hops_flat[cidx], 0).mean())
    b = 8 * q + 4 * words * hk * (spec.num_cores if spec.env_kind == S.ENV_RMCSA else 1) + 16 * nb_rows \
        + accept * (16 * chops + 2 * 32 * nb_shared) + 96
    note = (f"{q:.0f} running services x 8 B, rows of the k candidate paths, one gate evaluation = {nb_rows:.1f} neighbour "
            f"hops x 16 B, accepted ({accept:.2f}): own row + {nb_shared:.1f} shared-link sums read-modify-written at "
            "provision and again at release, 96 B fixed (DESIGN.md 6f)")
    return int(round(b)), note


def run_e2e(spec, w, n_env, parts, steps, local, rank, heap_capacity, block_wait, barrier):
    """The metric through the host-buffer C-ABI calls, `parts` VecEnv shards on their own streams (step_async /
    step_wait): per step and shard, the policy's action is produced (DeepRMSA: numpy K-SP-FF on the host observation;
    otherwise the device heuristic, copied to pinned host memory), orlg_step_host_async copies it H2D, steps, and
    copies reward / done / allocation (/ observation / GSNR pair) back D2H.  Returns a dict."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    dev = torch.device("cuda", local)
    hid, t = w["hid"], spec.tables
    obs_policy = spec.env_kind == S.ENV_DEEPRMSA and spec.obs_dim > 0
    gate = bool(spec.qot_gate)
    nh = n_env // parts
    odt = {S.OBS_F64: torch.float64, S.OBS_F32: torch.float32, S.OBS_MATRIX_U8: torch.uint8, S.OBS_PHY_INFO: torch.float64,
           S.OBS_PHY_CUTS: torch.float64}.get(spec.obs_kind)
    shards = []
    for hi in range(parts):
        he = Engine(spec, nh, device=local, global_env_offset=rank * n_env + hi * nh, heap_capacity=heap_capacity)
        he.set_host_wait(block_wait)
        he.reset(True)
        he.rollout(hid, w["preroll"])
        h = dict(e=he, st=torch.cuda.Stream(device=dev),
                 d_act=torch.zeros((nh, he.action_dim), dtype=torch.int32, device=dev),
                 act=torch.zeros((nh, he.action_dim), dtype=torch.int32).pin_memory(),
                 rew=torch.zeros(nh, dtype=torch.int32).pin_memory(), done=torch.zeros(nh, dtype=torch.uint8).pin_memory(),
                 alloc=torch.zeros((nh, 4), dtype=torch.int32).pin_memory(), obs=None, gs=None, d_gs=None)
        if he.obs_dim:
            dobs = he.new_obs()
            he.observation(dobs)
            h["obs"] = torch.zeros((he.obs_dim, nh), dtype=odt).pin_memory()  # feature-major, as the device buffer
            h["obs"].copy_(dobs.t())
        if gate:
            h["d_gs"] = torch.zeros((nh, 2), dtype=torch.float64, device=dev)
            h["gs"] = torch.zeros((nh, 2), dtype=torch.float64).pin_memory()
        shards.append(h)
    torch.cuda.synchronize()
    policy_s = [0.0]

    def propose(h):
        """Device policies: the heuristic launch and the copy of its proposal to pinned memory, not waited for."""
        if not obs_policy:
            with torch.cuda.stream(h["st"]):
                h["e"].heuristic(hid, h["d_act"])
                h["act"].copy_(h["d_act"], non_blocking=True)

    def submit(h):
        he, st = h["e"], h["st"]
        p0 = time.perf_counter()
        if obs_policy:
            host_policy_kspff(h["obs"].numpy(), t.n_nodes, t.k_paths, spec.j, out=h["act"].numpy())
        else:
            st.synchronize()  # the host "policy" has the proposal in hand before it answers
        policy_s[0] += time.perf_counter() - p0
        he.step_host_async(h["act"], h["rew"], h["done"], h["obs"], h["alloc"], stream=st)
        if gate:
            with torch.cuda.stream(st):
                he.qot_last(out=h["d_gs"])
                h["gs"].copy_(h["d_gs"], non_blocking=True)

    def run(nsteps):
        # two sweeps over the shards per step, so that the shards' proposal launches run side by side on the GPU and
        # so do their steps, instead of the host waiting for one shard's small launch at a time
        for h in shards:
            propose(h)
        for it in range(nsteps):
            for h in shards:
                submit(h)
            for h in shards:
                h["e"].step_host_wait(stream=h["st"])
                if it + 1 < nsteps:
                    propose(h)

    run(min(5, steps))
    barrier()
    policy_s[0] = 0.0
    w0 = time.perf_counter()
    run(steps)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - w0) * 1e3
    barrier()
    h0 = shards[0]
    per = lambda x: 0 if x is None else x.numel() * x.element_size()  # noqa: E731
    h2d = parts * per(h0["act"])
    d2h = parts * (per(h0["rew"]) + per(h0["done"]) + per(h0["alloc"]) + per(h0["obs"]) + per(h0["gs"])
                   + (0 if obs_policy else per(h0["act"])))
    for h in shards:
        if int(h["e"].counters()[:, 10].max()) != 0:
            raise RuntimeError("device fault bits set in the e2e run")
        h["e"].close()
    return dict(ms=ms, steps=steps, h2d=h2d, d2h=d2h, policy_ms=policy_s[0] * 1e3 / steps,
                mode=(f"{parts} VecEnv shard(s) of {nh} envs on their own streams, orlg_step_host_async / _wait"
                      + (" (a shard's host work overlaps the copies and kernels of the others; per step one sweep"
                         " submits every shard's step, one waits for it and launches the next proposal)" if parts > 1 else "")),
                policy=("numpy K-SP-FF from the host observation" if obs_policy
                        else "device heuristic -> pinned host (D2H) -> orlg_step_host (H2D)"))


def measure(name, args, ctx, full):
    """One workload on this rank's GPU.  full: the headline (clock sampling, pipelined e2e, extras); else the compact
    record of the `workloads` block.  Returns a dict of this rank's numbers (times in ms)."""
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    local, rank, world, barrier = ctx["local"], ctx["rank"], ctx["world"], ctx["barrier"]
    dev = torch.device("cuda", local)
    spec, w = make_workload(name)
    hid, auto = w["hid"], w["step"] == "auto"
    n_env, K, W = envs_of(name, args), args.steps, args.warmup
    eng = Engine(spec, n_env, device=local, global_env_offset=rank * n_env, heap_capacity=args.heap_capacity)
    eng.reset(True)
    eng.rollout(hid, w["preroll"])
    act = torch.zeros((n_env, eng.action_dim), dtype=torch.int32, device=dev)
    rew = torch.zeros(n_env, dtype=torch.int32, device=dev)
    done = torch.zeros(n_env, dtype=torch.uint8, device=dev)
    obs = eng.new_obs() if eng.obs_dim else None
    flush = ctx["flush"]
    if not auto:  # every step also writes the heuristic's action for the next request into `act`
        eng.heuristic(hid, act)
        eng.set_next_action_hint(hid, act)

    def timed(nsteps, fl):
        return eng.timed_steps(nsteps, actions=act, reward=rew, done=done, obs=obs, heuristic_id=hid if auto else None,
                               flush=fl, use_graph=not args.no_graph)

    timed(W, flush)
    sampler = ClockSampler(local) if full else None
    barrier()
    if sampler:
        sampler.start()
    l0 = eng.launch_count
    barrier()
    ms, first = timed(K, flush)
    barrier()
    launches = eng.launch_count - l0
    clocks = sampler.stop() if sampler else None
    r = dict(name=name, spec=spec, w=w, step_ms=float(ms.sum()), median_ms=float(np.median(ms)), launches=int(launches),
             clocks=clocks, hbm=eng.hbm_bytes)
    # the dominant kernel: the gate kernel (first launch) of the QoT workloads, else the step kernel (the only launch)
    if auto and spec.qot_gate:
        r["kern_name"], r["kern_ms"] = "qot_kernel (GN-model QoT gate, heuristic + verdict)", float(first.sum())
        r["bytes"], r["bytes_note"] = qot_gate_bytes(spec, eng, act)
        r["step_kernel_us"] = float((ms - first).sum()) * 1e3 / K
    else:
        r["kern_name"], r["kern_ms"] = "env_kernel (MODE_STEP, fused next-action hint)", float(ms.sum())
        r["bytes"], r["bytes_note"] = w["bytes"], ("SURVEY.md 8(d): every row of the k candidate paths read once, chosen "
                                                   "path written, one release")
        if (spec.env_kind == S.ENV_RMCSA and hid == S.H_SAP_BM_FC_FF) or (spec.env_kind == S.ENV_RMSA and hid == S.H_SAP_FF):
            r["bytes"], r["bytes_note"] = first_fit_bytes(spec, eng, act, w["bytes"])
    if full:  # warm-L2 variant (what a training loop sees when the state fits L2) and the fused rollout, informational
        r["warm_ms"] = float(timed(K, None)[0].sum())
        if not spec.qot_gate:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            eng.set_next_action_hint(hid, None)
            torch.cuda.synchronize()
            e0.record()
            eng.rollout(hid, K)
            e1.record()
            torch.cuda.synchronize()
            r["roll_ms"] = e0.elapsed_time(e1)
    cnt = eng.counters()
    r["stats"] = torch.stack([cnt[:, 0].sum(), cnt[:, 1].sum(), cnt[:, 4].sum(), cnt[:, 5].sum(), cnt[:, 11].sum(),
                              cnt[:, 10].max()]).to(torch.int64)
    r["queue"] = float(cnt[:, 9].float().mean())
    eng.close()
    del eng
    # ---- e2e ----
    ncpu = os.cpu_count() or 1
    block_wait = {"spin": False, "block": True}.get(args.host_wait, world * 2 > ncpu)
    # shards pay off when a step is long against the host's per-call work, or when the host computes the policy; the
    # 40-80 us step kernels of the workloads block are host bound with four times the calls
    slow_step = r["step_ms"] / K > 0.15 or (spec.env_kind == S.ENV_DEEPRMSA and spec.obs_dim > 0)
    parts = max(1, args.e2e_parts) if (full or slow_step) else 1
    while n_env % parts:
        parts -= 1
    r["e2e"] = run_e2e(spec, w, n_env, parts, max(1, min(K, args.e2e_steps if full else 10)), local, rank,
                       args.heap_capacity, block_wait, barrier)
    r["e2e"]["host_wait"] = "sleep on a blocking-sync event" if block_wait else "cudaStreamSynchronize (spin)"
    return r


def workload_record(r, n_total, K, peak):
    """The compact per-workload record of the `workloads` block."""
    kern_s = r["kern_ms"] * 1e-3 / K
    ach = r["bytes"] * (n_total) / kern_s / 1e9
    st = [int(x) for x in r["stats"].cpu()]
    e = r["e2e"]
    return {"what": r["w"]["what"], "env": r["w"]["env"], "envs_per_gpu": n_total, "value": n_total * K / (r["step_ms"] * 1e-3),
            "ms_per_step": r["step_ms"] / K, "median_ms_per_step": r["median_ms"], "launches_per_step": r["launches"] / K,
            "kernel": r["kern_name"], "kernel_us": kern_s * 1e6, "bytes_per_env_step": r["bytes"],
            "roofline_frac": ach / peak, "achieved_gbs": ach, "blocking": 1.0 - st[1] / max(1, st[0]),
            "e2e": n_total * e["steps"] / (e["ms"] * 1e-3), "e2e_d2h_bytes_per_step": e["d2h"],
            "hbm_bytes_per_env": r["hbm"] / max(1, n_total)}


def run_reference(args, name):
    """--impl reference: the reference's algorithm for the path on the box's host cores (the oracle C port: the
    reference is Python and cannot travel to the GPU box), all host threads, a bounded sample of the same workload."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    spec, w = make_workload(name)
    threads = os.cpu_count() or 1
    n_sample = args.ref_envs or w["cpu_envs"]
    v, ms, passes, blocking = cpu_arm(spec, w, n_sample, threads, steps=args.steps, warmup=args.warmup)
    sample = (f"{n_sample} envs of the {envs_of(name, args)}-env batch x {passes} steps in one pool call, oracle C port "
              f"(oracle/orlg_oracle.c) with its own heuristic, {threads} pthreads, {w['preroll']} pre-roll steps")
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 bitmask + f64 clock + f64 GSNR (oracle: u8 cells + f64)", "data": DATA,
        "config": {"workload": name, "env": w["env"], "envs_per_pass": n_sample, "heuristic": int(w["hid"]),
                   "blocking": blocking, "what": w["what"]},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample,
                         "python_reference": python_reference(name)},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args, name):
    import torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the engine has no CPU fallback")
    if args.gpus != world:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}: one process per GPU is launched by torchrun; "
              f"measuring {world} GPU(s)", file=sys.stderr)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = dict(local=local, rank=rank, world=world, barrier=barrier,
               flush=None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev))
    K = args.steps
    r = measure(name, args, ctx, full=True)
    spec, w, e = r["spec"], r["w"], r["e2e"]

    # ---- episode statistics: the only collective on this path (NCCL all-reduce, off the step path) ----
    stats = r["stats"]
    times = torch.tensor([r["step_ms"], r["kern_ms"], r["warm_ms"], r.get("roll_ms", 0.0), e["ms"], r["median_ms"]],
                         dtype=torch.float64, device=dev)
    mhz = torch.tensor([r["clocks"]["sm_mhz"] or 0.0], dtype=torch.float64, device=dev)
    if dist is not None:
        fault = stats[5:6].clone()
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        dist.all_reduce(fault, op=dist.ReduceOp.MAX)
        stats[5] = fault[0]
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(mhz, op=dist.ReduceOp.MIN)
    step_ms, kern_ms, warm_ms, roll_ms, e2e_ms, median_ms = [float(x) for x in times.cpu()]
    stats = [int(x) for x in stats.cpu()]
    if stats[5] != 0:
        raise RuntimeError(f"device fault bits set: {stats[5]}")
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    n_env = envs_of(name, args)
    total = n_env * world
    peak, peak_src = measured_peaks()
    kern_s = kern_ms * 1e-3 / K
    achieved = r["bytes"] * n_env / kern_s / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(REPO, "profiles", "traffic.json"))).get(name, {}).get("dram_bytes_per_launch")
    except Exception:
        pass
    clocks = dict(r["clocks"], min_sm_mhz_over_ranks=float(mhz[0]))
    line = {
        "metric": METRIC, "value": total * K / (step_ms * 1e-3), "unit": "env-steps/s", "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": step_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 bitmask + f64 clock + f64 GSNR (fixed-point 2^-44 running sums)", "data": DATA,
        "config": {
            "workload": name, "what": w["what"], "env": w["env"], "envs_per_gpu": n_env, "heuristic": int(w["hid"]),
            "obs": OBS_NAME[spec.obs_kind],
            "step": ("orlg_step_auto: the QoT gate kernel (candidates of heuristic 13 + verdict) then the step kernel, 2 launches"
                     if w["step"] == "auto" else "orlg_step with the fused next-action hint of the device heuristic (1 launch)")
                    + ", actions/obs/reward/done resident in HBM",
            "timing": ("one CUDA graph of K x (L2-evicting 256 MiB memset, event, step, event), launched by orlg_timed_steps"
                       if not args.no_graph else "one C loop of K x (memset, event, step, event) in orlg_timed_steps")
                      + "; value = envs x K / sum of the K event intervals",
            "l2": "flushed before every step (outside the event pairs)" if not args.no_flush else "not flushed",
            "qot": ("GN-model GSNR (calculate_osnr) of every candidate and of its neighbours against the SD-FEC thresholds "
                    "decides accept/block and the modulation" if spec.qot_gate else "as the env defines it (tables / reach / "
                    "crosstalk gate), no GN model"),
            "preroll_steps": w["preroll"], "parallelism": f"envs sharded x{world}, no data-path collective",
            "blocking": 1.0 - stats[1] / max(1, stats[0]), "episodes": stats[4], "running_services_per_env": r["queue"],
        },
        "median_step_value": total / (median_ms * 1e-3),
        "e2e": {"value": total * e["steps"] / (e2e_ms * 1e-3), "unit": "env-steps/s", "h2d_bytes_per_step": e["h2d"],
                "d2h_bytes_per_step": e["d2h"], "steps": e["steps"], "ms_per_step": e2e_ms / e["steps"], "mode": e["mode"],
                "policy": e["policy"], "host_policy_ms_per_step": e["policy_ms"], "host_cores": os.cpu_count(),
                "host_wait": e["host_wait"], "d2h_gbs": e["d2h"] * e["steps"] / (e2e_ms * 1e-3) / 1e9},
        "gpu_launches": r["launches"],
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "kernel": r["kern_name"], "bytes_per_env_step": r["bytes"],
                     "bytes_definition": r["bytes_note"], "kernel_ms": kern_s * 1e3,
                     "note": "latency / issue bound long before HBM (DESIGN.md 4, 6f; profiles/): the fraction is small by nature"},
        "extra": {"warm_l2_env_steps_per_s": total * K / (warm_ms * 1e-3),
                  "fused_rollout_env_steps_per_s": (total * K / (roll_ms * 1e-3)) if roll_ms else None,
                  "step_kernel_us": r.get("step_kernel_us"), "hbm_bytes_per_gpu": r["hbm"]},
    }
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        n_sample = args.ref_envs or w["cpu_envs"]
        v, ms, passes, blk = cpu_arm(spec, w, n_sample, threads, seconds=args.cpu_seconds)
        line["cpu_baseline"] = {
            "value": v, "unit": "env-steps/s", "cores": threads, "kind": "port",
            "sample": (f"{n_sample} envs x {passes} steps of the same workload (blocking {blk:.3f}), oracle C port on "
                       f"{threads} pthreads, {w['preroll']} pre-roll steps untimed"),
            "python_reference": python_reference(name)}
    if world == 1 and not args.no_workloads:  # one record per BASELINE.json config, compact
        blockargs = argparse.Namespace(**vars(args))
        block = {}
        for nm in BLOCK:
            if nm == name:
                continue
            rr = measure(nm, blockargs, ctx, full=False)
            if int(rr["stats"][5]) != 0:
                raise RuntimeError(f"device fault bits set in workload {nm}")
            rec = workload_record(rr, envs_of(nm, blockargs), K, peak)
            if not args.no_cpu_baseline:
                v, _, passes, _ = cpu_arm(rr["spec"], rr["w"], rr["w"]["cpu_envs"], os.cpu_count() or 1,
                                          seconds=args.block_cpu_seconds)
                rec["cpu_port"] = {"value": v, "cores": os.cpu_count() or 1, "sample": f"{rr['w']['cpu_envs']} envs x {passes} steps"}
                rec["python_reference"] = python_reference(nm)
            block[nm] = rec
        line["workloads"] = block
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1, help="informational: one process per GPU comes from torchrun (WORLD_SIZE)")
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=HEADLINE, choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (0 = the workload's own: 65 536 unless it names another)")
    ap.add_argument("--ref-envs", type=int, default=0, help="envs in the CPU arm's bounded sample (0 = the workload's default)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--block-cpu-seconds", type=float, default=2.5)
    ap.add_argument("--e2e-steps", type=int, default=50)
    ap.add_argument("--e2e-parts", type=int, default=4, help="VecEnv shards (streams) the pipelined e2e loop cycles through")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="time with the C loop instead of one CUDA graph")
    ap.add_argument("--host-wait", default="auto", choices=["auto", "spin", "block"],
                    help="orlg_step_host_wait: spin on the stream, sleep on an event, or decide from ranks x 2 > cores")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-workloads", action="store_true", help="skip the per-config `workloads` block (N = 1 only)")
    ap.add_argument("--heap-capacity", type=int, default=0,
                    help="release-queue entries per env (0 = sized by the offered load; overflow sets a fault bit, which fails the run)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args, args.workload)
    else:
        run_b200(args, args.workload)


if __name__ == "__main__":
    main()
