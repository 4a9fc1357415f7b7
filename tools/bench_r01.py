#!/usr/bin/env python
"""bench.py — env-steps/sec of the batched stepping engine on B200 (driver contract).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA engine
    python bench.py --impl reference --gpus N --steps K ...   # CPU arm (oracle port)

One "step" = one pass of the hot path over one batch of envs: the device
heuristic picks an action for every env (orlg_heuristic), then every env takes
one gym step (orlg_step: validate -> provision -> release queue -> next request
-> reward/done -> DeepRMSA observation).  The default workload is
BASELINE.json configs[1]: DeepRMSA-v0 on NSFNET, k=5, 100 slots, 65536 envs per
GPU (weak scaling: env count per GPU is fixed, envs are sharded with no
data-path collective; NCCL only all-reduces the episode statistics at the end).

`value`   device-resident throughput (actions, state and outputs stay in HBM).
`e2e`     the same metric through the host-buffer C-ABI call (orlg_step_host):
          a numpy policy on the host picks actions from the previous host
          observation, actions go H2D from pinned memory, obs/reward/done come
          back D2H every step; copies and host policy are inside the timed region.
`roofline` the step kernel: algorithmic bytes per env-step x envs /
          its CUDA-event duration, against MEASURED_PEAKS.json hbm_gbs.
`cpu_baseline` the C oracle port (oracle/, test infrastructure) stepping a
          bounded sample of the same workload on all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
for _p in (REPO, os.path.join(REPO, "oracle"), os.path.join(REPO, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

from optical_rl_gym_qot_aware_b200 import spec as S  # noqa: E402
from optical_rl_gym_qot_aware_b200.tables import TopologyTables  # noqa: E402

GOLDEN = os.path.join(REPO, "tests", "golden")
METRIC = "env-steps/sec"

# node probabilities of the reference's DeepRMSA driver (tests/test_deeprmsa.py:30-47)
DEEPRMSA_NODE_PROBS = [0.01801802, 0.04004004, 0.05305305, 0.01901902, 0.04504505, 0.02402402, 0.06706707,
                       0.08908909, 0.13813814, 0.12212212, 0.07607608, 0.12012012, 0.01901902, 0.16916917]

# name -> (env id, topology fixture, kwargs, heuristic, obs kind, algorithmic bytes per env-step)
# bytes/env-step are SURVEY.md section 8(d) / DESIGN.md "Algorithmic bytes".
WORKLOADS = {
    "c1_rmsa_nsfnet_100": ("RMSA-v0", "nsfnet", dict(load=50, num_spectrum_resources=100, episode_length=1000),
                           S.H_SAP_FF, S.OBS_NONE, 609),
    "c2_deeprmsa_nsfnet_100": ("DeepRMSA-v0", "nsfnet",
                               dict(j=1, mean_service_holding_time=7.5, mean_service_inter_arrival_time=1 / 12.0,
                                    num_spectrum_resources=100, episode_length=50,
                                    node_request_probabilities=DEEPRMSA_NODE_PROBS),
                               S.H_SAP_FF, S.OBS_F32, 825),
    "c3_rmcsa_jpn12_320": ("RMCSA-v0", "jpn12_k5",
                           dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                num_spatial_resources=7, worst_xt=-84.7, episode_length=1000),
                           S.H_SAP_BM_FC_FF, S.OBS_NONE, 6454),
    "c5_rmsa_spn_100": ("RMSA-v0", "spn", dict(load=100, mean_service_holding_time=25,
                                               num_spectrum_resources=100, episode_length=1000),
                        S.H_SAP_FF, S.OBS_NONE, 517),
    "c5_rmsa_spn_320": ("RMSA-v0", "spn", dict(load=400, mean_service_holding_time=25,
                                               num_spectrum_resources=320, episode_length=1000),
                        S.H_SAP_FF, S.OBS_NONE, 1039),
}
# C4: PhyRMSA-v0 (phy-frag) on US14 with 320 channels.  The shipped QoT tables have 268 channels
# (SURVEY.md F4); the synthetic 320-channel tables keep them and replicate the S-band tail.
WORKLOADS["c4_phyrmsa_us14_320"] = ("PhyRMSA-v0", "us14", dict(load=1400, mean_service_holding_time=25,
                                                               episode_length=200, number_spectrum_channels=160,
                                                               s_band=False),
                                    S.H_PHY_BM_FA_CUT, S.OBS_PHY_CUTS, 1426)
GSNR_WORKLOADS = {"c3_rmcsa_jpn12_320"}
DEFAULT_WORKLOAD = "c2_deeprmsa_nsfnet_100"
PREROLL = 600  # steps run before any timing so that the networks are in steady state
PREROLL_OF = {"c4_phyrmsa_us14_320": 6000, "c3_rmcsa_jpn12_320": 1500}  # loads of 1400 / 250 Erlang fill up slowly


def make_workload(name, seed=0x5EED):
    env_id, topo, kw, hid, obs_kind, bytes_per_step = WORKLOADS[name]
    tables = TopologyTables.load(os.path.join(GOLDEN, f"topology_{topo}.npz"))
    if env_id == "PhyRMSA-v0":
        z = np.load(os.path.join(GOLDEN, "phytab_Results_K3SP_FRP_SLC_CBG_USB14.npz"))
        cols = np.concatenate([np.arange(268), np.arange(268 - 52, 268)])  # 320 channels
        kw = dict(kw, modulation_level=z["phy_mod"][:, cols], gsnr=z["phy_gsnr"][:, cols],
                  connections_detail=z["phy_conn"])
    spec = S.make_spec(env_id, tables, seed=seed, request_source=S.SRC_PHILOX, auto_reset=True,
                       obs_kind=obs_kind, **kw)
    return spec, hid, bytes_per_step, env_id


def measured_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock and throttle reasons sampled every 5 ms during the timed region (NVML; the
    nvidia-smi -lms loop of the recipe starts too slowly for a sub-second region)."""

    def __init__(self, index):
        self.index, self.sm, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self.thread = None

    def start(self):
        try:
            import pynvml as N

            N.nvmlInit()
            # NVML enumerates all GPUs of the box; honour CUDA_VISIBLE_DEVICES if it is a plain list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
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
    """K-SP-FF (deeprmsa_env.py:146-155) evaluated on the host from the observation:
    path p has a fitting block iff its first block-length feature is not -1.
    obs_fm: [obs_dim, n_env] (feature-major host buffer).  Returns int32 [n_env, 1].
    The k availability bits of an env are packed into one byte and mapped through a table."""
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
    # Horner over the paths, highest first: code = 2*code | (path p has a fitting block)
    np.not_equal(obs_fm[base + (k - 1) * rowlen + 1], -1.0, out=code.view(np.bool_))
    for p in range(k - 2, -1, -1):
        np.not_equal(obs_fm[base + p * rowlen + 1], -1.0, out=bit)
        np.add(code, code, out=code)
        np.bitwise_or(code, bit.view(np.uint8), out=code)
    np.copyto(idx, code)
    if out is None:
        return lut.take(idx).reshape(-1, 1)
    np.take(lut, idx, out=out.reshape(-1), mode="clip")  # codes are < 2**k by construction
    return out


def cpu_arm(spec, hid, n_sample, min_seconds, threads, steps=None, warmup=0, preroll=PREROLL, gsnr=False):
    """Times the oracle port on host threads.  Returns (steps/s, ms per pass, passes)."""
    import oracle_binding as OB

    pool = OB.OraclePool(spec, 0, n_sample)
    pool.rollout(preroll, hid, threads)
    if gsnr:  # the GN model for every decision, like the GPU arm's orlg_gsnr_for_actions launch
        pool.set_gsnr(S.gsnr_params(spec.tables, spec.channel_width))
    for _ in range(warmup):
        pool.rollout(1, hid, threads)
    done_steps, passes = 0, 0
    t0 = time.perf_counter()
    if steps is not None:
        for _ in range(steps):
            done_steps += pool.rollout(1, hid, threads)
            passes += 1
    else:
        while time.perf_counter() - t0 < min_seconds:
            done_steps += pool.rollout(50, hid, threads)
            passes += 50
    dt = time.perf_counter() - t0
    cnt = pool.counters()
    blocking = 1.0 - cnt[:, 1].sum() / max(1, cnt[:, 0].sum())
    pool.close()
    return done_steps / dt, dt * 1e3 / max(1, passes), passes, float(blocking)


def run_reference(args, spec, hid, env_id, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_sample = args.ref_envs
    PREROLL = PREROLL_OF.get(wl, 600)
    v, ms, passes, blocking = cpu_arm(spec, hid, n_sample, 0, threads, steps=args.steps, warmup=args.warmup,
                                      preroll=PREROLL, gsnr=wl in GSNR_WORKLOADS and not args.no_gsnr)
    sample = (f"{n_sample} envs of the {args.envs}-env batch per pass, oracle C port (oracle/orlg_oracle.c) "
              f"with its own heuristic, {threads} pthreads, {PREROLL} pre-roll steps")
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32 bitmask + f64 clock (oracle: u8 cells + f64)",
        "data": "synthetic (Philox request streams, seed 0x5eed, keyed by global env id)",
        "config": {"workload": wl, "env": env_id, "envs_per_pass": n_sample, "heuristic": int(hid),
                   "blocking": blocking},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args, spec, hid, bytes_per_step, env_id, wl):
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    PREROLL = PREROLL_OF.get(wl, 600)
    n_env = args.envs
    eng = Engine(spec, n_env, device=local, global_env_offset=rank * n_env, heap_capacity=args.heap_capacity)
    eng.reset(True)
    eng.rollout(hid, PREROLL)
    act = torch.zeros((n_env, eng.action_dim), dtype=torch.int32, device=dev)
    rew = torch.zeros(n_env, dtype=torch.int32, device=dev)
    done = torch.zeros(n_env, dtype=torch.uint8, device=dev)
    obs = eng.new_obs() if eng.obs_dim else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if not args.no_flush else None

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # BASELINE configs[2] "QoT-aware RMCSA with GN-model GSNR": every decision's candidate lightpath and its
    # neighbours are evaluated with the GN model (orlg_gsnr_for_actions on the action tensor, one more launch)
    gs_out = None
    if wl in GSNR_WORKLOADS and not args.no_gsnr:
        eng.gsnr_configure()
        gs_out = torch.zeros((n_env, 2), dtype=torch.float64, device=dev)
    fused = not args.two_launch
    eng.heuristic(hid, act)  # action for the first pending request
    if fused:  # every step also writes the heuristic's action for the next request into `act`
        eng.set_next_action_hint(hid, act)

    def one_step(evs=None):
        """evs = [start, after the heuristic launch, after the step launch, end]; only the events that separate
        launches of this configuration are recorded (an event record costs GPU front-end time too)."""
        if flush is not None:
            flush.fill_(1)  # evict the 126 MB L2 (untimed: outside the event pairs)
        if evs is not None:
            evs[0].record()
        if not fused:
            eng.heuristic(hid, act)
            if gs_out is not None:
                eng.gsnr_for_actions(act, gs_out)
            if evs is not None:
                evs[1].record()
        eng.step(act, rew, done, obs, None)
        if evs is not None:
            evs[2].record()
        if fused and gs_out is not None:  # the hint is the action for the request that has just become pending
            eng.gsnr_for_actions(act, gs_out)
            if evs is not None:
                evs[3].record()

    def ev_times(evs):
        """(whole step, step kernel, GN-model kernel or None) in ms summed over the recorded steps"""
        k0 = 0 if fused else 1
        end = 3 if (fused and gs_out is not None) else 2
        return (sum(e[0].elapsed_time(e[end]) for e in evs), sum(e[k0].elapsed_time(e[2]) for e in evs),
                sum(e[2].elapsed_time(e[3]) for e in evs) if end == 3 else None)

    for _ in range(args.warmup):
        one_step()
    K = args.steps
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = eng.launch_count
    barrier()
    for t in range(K):
        one_step(evs[t])
    barrier()
    launches = eng.launch_count - launches0
    step_ms, kern_ms, gsnr_ms = ev_times(evs)
    bytes_note = "SURVEY.md 8(d): every row of the k candidate paths read once, chosen path written, one release"
    if fused and ((spec.env_kind == S.ENV_RMCSA and hid == S.H_SAP_BM_FC_FF) or
                  (spec.env_kind == S.ENV_RMSA and hid == S.H_SAP_FF)):
        # shortest-available-path first-fit stops at the first path (RMCSA: first (path, core)) that fits, so the rows
        # it has to read follow from the decisions actually taken; SURVEY 8(d)'s figures assume all k paths (x all cores)
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
        cnt_now = eng.counters()
        accept = float(cnt_now[:, 1].sum()) / max(1.0, float(cnt_now[:, 0].sum()))
        rows_rw = accept * 3.0 * float(np.where(ok, hops[idx, pth], 0).mean())  # write the chosen rows, read+write a release
        words = (spec.num_slots + 31) // 32
        upper = bytes_per_step
        bytes_per_step = int(round(4 * words * (float(rows_read.mean()) + rows_rw) + 168))
        bytes_note = (f"rows of the paths{' x cores' if cores > 1 else ''} first-fit examined for the decisions taken "
                      f"({float(rows_read.mean()):.2f} rows read per step on average), chosen rows written, one release "
                      f"per accepted service, 168 B fixed (SURVEY.md 8(d) upper bound: {upper} B)")
    if gs_out is not None and not bool(torch.isfinite(gs_out[:, 0]).any()):
        raise RuntimeError("GSNR evaluation produced no finite value")

    # warm-L2 variant (what a training loop sees when the state fits L2): one event pair around K steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    saved, flush = flush, None
    barrier()
    e0.record()
    for t in range(K):
        one_step()
    e1.record()
    barrier()
    warm_ms = e0.elapsed_time(e1)
    flush = saved

    # the same step as two launches (orlg_heuristic, then orlg_step without the hint), informational
    two_ms = None
    if fused:
        eng.set_next_action_hint(hid, None)
        fused = False
        for _ in range(3):
            one_step()
        barrier()
        ev2 = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
        for t in range(K):
            one_step(ev2[t])
        barrier()
        two_ms = ev_times(ev2)[0]
        fused = True
        eng.heuristic(hid, act)
        eng.set_next_action_hint(hid, act)

    # fused rollout (heuristic + step in one launch, state on chip between steps), informational
    e0.record()
    eng.rollout(hid, K)
    e1.record()
    torch.cuda.synchronize()
    roll_ms = e0.elapsed_time(e1)

    # ---- e2e: host buffers through orlg_step_host, host policy, copies inside the timed region ----
    t_nodes, t_k = spec.tables.n_nodes, spec.tables.k_paths
    h_act = torch.zeros((n_env, eng.action_dim), dtype=torch.int32).pin_memory()
    h_rew = torch.zeros(n_env, dtype=torch.int32).pin_memory()
    h_done = torch.zeros(n_env, dtype=torch.uint8).pin_memory()
    h_gs = torch.zeros((n_env, 2), dtype=torch.float64).pin_memory() if gs_out is not None else None
    use_obs_policy = env_id == "DeepRMSA-v0" and eng.obs_dim > 0
    h_obs = None
    if eng.obs_dim:
        odt = {S.OBS_F64: torch.float64, S.OBS_F32: torch.float32, S.OBS_MATRIX_U8: torch.uint8,
               S.OBS_PHY_INFO: torch.float64, S.OBS_PHY_CUTS: torch.float64}[spec.obs_kind]
        h_obs = torch.zeros((eng.obs_dim, n_env), dtype=odt).pin_memory()  # feature-major, as the device buffer
        eng.observation(obs)
        h_obs.copy_(obs.t())
    torch.cuda.synchronize()

    policy_s = [0.0]

    def e2e_step():
        if use_obs_policy:
            p0 = time.perf_counter()
            host_policy_kspff(h_obs.numpy(), t_nodes, t_k, spec.j, out=h_act.numpy())
            policy_s[0] += time.perf_counter() - p0
        else:  # no observation to decide from: device heuristic (or its fused hint), actions D2H then H2D
            if not fused:
                eng.heuristic(hid, act)
            h_act.copy_(act)
        eng.step_host(h_act, h_rew, h_done, h_obs, None)
        if gs_out is not None:
            if fused:
                eng.gsnr_for_actions(act, gs_out)
            h_gs.copy_(gs_out)  # synchronous D2H of the step's GSNR values
        return int(h_rew.numpy()[0])

    eng.heuristic(hid, act)  # the rollout above moved the envs on: refresh the pending actions
    Ke = max(1, min(K, args.e2e_steps))
    for _ in range(min(args.warmup, 5)):
        e2e_step()
    barrier()
    policy_s[0] = 0.0
    e0.record()
    w0 = time.perf_counter()
    for t in range(Ke):
        e2e_step()
    e1.record()
    barrier()
    e2e_wall_ms = (time.perf_counter() - w0) * 1e3
    e2e_ms = max(e0.elapsed_time(e1), e2e_wall_ms)
    e2e_serial_ms, e2e_mode = e2e_ms, "one orlg_step_host call per step (serial: policy, H2D, step, D2H)"

    # ---- e2e, pipelined: the batch as two VecEnv halves on two streams (step_async / step_wait);
    # the host policy of one half overlaps the copies and the kernel of the other ----
    parts = max(2, args.e2e_parts)
    # ranks x (a waiting main thread + the driver's helper threads) against the host's cores: spin-waiting is the
    # fastest when every rank has cores to spare, and it starves the policy when it has not
    ncpu = os.cpu_count() or 1
    block_wait = {"spin": False, "block": True}.get(args.host_wait, world * 2 > ncpu)
    if use_obs_policy and n_env % parts == 0 and not args.e2e_serial:
        nh = n_env // parts
        halves = []
        for hi in range(parts):
            he = Engine(spec, nh, device=local, global_env_offset=rank * n_env + hi * nh, heap_capacity=args.heap_capacity)
            he.set_host_wait(block_wait)
            he.reset(True)
            he.rollout(hid, PREROLL)
            dobs = he.new_obs()
            he.observation(dobs)
            ho = torch.zeros((he.obs_dim, nh), dtype=h_obs.dtype).pin_memory()
            ho.copy_(dobs.t())
            halves.append(dict(e=he, st=torch.cuda.Stream(device=dev), act=torch.zeros((nh, 1), dtype=torch.int32).pin_memory(),
                               rew=torch.zeros(nh, dtype=torch.int32).pin_memory(),
                               done=torch.zeros(nh, dtype=torch.uint8).pin_memory(), obs=ho))
        torch.cuda.synchronize()

        def submit(h):
            p0 = time.perf_counter()
            host_policy_kspff(h["obs"].numpy(), t_nodes, t_k, spec.j, out=h["act"].numpy())
            policy_s[0] += time.perf_counter() - p0
            h["e"].step_host_async(h["act"], h["rew"], h["done"], h["obs"], None, stream=h["st"])

        def pipelined(nsteps):
            for h in halves:
                submit(h)
            for _ in range(nsteps - 1):
                for h in halves:
                    h["e"].step_host_wait(stream=h["st"])
                    submit(h)
            for h in halves:
                h["e"].step_host_wait(stream=h["st"])

        pipelined(5)
        barrier()
        policy_s[0] = 0.0
        e0.record()
        w0 = time.perf_counter()
        pipelined(Ke)
        e1.record()
        barrier()
        e2e_ms = max(e0.elapsed_time(e1), (time.perf_counter() - w0) * 1e3)
        e2e_mode = (f"{parts} VecEnv shards of {nh} envs on {parts} streams, orlg_step_host_async / _wait "
                    "(the host policy of one shard overlaps the copies and kernels of the others)")
        for h in halves:
            if int(h["e"].counters()[:, 10].max()) != 0:
                raise RuntimeError("device fault bits set in the pipelined e2e run")
            h["e"].close()
    clocks = sampler.stop() if rank == 0 else None
    # what the host link can do with the same buffer: pinned D2H copies of the observation alone
    pcie_gbs = None
    if h_obs is not None:
        dsrc = obs.t().contiguous() if obs.dim() == 2 else obs
        for _ in range(3):
            h_obs.copy_(dsrc, non_blocking=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            h_obs.copy_(dsrc, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        pcie_gbs = 20 * h_obs.numel() * h_obs.element_size() / (e0.elapsed_time(e1) * 1e-3) / 1e9
    h2d = h_act.numel() * 4
    d2h = h_rew.numel() * 4 + h_done.numel() + (h_obs.numel() * h_obs.element_size() if h_obs is not None else 0)
    d2h += h_gs.numel() * 8 if h_gs is not None else 0

    # ---- episode statistics: the only collective on this path (NCCL all-reduce, off the step path) ----
    cnt = eng.counters()
    stats = torch.stack([cnt[:, 0].sum(), cnt[:, 1].sum(), cnt[:, 4].sum(), cnt[:, 5].sum(), cnt[:, 11].sum(),
                         cnt[:, 10].max()]).to(torch.int64)
    times = torch.tensor([step_ms, kern_ms, warm_ms, roll_ms, e2e_ms, e2e_serial_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        fault = stats[5:6].clone()
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        dist.all_reduce(fault, op=dist.ReduceOp.MAX)
        stats[5] = fault[0]
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    step_ms, kern_ms, warm_ms, roll_ms, e2e_ms, e2e_serial_ms = [float(x) for x in times.cpu()]
    stats = [int(x) for x in stats.cpu()]
    if stats[5] != 0:
        raise RuntimeError(f"device fault bits set: {stats[5]}")
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    total_envs = n_env * world
    value = total_envs * K / (step_ms * 1e-3)
    peak, peak_src = measured_peaks()
    kern_avg_s = kern_ms * 1e-3 / K
    achieved = bytes_per_step * n_env / kern_avg_s / 1e9
    traffic = None
    tp = os.path.join(REPO, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(wl, {}).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    line = {
        "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": step_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 bitmask + f64 clock",
        "data": "synthetic (Philox request streams, seed 0x5eed, keyed by global env id)",
        "config": {
            "workload": wl, "env": env_id, "envs_per_gpu": n_env, "heuristic": int(hid),
            "obs": {0: "none", 1: "f64", 2: "f32", 3: "matrix_u8", 4: "phy cuts+rss f64", 5: "phy cuts f64"}[
                spec.obs_kind],
            "step": ("orlg_step with the fused next-action hint of the device heuristic (1 launch)" if fused
                     else "orlg_heuristic + orlg_step (2 launches)") + ", actions/obs/reward/done resident in HBM",
            "l2": ("flushed before every step by a 256 MiB fill outside the per-step event pairs"
                   if not args.no_flush else "not flushed (state may be L2 resident)"),
            "qot": ("GN-model GSNR of every decision's candidate and its neighbours (orlg_gsnr_for_actions, +1 launch)"
                    if gs_out is not None else "as the env defines it (tables / reach / crosstalk gate), no GN-model launch"),
            "preroll_steps": PREROLL, "parallelism": f"envs sharded x{world}, no data-path collective",
            "blocking": 1.0 - stats[1] / max(1, stats[0]), "episodes": stats[4],
        },
        "e2e": {"value": total_envs * Ke / (e2e_ms * 1e-3), "unit": "env-steps/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "steps": Ke, "ms_per_step": e2e_ms / Ke, "mode": e2e_mode,
                "serial_value": total_envs * Ke / (e2e_serial_ms * 1e-3),
                "host_policy_ms_per_step": policy_s[0] * 1e3 / Ke, "host_cores": ncpu,
                "host_wait": "sleep on a blocking-sync event" if block_wait else "cudaStreamSynchronize (spin)",
                "d2h_gbs": d2h * Ke / (e2e_ms * 1e-3) / 1e9, "pinned_d2h_copy_gbs": pcie_gbs,
                "policy": "numpy K-SP-FF from the host observation" if use_obs_policy
                else "device heuristic, actions D2H then H2D"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src,
                     "kernel": "env_kernel (MODE_STEP)" + ("; this workload's step is dominated by the GN-model launch "
                                                           "(extra.gsnr_kernel_us, fp64 issue bound, DESIGN.md section 6)"
                                                           if gsnr_ms is not None else ""),
                     "bytes_per_env_step": bytes_per_step, "bytes_definition": bytes_note,
                     "kernel_ms": kern_avg_s * 1e3},
        "extra": {"two_launch_env_steps_per_s": (total_envs * K / (two_ms * 1e-3)) if two_ms else None,
                  "warm_l2_env_steps_per_s": total_envs * K / (warm_ms * 1e-3),
                  "fused_rollout_env_steps_per_s": total_envs * K / (roll_ms * 1e-3),
                  "gsnr_kernel_us": (gsnr_ms * 1e3 / K) if gsnr_ms is not None else None,
                  "hbm_bytes_per_gpu": eng.hbm_bytes},
    }
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, ms, passes, blk = cpu_arm(spec, hid, args.ref_envs, args.cpu_seconds, threads, preroll=PREROLL,
                                     gsnr=gs_out is not None)
        line["cpu_baseline"] = {
            "value": v, "unit": "env-steps/s", "cores": threads, "kind": "port",
            "sample": (f"{args.ref_envs} envs x {passes} steps of the same workload (blocking {blk:.3f}), oracle C port "
                       f"on {threads} pthreads, {PREROLL} pre-roll steps untimed"),
        }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU")
    ap.add_argument("--ref-envs", type=int, default=8192, help="envs in the CPU arm's bounded sample")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--e2e-steps", type=int, default=100)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--e2e-serial", action="store_true", help="no pipelining of the e2e loop")
    ap.add_argument("--host-wait", default="auto", choices=["auto", "spin", "block"],
                    help="orlg_step_host_wait: spin on the stream, sleep on an event, or decide from ranks x 2 > cores")
    ap.add_argument("--e2e-parts", type=int, default=4, help="VecEnv shards (streams) the pipelined e2e loop cycles through")
    ap.add_argument("--two-launch", action="store_true", help="orlg_heuristic + orlg_step instead of the fused hint")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gsnr", action="store_true", help="c3 only: skip the per-decision GN-model launch")
    ap.add_argument("--heap-capacity", type=int, default=0,
                    help="release-queue entries per env (0 = the provable bound; overflow sets a fault bit, which fails the run)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    spec, hid, bytes_per_step, env_id = make_workload(args.workload)
    if args.impl == "reference":
        run_reference(args, spec, hid, env_id, args.workload)
    else:
        run_b200(args, spec, hid, bytes_per_step, env_id, args.workload)


if __name__ == "__main__":
    main()
