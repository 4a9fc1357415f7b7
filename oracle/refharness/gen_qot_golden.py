"""Golden traces for the GN-model QoT gate (TEST INFRASTRUCTURE).

    python oracle/refharness/gen_qot_golden.py      # writes tests/golden/qot_*.npz

The reference has no env that calls its GN model, so the *gate* — accept iff the
candidate's GSNR and every neighbour's GSNR, re-evaluated with the candidate present, meet the SD-FEC
threshold of their modulations (examples/create_topology_gn.py:65-110) — is this repo's definition
(include/orlg.h "QoT gate").  What IS the reference's here:

* every GSNR is one call of the unmodified ``calculate_osnr`` (examples/calculate_osnr.py:9-56, imported
  as gen_gsnr_golden.py imports it);
* the envs are the unmodified RMSAEnv / RMCSAEnv: slot counts (``get_number_slots``), first fit
  (``is_path_free`` over ``range(0, S - n)``), RMCSA's closed-form gate (``_crosstalk_is_acceptable``),
  provisioning, release queue and counters all run in the reference.  A blocked candidate is handed to
  ``env.step`` as the reject action; an adaptive RMSA-v0 decision is handed over with the path's
  ``best_modulation`` pointing at the chosen modulation for the duration of the call (the env reads the
  spectral efficiency nowhere else).

Stored per step: the pending request, the action given to the gated step, accepted, slots, and the two
GSNR values the gate looked at last (candidate, lowest neighbour).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import gen_gsnr_golden as GG  # noqa: E402  (installs the stubs, injects Span / Link, imports calculate_osnr)

from optical_rl_gym_qot_aware_b200 import spec as S  # noqa: E402
from optical_rl_gym_qot_aware_b200.tables import tables_from_topology  # noqa: E402

gym, refenv = GG.gym, GG.refenv
THR = {se: v for se, v in enumerate(S.SD_FEC_SNR_DB, start=1)}


class Gate:
    def __init__(self, env, kind, gp):
        self.env, self.kind = env, kind
        self.scene = GG.Scene(env, gp, kind)
        self.last = [np.nan, np.nan]

    def cand(self, path, core, start, n, se):
        return self.scene.shadow(-1, path, core, start, n, se)

    def alone_ok(self, path, start, n, se):
        c = self.cand(path, 0, start, n, se)
        return self.scene.evaluate(c, []) >= THR[se]

    def ok(self, path, core, start, n, se):
        running = self.scene.running()
        c = self.cand(path, core, start, n, se)
        g0 = self.scene.evaluate(c, running)
        self.last = [g0, np.nan]
        if not g0 >= THR[se]:
            return False
        good, mn = True, np.inf
        for j in running:
            if j.core == core and any(lk in j.path.links for lk in c.path.links):
                v = self.scene.evaluate(j, running + [c])
                mn = min(mn, v)
                if not v >= THR[j.current_modulation.spectral_efficiency]:
                    good = False
        self.last[1] = mn
        return good


def mods_desc(env):  # the order get_best_modulation_format walks (utils.py:105-117)
    mods = env.topology.graph["modulations"] if not hasattr(env, "modulation_formats") else env.modulation_formats
    return sorted(mods, key=lambda x: x.spectral_efficiency, reverse=True)


def slots_with(env, path, mod):
    """env.get_number_slots with `mod` as the path's modulation (RMSA-v0 reads path.best_modulation)."""
    keep = path.best_modulation
    path.best_modulation = mod
    try:
        return env.get_number_slots(path)
    finally:
        path.best_modulation = keep


def first_fit_rmsa(env, path, n):
    for s in range(0, env.topology.graph["num_spectrum_resources"] - n):  # rmsa_env.py:908-910 (F8)
        if env.is_path_free(path, s, n):
            return s
    return -1


def rmsa_heuristic(env, gate, adaptive):
    """heuristic 13 on RMSA-v0"""
    svc = env.current_service
    gate.last = [np.nan, np.nan]
    for p, path in enumerate(env.k_shortest_paths[svc.source, svc.destination]):
        mods = mods_desc(env) if adaptive else [path.best_modulation]
        for mod in mods:
            n = slots_with(env, path, mod)
            s = first_fit_rmsa(env, path, n)
            if s < 0:
                break
            se = mod.spectral_efficiency
            if adaptive and not gate.alone_ok(path, s, n, se):
                continue
            if gate.ok(path, 0, s, n, se):
                return [p, s]
    return [env.k_paths, env.num_spectrum_resources]


def rmsa_gated_step(env, gate, adaptive, a):
    """RMSAEnv.step behind the gate; returns (env.step output, accepted modulation or None)."""
    k, S_ = env.k_paths, env.num_spectrum_resources
    svc = env.current_service
    reject = [k, S_]
    gate.last = [np.nan, np.nan]
    if not (0 <= a[0] < k and 0 <= a[1] < S_):
        return env.step(reject), None
    path = env.k_shortest_paths[svc.source, svc.destination][a[0]]
    mods = mods_desc(env) if adaptive else [path.best_modulation]
    for mod in mods:
        n = slots_with(env, path, mod)
        if not env.is_path_free(path, a[1], n):
            break
        se = mod.spectral_efficiency
        if adaptive and not gate.alone_ok(path, a[1], n, se):
            continue
        if gate.ok(path, 0, a[1], n, se):
            keep = path.best_modulation
            path.best_modulation = mod
            try:
                svc.gn_se = se
                return env.step(a), mod
            finally:
                path.best_modulation = keep
    return env.step(reject), None


def rmcsa_heuristic(env, gate, adaptive):
    """heuristic 13 on RMCSA-v0 (walk of rmcsa_env.py:882-911 with the gate)"""
    from optical_rl_gym.utils import get_best_modulation_format

    svc = env.current_service
    k, S_, C_ = env.k_paths, env.num_spectrum_resources, env.num_spatial_resources
    gate.last = [np.nan, np.nan]
    for p, path in enumerate(env.k_shortest_paths[svc.source, svc.destination]):
        if adaptive:
            mods = mods_desc(env)
        else:
            mods = [get_best_modulation_format(path.length, env.modulation_formats)]
        for mod in mods:
            n = env.get_number_slots(path, mod)
            core, s = -1, -1
            for c in range(C_):
                for s0 in range(0, S_ - n):  # rmcsa_env.py:900-902 (F8)
                    if env.is_path_free(path, c, s0, n):
                        core, s = c, s0
                        break
                if s >= 0:
                    break
            if s < 0:
                break
            if not env._crosstalk_is_acceptable(mod, path.length):
                continue
            se = mod.spectral_efficiency
            if adaptive and not gate.alone_ok(path, s, n, se):
                continue
            if gate.ok(path, core, s, n, se):
                return [p, env.modulation_formats.index(mod), core, s]
    return [k, len(env.modulation_formats), C_, S_]


def rmcsa_gated_step(env, gate, a):
    k, S_, C_ = env.k_paths, env.num_spectrum_resources, env.num_spatial_resources
    M = len(env.modulation_formats)
    svc = env.current_service
    reject = [k, M, C_, S_]
    gate.last = [np.nan, np.nan]
    if not (0 <= a[0] < k and 0 <= a[1] < M and 0 <= a[2] < C_ and 0 <= a[3] < S_):
        return env.step(reject)
    path = env.k_shortest_paths[svc.source, svc.destination][a[0]]
    mod = env.modulation_formats[a[1]]
    n = env.get_number_slots(path, mod)
    if (env.is_path_free(path, a[2], a[3], n) and env._crosstalk_is_acceptable(mod, path.length)
            and gate.ok(path, a[2], a[3], n, mod.spectral_efficiency)):
        svc.modulation_se = mod.spectral_efficiency
        return env.step(a)
    return env.step(reject)


class AdaptiveScene(GG.Scene):
    """RMSA-v0 services remember the spectral efficiency the gate gave them (gn_se)."""

    def running(self):
        out = []
        for i, s in enumerate(self.env.topology.graph["running_services"]):
            if self.kind == 0:
                se = getattr(s, "gn_se", None) or s.path.best_modulation.spectral_efficiency
            else:
                se = s.modulation_se
            out.append(self.shadow(1000 + i, s.path, s.core if self.kind == 2 else 0, s.initial_slot,
                                   s.number_slots, se))
        return out


def run(name, env_id, topo_file, kwargs, adaptive, propose, n_steps):
    topo = refenv.load_topology(topo_file)
    env = gym.make(env_id, topology=topo, **kwargs)
    kind = {"RMSA-v0": 0, "RMCSA-v0": 2}[env_id]
    tables = tables_from_topology(topo)
    gp = S.gsnr_params(tables)
    gate = Gate(env, kind, gp)
    gate.scene = AdaptiveScene(env, gp, kind)
    adim = 2 if kind == 0 else 4
    reqs = np.zeros(n_steps + 1, GG.REQ_DTYPE)
    actions = np.zeros((n_steps, adim), np.int32)
    accepted = np.zeros(n_steps, np.int8)
    slots = np.zeros(n_steps, np.int32)
    rewards = np.zeros(n_steps, np.int32)
    dones = np.zeros(n_steps, np.int8)
    gs = np.full((n_steps, 2), np.nan)
    hgs = np.full((n_steps, 2), np.nan)
    reqs[0] = GG.req_of(env)
    for t in range(n_steps):
        if propose == "gn":  # heuristic 13
            a = rmsa_heuristic(env, gate, adaptive) if kind == 0 else rmcsa_heuristic(env, gate, adaptive)
            hgs[t] = gate.last
        else:  # the reference's own first-fit heuristic proposes, the gate decides
            a = list(propose(env))
            if kind == 2 and len(a) == 3:
                a = [env.k_paths, len(env.modulation_formats), env.num_spatial_resources, env.num_spectrum_resources]
        actions[t] = a
        svc = env.current_service
        if kind == 0:
            out, _ = rmsa_gated_step(env, gate, adaptive, a)
        else:
            out = rmcsa_gated_step(env, gate, a)
        gs[t] = gate.last
        accepted[t] = int(svc.accepted)
        slots[t] = int(svc.number_slots) if svc.accepted else 0
        rewards[t] = int(out[1])
        dones[t] = int(out[2])
        reqs[t + 1] = GG.req_of(env)
        if out[2]:
            env.reset()
    final = np.asarray(env.topology.graph["available_slots"]).astype(np.uint8)
    res = dict(env_kind=kind, topo_file=topo_file, requests=reqs, actions=actions, accepted=accepted, slots=slots,
               rewards=rewards, dones=dones, gsnr=gs, heur_gsnr=hgs, final_state=final, adaptive=int(adaptive),
               by_heuristic=int(propose == "gn"),
               kwargs_keys=np.array(sorted(k_ for k_, v in kwargs.items() if v is not None)))
    for key, v in kwargs.items():
        if v is not None:
            res["kw_" + key] = np.asarray(v)
    np.savez_compressed(os.path.join(GG.GOLDEN, name + ".npz"), **res)
    ev = np.isfinite(gs[:, 0])
    print(f"{name}: {n_steps} steps, accepted {accepted.mean():.3f}, gate evaluated {ev.mean():.3f}, "
          f"blocked by the gate {(ev & (accepted == 0)).mean():.3f}")


def main():
    us, jpn = "us14_3-paths_6-modulations.h5", "jpn12_5-paths_6-modulations.h5"
    from optical_rl_gym.envs import rmcsa_env, rmsa_env

    kw_us = dict(seed=61, load=50, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=200)
    run("qot_rmsa_us14_gate", "RMSA-v0", us, kw_us, False, rmsa_env.shortest_available_path_first_fit, 700)
    run("qot_rmsa_us14_adaptive", "RMSA-v0", us, dict(kw_us, seed=62), True, "gn", 700)
    kw_j = dict(seed=63, load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                num_spatial_resources=7, worst_xt=-84.7, episode_length=200)
    run("qot_rmcsa_jpn12_adaptive", "RMCSA-v0", jpn, kw_j, True, "gn", 500)
    run("qot_rmcsa_jpn12_gate", "RMCSA-v0", jpn, dict(kw_j, seed=64), False,
        rmcsa_env.shortest_available_path_best_modulation_first_core_first_fit, 500)


if __name__ == "__main__":
    main()
