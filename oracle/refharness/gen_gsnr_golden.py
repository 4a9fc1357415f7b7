"""GSNR golden vectors from the reference's own calculate_osnr (TEST INFRASTRUCTURE).

    python oracle/refharness/gen_gsnr_golden.py      # writes tests/golden/gsnr_*.npz

`examples/calculate_osnr.py` does `from optical_rl_gym.utils import Service, Span, Link`, but
`Span`/`Link` are commented out in the reference, so the import fails as
shipped.  This script injects the two dataclasses (fields of the commented-out sketch,
optical_rl_gym/utils.py:38-54, plus the `*_normalized` attributes the function reads) into
`optical_rl_gym.utils` and then imports the UNMODIFIED file, so the numbers below come from
the reference's own function body.

Scenario: a reference RMSA-v0 / RMCSA-v0 env is driven by its own heuristic; every few steps
the heuristic's candidate (path, first slot, slots) for the pending request is evaluated
against the env's live services:
  out[0] = calculate_osnr(candidate | running services on its links)
  out[1] = min over neighbours j (running services sharing a link and core with the candidate)
           of calculate_osnr(j | the other running services on j's links + the candidate)
The objects handed to calculate_osnr are duck-typed shadows with unique service ids (the
reference's own ids repeat across episodes, F7) carrying the physical attributes the reference
never defines (bandwidth, center_frequency, launch_power; spec.GsnrParams states them).
The request/action trace is stored too, so the CUDA engine and the C oracle can replay the
run to the same network state and be compared at the same steps.
"""
import importlib.util
import os
import sys
from dataclasses import dataclass
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, REPO)

import refenv  # noqa: E402

gym = refenv.install()

import optical_rl_gym.utils as ref_utils  # noqa: E402


@dataclass
class Span:  # optical_rl_gym/utils.py:38-44 (commented out there)
    length: float
    attenuation: float
    default_attenuation: float
    noise_figure: float
    attenuation_normalized: float = 0.0
    noise_figure_normalized: float = 0.0


@dataclass
class Link:  # optical_rl_gym/utils.py:47-54 (commented out there)
    id: int
    length: float
    spans: tuple
    node1: str = None
    node2: str = None


ref_utils.Span = Span
ref_utils.Link = Link
_spec = importlib.util.spec_from_file_location(
    "ref_calculate_osnr", os.path.join(refenv.REFERENCE_ROOT, "examples", "calculate_osnr.py"))
ref_osnr = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(ref_osnr)

from optical_rl_gym.envs import rmcsa_env, rmsa_env  # noqa: E402

from optical_rl_gym_qot_aware_b200 import spec as S  # noqa: E402
from optical_rl_gym_qot_aware_b200.tables import tables_from_topology  # noqa: E402

GOLDEN = os.path.join(REPO, "tests", "golden")
REQ_DTYPE = np.dtype([("arrival", "<f8"), ("holding", "<f8"), ("bit_rate", "<i4"), ("src", "<i2"), ("dst", "<i2")])


def req_of(env):
    s = env.current_service
    return (float(s.arrival_time), float(s.holding_time), int(s.bit_rate), int(s.source_id), int(s.destination_id))


class Scene:
    """Shadow of the env's network for calculate_osnr: topology[n1][n2] -> {"link", "running_services"}."""

    def __init__(self, env, gp, kind):
        self.env, self.gp, self.kind = env, gp, kind
        self.links = {}
        for u, v, d in env.topology.edges(data=True):
            e = int(d["index"])
            spans = tuple(Span(length=float(gp.span_length_km[e]), attenuation=0.2, default_attenuation=0.2,
                               noise_figure=4.5, attenuation_normalized=gp.attenuation_per_m,
                               noise_figure_normalized=gp.noise_figure_lin) for _ in range(int(gp.span_count[e])))
            self.links[e] = Link(id=e, length=float(d["length"]), spans=spans, node1=u, node2=v)

    def shadow(self, uid, path, core, start, n, se):
        gp = self.gp
        nl = path.node_list
        links = [self.links[int(self.env.topology[nl[i]][nl[i + 1]]["index"])] for i in range(len(nl) - 1)]
        return SimpleNamespace(
            service_id=uid, path=SimpleNamespace(links=links), core=core,
            bandwidth=n * gp.slot_hz, center_frequency=gp.f0_hz + (start + 0.5 * n) * gp.slot_hz,
            launch_power=gp.launch_power_w, current_modulation=SimpleNamespace(spectral_efficiency=int(se)))

    def running(self):
        out = []
        for i, s in enumerate(self.env.topology.graph["running_services"]):
            se = s.path.best_modulation.spectral_efficiency if self.kind == 0 else s.modulation_se
            out.append(self.shadow(1000 + i, s.path, s.core if self.kind == 2 else 0, s.initial_slot,
                                   s.number_slots, se))
        return out

    def evaluate(self, target, services):
        """calculate_osnr(target | services), restricted to the target's core."""
        topo = {}
        for lk in self.links.values():
            lst = [s for s in services if s is not target and s.core == target.core and lk in s.path.links]
            for a, b in ((lk.node1, lk.node2), (lk.node2, lk.node1)):
                topo.setdefault(a, {})[b] = {"link": lk, "running_services": lst}
        return float(ref_osnr.calculate_osnr(SimpleNamespace(topology=topo), target))


def run(name, env_id, topo_file, kwargs, policy, n_steps, every):
    topo = refenv.load_topology(topo_file)
    env = gym.make(env_id, topology=topo, **kwargs)
    kind = {"RMSA-v0": 0, "RMCSA-v0": 2}[env_id]
    tables = tables_from_topology(topo)
    gp = S.gsnr_params(tables)
    scene = Scene(env, gp, kind)
    adim = 2 if kind == 0 else 4
    reqs = np.zeros(n_steps + 1, REQ_DTYPE)
    actions = np.zeros((n_steps, adim), np.int32)
    q_step, q_query, q_out = [], [], []
    reqs[0] = req_of(env)
    S_, k = env.num_spectrum_resources, env.k_paths
    for t in range(n_steps):
        a = list(policy(env))
        if kind == 2 and len(a) == 3:
            a = [k, len(env.modulation_formats), env.num_spatial_resources, S_]
        actions[t] = a
        svc = env.current_service
        if t % every == 0 and a[0] < k:
            path = env.k_shortest_paths[svc.source, svc.destination][a[0]]
            if kind == 0:
                core, start, se = 0, a[1], path.best_modulation.spectral_efficiency
                n = env.get_number_slots(path)
            else:
                mod = env.modulation_formats[a[1]]
                core, start, se = a[2], a[3], mod.spectral_efficiency
                n = env.get_number_slots(path, mod)
            running = scene.running()
            cand = scene.shadow(-1, path, core, start, n, se)
            g0 = scene.evaluate(cand, running)
            g1 = np.inf
            for j in running:
                if j.core == core and any(lk in j.path.links for lk in cand.path.links):
                    g1 = min(g1, scene.evaluate(j, running + [cand]))
            q_step.append(t)
            q_query.append((a[0] | (core << 16), start, n, se))
            q_out.append((g0, g1))
        if kind == 2:  # remember the modulation of the service about to be provisioned
            svc.modulation_se = env.modulation_formats[a[1]].spectral_efficiency if a[1] < len(
                env.modulation_formats) else 0
        out = env.step(a)
        reqs[t + 1] = req_of(env)
        if out[2]:
            env.reset()
    res = dict(env_kind=kind, topo_file=topo_file, requests=reqs, actions=actions,
               q_step=np.array(q_step, np.int32), q_query=np.array(q_query, np.int32),
               q_out=np.array(q_out, np.float64),
               kwargs_keys=np.array(sorted(k_ for k_, v in kwargs.items() if v is not None)))
    for key, v in kwargs.items():
        if v is not None:
            res["kw_" + key] = np.asarray(v)
    np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **res)
    o = np.array(q_out)
    print(f"{name}: {len(q_step)} GSNR queries, candidate {o[:, 0].min():.3f}..{o[:, 0].max():.3f} dB, "
          f"neighbour min {o[np.isfinite(o[:, 1]), 1].min():.3f} dB")


def main():
    nsf, jpn = "nsfnet_chen_5-paths_6-modulations.h5", "jpn12_5-paths_6-modulations.h5"
    run("gsnr_rmsa_nsfnet", "RMSA-v0", nsf, dict(seed=51, load=50, num_spectrum_resources=100, episode_length=150),
        rmsa_env.shortest_available_path_first_fit, 600, 7)
    run("gsnr_rmcsa_jpn12", "RMCSA-v0", jpn,
        dict(seed=52, load=250, mean_service_holding_time=25, num_spectrum_resources=320, num_spatial_resources=7,
             worst_xt=-84.7, episode_length=200),
        rmcsa_env.shortest_available_path_best_modulation_first_core_first_fit, 500, 9)


if __name__ == "__main__":
    main()
