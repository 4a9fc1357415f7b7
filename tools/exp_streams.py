"""Scratch experiment: does stepping the batch as P sub-batches on P streams beat one launch sequence?"""
import os, sys, time
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "tests")]
import golden_util as G
from optical_rl_gym_qot_aware_b200 import spec as S
from optical_rl_gym_qot_aware_b200.engine import Engine

N = int(os.environ.get("N", 65536))
spec = S.make_spec("RMSA-v0", G.topology("us14"), seed=0x5EED, auto_reset=True, qot="gn", qot_adaptive=True,
                   load=100, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=1000)
for P in (1, 2, 4):
    n = N // P
    engs = [Engine(spec, n, global_env_offset=i * n) for i in range(P)]
    sts = [torch.cuda.Stream() for _ in range(P)]
    bufs = []
    for e in engs:
        e.reset(True); e.rollout(S.H_SAP_GN_FF, 1500)
        dev = e.device
        bufs.append((torch.zeros((n, e.action_dim), dtype=torch.int32, device=dev), torch.zeros(n, dtype=torch.int32, device=dev),
                     torch.zeros(n, dtype=torch.uint8, device=dev)))
    torch.cuda.synchronize()
    K = 40
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for k in range(K):
            for e, st, (a, r, d) in zip(engs, sts, bufs):
                with torch.cuda.stream(st):
                    e.step_auto(S.H_SAP_GN_FF, a, r, d, None, None)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("streams", P, "ms/step", dt * 1e3 / K, "steps/s", N * K / dt)
    del engs
