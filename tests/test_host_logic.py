"""CPU-side checks: the C-ABI library loads and exports every symbol include/orlg.h declares,
the config struct layouts agree, env sharding and the episode-statistics all-reduce (gloo,
world_size 2) behave.  No compute call is made: there is no GPU here."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from optical_rl_gym_qot_aware_b200 import engine

    hdr = open(os.path.join(REPO, "include", "orlg.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(orlg_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    L = engine.load_library()
    missing = [n for n in sorted(declared) if not hasattr(L, n)]
    assert not missing, missing
    assert set(engine.EXPORTS) >= declared - {"orlg_engine"}
    assert L.orlg_abi_version() == 2
    assert b"sm_100a" in L.orlg_build_info()


def test_config_struct_matches_header_size(tmp_path):
    """sizeof/offsetof of orlg_config as gcc sees the header == the ctypes mirror."""
    from optical_rl_gym_qot_aware_b200 import engine

    src = tmp_path / "sz.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "orlg.h"\n'
        "int main(void){printf(\"%zu %zu %zu %zu %zu\\n\", sizeof(orlg_config), offsetof(orlg_config, path_links),"
        " offsetof(orlg_config, channel_width), offsetof(orlg_config, seed), sizeof(orlg_request));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(REPO, "include"), str(src), "-o", str(exe)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    cfg = engine.OrlgConfig
    assert out == [C.sizeof(cfg), cfg.path_links.offset, cfg.channel_width.offset, cfg.seed.offset,
                   engine.REQUEST_DTYPE.itemsize]


def test_engine_refuses_to_run_without_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import golden_util as G
    from optical_rl_gym_qot_aware_b200 import spec as S
    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec("RMSA-v0", G.topology("nsfnet"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine(spec, 4)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(REPO, "optical_rl_gym_qot_aware_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(root, f)).read()
                assert "oracle_binding" not in txt and "orlg_oracle" not in txt, f


def test_shard_bounds_partition():
    from optical_rl_gym_qot_aware_b200.sharding import shard_bounds

    for n in (0, 1, 7, 64, 65536, 1000003):
        for w in (1, 2, 3, 4, 8):
            b = [shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [y - x for x, y in b]
            assert max(sizes) - min(sizes) <= 1


_WORKER = r"""
import os, sys, json
sys.path.insert(0, {repo!r})
import torch, torch.distributed as dist
from optical_rl_gym_qot_aware_b200.sharding import shard_bounds, reduce_episode_stats, init_from_env
rank, world, _ = init_from_env("gloo")
n_global = 37
lo, hi = shard_bounds(n_global, world, rank)
g = torch.arange(n_global, dtype=torch.int64)
cnt = torch.zeros((n_global, 12), dtype=torch.int64)
cnt[:, 0] = 100 + g; cnt[:, 1] = 40 + (g % 7); cnt[:, 4] = 1000 * (g + 1); cnt[:, 5] = 300 * (g + 1)
cnt[:, 11] = g % 3; cnt[5, 10] = 2
rs = (g * 3 - 10)
out = reduce_episode_stats(cnt[lo:hi], rs[lo:hi])
print("RESULT", rank, json.dumps(out), flush=True)
dist.destroy_process_group()
"""


def test_stats_all_reduce_gloo_world2(tmp_path):
    """Two gloo ranks each summarise their shard; the reduced totals equal the 1-rank totals."""
    import json
    import socket

    script = tmp_path / "w.py"
    script.write_text(_WORKER.format(repo=REPO))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    res = [json.loads(o.split("RESULT", 1)[1].split(" ", 2)[2]) for o in outs]
    assert res[0] == res[1]
    g = np.arange(37)
    assert res[0]["services_processed"] == int((100 + g).sum())
    assert res[0]["services_accepted"] == int((40 + g % 7).sum())
    assert res[0]["bit_rate_requested"] == int((1000 * (g + 1)).sum())
    assert res[0]["episodes"] == int((g % 3).sum())
    assert res[0]["reward_sum"] == int((g * 3 - 10).sum())
    assert res[0]["envs"] == 37 and res[0]["fault"] == 2
    assert abs(res[0]["service_blocking_rate"] - (1 - (40 + g % 7).sum() / (100 + g).sum())) < 1e-15


def test_bench_reference_arm_and_host_policy():
    """`bench.py --impl reference` runs without a GPU (it times the oracle port) and prints the contract's line;
    the host K-SP-FF policy of the e2e loop picks the first path whose block-length feature is not -1."""
    import json
    import subprocess
    import sys

    import numpy as np

    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(repo, "bench.py"), "--impl", "reference", "--steps", "2",
                          "--warmup", "3", "--ref-envs", "64"], capture_output=True, text=True, timeout=300, cwd=repo)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "env-steps/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # the default workload is the QoT configuration BASELINE.json's metric is quoted on
    assert line["config"]["workload"] == "c6_rmsa_us14_320_gn" and line["gpu_launches"] == 0

    sys.path.insert(0, repo)
    import bench

    n, k, j, nodes = 1000, 5, 1, 14
    rng = np.random.default_rng(0)
    obs = rng.random((1 + 2 * nodes + (2 * j + 3) * k, n), dtype=np.float32)
    blocked = rng.random((k, n)) < 0.6
    for p in range(k):
        obs[1 + 2 * nodes + p * (2 * j + 3) + 1][blocked[p]] = -1.0
    want = np.array([next((p * j for p in range(k) if not blocked[p, i]), k * j) for i in range(n)])
    out_buf = np.zeros((n, 1), np.int32)
    assert np.array_equal(bench.host_policy_kspff(obs, nodes, k, j, out=out_buf)[:, 0], want)
    assert np.array_equal(bench.host_policy_kspff(obs, nodes, k, j)[:, 0], want)
