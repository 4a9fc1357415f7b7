"""orlg_step_auto_noalloc (include/orlg.h): the allocation-free variant of orlg_step_auto, whose action scratch is a
static device array of liborlg.so.  Called through the C-ABI and checked against the existing entry point on the
same Philox streams, two calls in a row (bit-exact: integer outputs, the tolerance every parity test here uses).
Collected last on purpose."""
import numpy as np
import pytest

import golden_util as G
from optical_rl_gym_qot_aware_b200 import spec as S

pytestmark = pytest.mark.gpu

CASES = [
    ("RMSA-v0", "nsfnet", dict(load=50, num_spectrum_resources=100, episode_length=60), S.H_SAP_FF, {}),
    ("RMSA-v0", "us14", dict(load=60, mean_service_holding_time=25, num_spectrum_resources=320, episode_length=60),
     S.H_SAP_GN_FF, dict(qot="gn", qot_adaptive=True)),
    ("RMCSA-v0", "jpn12_k5", dict(load=250, mean_service_holding_time=25, num_spectrum_resources=320,
                                  num_spatial_resources=7, worst_xt=-84.7, episode_length=60), S.H_SAP_BM_FC_FF, {}),
]


@pytest.mark.parametrize("env_id,topo,kw,hid,extra", CASES)
def test_no_alloc_step_equals_step_auto(env_id, topo, kw, hid, extra):
    import torch

    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec(env_id, G.topology(topo), seed=123, auto_reset=True, **kw, **extra)
    n_env = 500
    a, b = Engine(spec, n_env), Engine(spec, n_env)
    a.reset(True)
    b.reset(True)
    dev = a.device

    def bufs():
        return (torch.zeros(n_env, dtype=torch.int32, device=dev), torch.zeros(n_env, dtype=torch.uint8, device=dev),
                torch.zeros((n_env, 4), dtype=torch.int32, device=dev))

    ra, da, ala = bufs()
    rb, db, alb = bufs()
    act = torch.zeros((n_env, a.action_dim), dtype=torch.int32, device=dev)
    for call in range(80):  # "two calls in a row" and then some: the shared array keeps the last call's contents
        a.step_auto(hid, act, ra, da, None, ala)
        free0 = torch.cuda.mem_get_info(dev)[0]
        b.step_auto_noalloc(hid, rb, db, None, alb)
        torch.cuda.synchronize()
        if call >= 2:  # the first calls may load the kernels' modules; after that the device's free memory is untouched
            assert torch.cuda.mem_get_info(dev)[0] == free0
        assert torch.equal(ra, rb) and torch.equal(da, db) and torch.equal(ala, alb), call
    assert torch.equal(a.counters(), b.counters())
    assert torch.equal(a.export_state(), b.export_state())
    assert int(a.counters()[:, 1].sum()) > 0 and int(b.counters()[:, 10].max()) == 0


def test_no_alloc_rejects_engines_beyond_the_static_scratch():
    from optical_rl_gym_qot_aware_b200.engine import Engine

    spec = S.make_spec("RMSA-v0", G.topology("nsfnet"), seed=1, load=50, num_spectrum_resources=100)
    eng = Engine(spec, 262144 // 2 + 1)  # action_dim 2: one env too many
    eng.reset(True)
    with pytest.raises(RuntimeError, match="ORLG_NOALLOC_MAX_INTS"):
        eng.step_auto_noalloc(S.H_SAP_FF)
    ok = Engine(spec, 262144 // 2)
    ok.reset(True)
    ok.step_auto_noalloc(S.H_SAP_FF)
    assert int(ok.counters()[:, 0].min()) == 2
