"""The reference's request stream regenerated from its seed (TEST INFRASTRUCTURE).

The reference draws every request from CPython's ``random.Random(seed)`` (optical_network_env.py:266-271) with
exactly five calls (rmsa_env.py:646-659, optical_network_env.py:191-208): ``expovariate`` for the inter-arrival and
the holding time, ``choices`` for the source, ``choices`` for the destination over the weights renormalised with
numpy, and ``choices`` (discrete) or ``randint`` (continuous) for the bit rate.  Repeating those calls gives the
reference's trace bit for bit without the reference (CPython's MT19937 and its ``random`` module are stdlib), so
multi-seed goldens need to store only hashes.  Pinned by the traces stored in tests/golden/*.npz
(tests/test_multiseed.py::test_regenerated_traces_equal_the_stored_ones)."""
import random

import numpy as np

REQ_DTYPE = np.dtype([("arrival", "<f8"), ("holding", "<f8"), ("bit_rate", "<i4"), ("src", "<i2"), ("dst", "<i2")])


def reference_trace(spec, seed, n):
    rng = random.Random(seed)
    probs = np.asarray(spec.node_probs, np.float64)
    ids = list(range(len(probs)))
    out = np.zeros(n, REQ_DTYPE)
    now = 0.0
    for r in range(n):
        now = now + rng.expovariate(1 / spec.mean_iat)
        ht = rng.expovariate(1 / spec.mean_holding)
        src = rng.choices(ids, weights=probs)[0]
        newp = np.copy(probs)
        newp[src] = 0.0
        newp = newp / np.sum(newp)
        dst = rng.choices(ids, weights=newp)[0]
        if spec.bit_rates is None:
            br = rng.randint(spec.bit_rate_lo, spec.bit_rate_hi)
        else:
            nb = len(spec.bit_rates)
            w = np.diff(np.concatenate([[0.0], spec.bit_rate_cum]))  # the probabilities behind the cumulative weights
            br = rng.choices([int(b) for b in spec.bit_rates], getattr(spec, "bit_rate_probs", None) or list(w), k=1)[0]
            del nb
        out[r] = (now, ht, br, src, dst)
    return out


def fold(h, *cols):
    """rolling 64-bit hash of per-step columns (uint64 wrap-around), one value per env"""
    for c in cols:
        h = h * np.uint64(1000003) + np.asarray(c).astype(np.int64).astype(np.uint64)
    return h
