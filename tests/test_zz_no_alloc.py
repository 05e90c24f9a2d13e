"""klm_state_l2_distance_noalloc (include/keraslm_b200.h): the state distances of the beam clustering through the
STATIC device scratch, for callers that cannot allocate on the device.  No counterpart in the reference;
checked through the C-ABI against the existing klm_state_l2_distance and against
numpy on the downloaded states, on two calls in a row (the second smaller, so that contents left by the first call
would show), plus the size limit.

Collected last on purpose (file name).
"""
import logging

import numpy as np
import pytest

from ocrd_keraslm_b200 import synth

pytestmark = pytest.mark.gpu

quiet = logging.getLogger('gpu-test-noalloc')
quiet.setLevel(logging.CRITICAL)

RTOL = 1e-6         # same kernel behind both entry points; fp32 sums against numpy's fp64: 1e-5


def _rater_with_states(n_streams):
    """A width-128 / depth-2 model whose first n_streams slots hold the states after rating n_streams texts."""
    from ocrd_keraslm_b200.lib import Rater
    wd = synth.weights(31, width=128, depth=2, voc_size=256, contexts=((200, 10),), style='trained')
    r = Rater.from_weights(wd, synth.vocabulary(), 64, stateful=True, incremental=False, logger=quiet)
    nat = r._native
    slots = np.asarray(r._pool.acquire(n_streams), dtype=np.int32)
    idx = [r._indices(synth.text(900 + i, 30 + (i % 17)))[0] for i in range(n_streams)]
    lens = np.array([len(i) - 1 for i in idx])
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    chars = np.concatenate([i[:-1] for i in idx]).astype(np.int32)
    targets = np.concatenate([i[1:] for i in idx]).astype(np.int32)
    logp = np.empty(len(chars), dtype=np.float32)
    nat.state_reset(slots)
    nat.rate(slots, offsets, chars, targets, None, [[(7 * i) % 200] for i in range(n_streams)], logp)
    assert np.isfinite(logp).all()
    return r, nat, slots


def _numpy_distances(states, ia, ib, n_arrays):
    d = states[ia, :n_arrays].astype(np.float64) - states[ib, :n_arrays].astype(np.float64)
    return np.sqrt((d * d).sum(axis=-1))


def test_noalloc_distances_match_allocating_entry_point_twice():
    n = 96
    r, nat, slots = _rater_with_states(n)
    states = nat.state_to_host(slots)                              # [n, 2*depth, width]
    rng = np.random.default_rng(12)
    # first call: all four arrays (h1, c1, h2, c2) of n pairs; second call: fewer pairs, fewer arrays, other pairing
    for n_pairs, n_arrays in ((n, 4), (37, 2)):
        ia = rng.integers(0, n, n_pairs)
        ib = rng.integers(0, n, n_pairs)
        ib[0] = ia[0]                                              # one pair of identical states: distance exactly 0
        want = nat.state_l2_distance(slots[ia], slots[ib], n_arrays)
        got = np.full((n_pairs, n_arrays), np.nan, dtype=np.float32)
        nat.state_l2_distance_noalloc(slots[ia], slots[ib], n_arrays, out=got)
        assert got.shape == want.shape and np.isfinite(got).all()
        np.testing.assert_allclose(got, want, rtol=RTOL, atol=0)
        assert (got[0] == 0).all() and (got[1:] > 0).any()
        np.testing.assert_allclose(got, _numpy_distances(states, ia, ib, n_arrays), rtol=1e-5, atol=1e-7)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    from ocrd_keraslm_b200 import _native
    r, nat, slots = _rater_with_states(4)
    n_arrays = 4
    n_max = (_native.NOALLOC_SCRATCH_BYTES - 256) // (8 + 4 * n_arrays)
    big = np.resize(slots, n_max + 64).astype(np.int32)
    with pytest.raises(ValueError, match='scratch'):                     # KLM_EINVAL, the message names the sizes
        nat.state_l2_distance_noalloc(big, big, n_arrays)
    # the allocating entry point takes the same request, and the static one still works afterwards
    assert (nat.state_l2_distance(big, big, n_arrays) == 0).all()
    ok = np.resize(slots, n_max - 64).astype(np.int32)
    got = nat.state_l2_distance_noalloc(ok, np.roll(ok, 1), n_arrays)
    np.testing.assert_allclose(got, nat.state_l2_distance(ok, np.roll(ok, 1), n_arrays), rtol=RTOL, atol=0)
