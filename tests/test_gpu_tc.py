"""The tcgen05 sequence kernel (csrc/klm_seq_tc.cu): building-block self-test and parity with the
oracle for 9..64 concurrent streams (same tolerances as test_gpu_parity.py)."""
import logging

import numpy as np
import pytest

from oracle.rater_oracle import OracleRater
from ocrd_keraslm_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu

quiet = logging.getLogger('gpu-test-tc')
quiet.setLevel(logging.CRITICAL)


def product(wd, maparr, length, stateful=True, incremental=False):
    from ocrd_keraslm_b200.lib import Rater
    return Rater.from_weights(wd, maparr, length, stateful=stateful, incremental=incremental, logger=quiet)


def logp_err(got, ref):
    got = np.log(np.maximum(np.asarray(got, dtype=np.float64), 1e-300))
    ref = np.log(np.maximum(np.asarray(ref, dtype=np.float64), 1e-300))
    assert got.shape == ref.shape
    return np.abs(got - ref).max()


@pytest.mark.parametrize('N,K', [(16, 16), (64, 64), (160, 128), (144, 256), (256, 64)])
def test_umma_selftest_matches_numpy(N, K):
    """tcgen05.mma with the no-swizzle K-major descriptors, bulk async copy and tcgen05.ld: exact for
    small-integer fp16 operands, fp32-accurate for random ones."""
    from ocrd_keraslm_b200 import _native
    rng = np.random.default_rng(N * 1000 + K)
    A = rng.integers(-4, 5, (128, K)).astype(np.float16)
    B = rng.integers(-4, 5, (N, K)).astype(np.float16)
    D = _native.selftest_umma(A, B)
    np.testing.assert_array_equal(D, A.astype(np.float32) @ B.astype(np.float32).T)
    A = rng.standard_normal((128, K)).astype(np.float16)
    B = rng.standard_normal((N, K)).astype(np.float16)
    D = _native.selftest_umma(A, B)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    assert np.abs(D - ref).max() <= 1e-4 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize('width,depth,n_docs,nctx', [(64, 1, 9, 1), (128, 2, 33, 1), (256, 3, 64, 3), (512, 3, 40, 1)])
def test_tc_rate_many_matches_oracle(width, depth, n_docs, nctx):
    wd = synth.weights(31 + width, width=width, depth=depth, voc_size=256, contexts=((200, 10),) * nctx, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 32)
    info = r._native.info()
    assert info['tc_enabled'] >= 1
    rng = np.random.default_rng(width + n_docs)
    sizes = rng.integers(2, 40 if width < 512 else 20, n_docs)
    sizes[1] = 1
    texts = [synth.text(3000 + i, int(s)) for i, s in enumerate(sizes)]
    ctxs = [[int(x) for x in rng.integers(0, 200, nctx)] for _ in texts]
    got = r.rate_many(texts, ctxs)
    assert r._native.info()['tc_launches'] >= 1
    o = OracleRater(util.to_model(wd), synth.mapping_from_array(maparr), 32, logger=quiet)
    worst = 0.0
    for i, (t, c) in enumerate(zip(texts, ctxs)):
        ref, _ = o.rate2(t, c)
        worst = max(worst, logp_err(got[i], [p for _, p in ref]))
    assert worst <= 1e-3, worst


@pytest.mark.parametrize('depth,n_docs,nctx', [(1, 9, 1), (3, 64, 1), (2, 40, 2)])
def test_tc_streamed_weights_width_1024_matches_oracle(depth, n_docs, nctx):
    """Width 1024 (BASELINE config 5): the B images do not fit shared memory, klm_seq_tcs streams them through the
    operand ring (slices of up to 22 pairs, two epilogue rounds, split hi / lo MMAs).  Ragged lengths, carried
    state out of the launch, same tolerance as the resident form."""
    wd = synth.weights(77 + depth, width=1024, depth=depth, voc_size=256, contexts=((200, 10),) * nctx, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 32)
    assert r._native.info()['tc_enabled'] == (3 if depth > 1 else 1)      # one layer of width 1024 still fits resident
    rng = np.random.default_rng(1024 + n_docs)
    sizes = rng.integers(2, 14, n_docs)
    sizes[1] = 1
    texts = [synth.text(3500 + i, int(s)) for i, s in enumerate(sizes)]
    ctxs = [[int(x) for x in rng.integers(0, 200, nctx)] for _ in texts]
    before = r._native.info()['tc_launches']
    got = r.rate_many(texts, ctxs)
    assert r._native.info()['tc_launches'] > before
    o = OracleRater(util.to_model(wd), synth.mapping_from_array(maparr), 32, logger=quiet)
    worst = 0.0
    for i in sorted({0, 1, 2, n_docs // 2, n_docs - 1}):
        ref, _ = o.rate2(texts[i], ctxs[i])
        worst = max(worst, logp_err(got[i], [p for _, p in ref]))
    assert worst <= 1e-3, worst


@pytest.mark.parametrize('width,depth,n_docs', [(512, 4, 64), (768, 2, 30), (1536, 2, 12)])
def test_tc_streamed_weights_other_widths_match_oracle(width, depth, n_docs):
    """Models that do not fit the resident plan at other widths (ring geometries of 8, 12, 24 chunks)."""
    wd = synth.weights(13 + width, width=width, depth=depth, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 32)
    assert r._native.info()['tc_enabled'] == 3
    rng = np.random.default_rng(width + depth)
    sizes = rng.integers(2, 12, n_docs)
    texts = [synth.text(3700 + i, int(s)) for i, s in enumerate(sizes)]
    ctxs = [[int(rng.integers(0, 200))] for _ in texts]
    before = r._native.info()['tc_launches']
    got = r.rate_many(texts, ctxs)
    assert r._native.info()['tc_launches'] > before
    o = OracleRater(util.to_model(wd), synth.mapping_from_array(maparr), 32, logger=quiet)
    worst = 0.0
    for i in sorted({0, 1, n_docs // 2, n_docs - 1}):
        ref, _ = o.rate2(texts[i], ctxs[i])
        worst = max(worst, logp_err(got[i], [p for _, p in ref]))
    assert worst <= 1e-3, worst


def test_tc_states_and_carry_match_oracle():
    """Hidden and cell states after a launch (1e-4 relative) and a second launch continuing from them."""
    width, depth, n = 128, 2, 12
    wd = synth.weights(5, width=width, depth=depth, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 16)
    nat = r._native
    texts = [synth.text(500 + i, 30 + i) for i in range(n)]
    slots = r._pool.acquire(n)
    nat.state_reset(slots)
    oracles = [OracleRater(util.to_model(wd), synth.mapping_from_array(maparr), 16, logger=quiet) for _ in range(n)]
    for piece in (slice(0, 17), slice(16, None)):          # second piece continues from the carried state
        idx = [r._indices(t[piece])[0] for t in texts]
        lens = np.array([len(i) - 1 for i in idx])
        offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        chars = np.concatenate([i[:-1] for i in idx]).astype(np.int32)
        targets = np.concatenate([i[1:] for i in idx]).astype(np.int32)
        logp = np.empty(len(chars), dtype=np.float32)
        nat.rate(slots, offsets, chars, targets, None, [[7]] * n, logp)
        for k in range(n):
            # oracle: same steps without reset / padding
            o = oracles[k]
            from oracle import lstm_numpy as ln
            x = idx[k][:-1][None, :].astype(np.int64)
            z = [np.full_like(x, 7)]
            out, o.states = ln.run_sequence(o.w, x, z, o.states, want='target', targets=idx[k][1:][None, :].astype(np.int64))
            seg = logp[offsets[k]:offsets[k + 1]]
            assert np.abs(seg - np.log(out[0].astype(np.float64))).max() <= 1e-3
        got = nat.state_to_host(slots)
        for k in range(n):
            ref = np.concatenate([s_ for s_ in oracles[k].states], axis=0)
            assert np.abs(got[k] - ref).max() <= 1e-4 * np.abs(ref).max()


def test_tc_step_full_rows_match_fma():
    """klm_step through the tensor-core kernel (n = 40 hypotheses, full softmax rows) against the
    FMA K-parallel kernel (n <= 8) on the same inputs."""
    wd = synth.weights(77, width=128, depth=2, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    cands = [chr(97 + i % 26) for i in range(40)]
    p1, s1 = r.predict(list(cands), [None] * 40, [3])
    p2, s2 = r.predict(list(cands), list(s1), [3])
    for lo in range(0, 40, 8):
        q1, t1 = r.predict(cands[lo:lo + 8], [None] * 8, [3])
        q2, t2 = r.predict(cands[lo:lo + 8], list(t1), [3])
        assert logp_err(np.stack(p1[lo:lo + 8]), np.stack(q1)) <= 1e-3
        assert logp_err(np.stack(p2[lo:lo + 8]), np.stack(q2)) <= 1e-3
        a = np.stack([np.concatenate(list(s), 0) for s in s2[lo:lo + 8]])
        b = np.stack([np.concatenate(list(s), 0) for s in t2])
        assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max()
    np.testing.assert_allclose(np.stack(p2).sum(axis=1), 1.0, rtol=1e-5)


def test_tcs_step_full_rows_and_states_match_fma():
    """klm_step (Rater.predict: one tick, full softmax rows, states handed back) through the streamed-weights kernel at
    width 1024 against the FMA K-parallel kernel (n <= 8) on the same inputs."""
    wd = synth.weights(79, width=1024, depth=2, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    assert r._native.info()['tc_enabled'] == 3
    cands = [chr(97 + i % 26) for i in range(24)]
    p1, s1 = r.predict(list(cands), [None] * 24, [3])
    p2, s2 = r.predict(list(cands), list(s1), [3])
    assert r._native.info()['tc_launches'] >= 2
    for lo in range(0, 24, 8):
        q1, t1 = r.predict(cands[lo:lo + 8], [None] * 8, [3])
        q2, t2 = r.predict(cands[lo:lo + 8], list(t1), [3])
        assert logp_err(np.stack(p1[lo:lo + 8]), np.stack(q1)) <= 1e-3
        assert logp_err(np.stack(p2[lo:lo + 8]), np.stack(q2)) <= 1e-3
        a = np.stack([np.concatenate(list(s), 0) for s in s2[lo:lo + 8]])
        b = np.stack([np.concatenate(list(s), 0) for s in t2])
        assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max()
    np.testing.assert_allclose(np.stack(p2).sum(axis=1), 1.0, rtol=1e-5)


def test_umma2_selftest_cta_pair_layout():
    """tcgen05.mma.cta_group::2 (M = 128 over a CTA pair): operand split, multicast commit and the 2x2 accumulator layout."""
    from ocrd_keraslm_b200 import _native
    rng = np.random.default_rng(3)
    for N, K in [(32, 16), (160, 128), (256, 64)]:
        A = rng.integers(-4, 5, (128, K)).astype(np.float16)
        B = rng.integers(-4, 5, (N, K)).astype(np.float16)
        np.testing.assert_array_equal(_native.selftest_umma2(A, B), A.astype(np.float32) @ B.astype(np.float32).T)


@pytest.mark.parametrize('width,depth,nctx', [(128, 2, 1), (256, 3, 3), (512, 3, 1)])
def test_tc_skew3_early_signal_matches_skew2_and_oracle(width, depth, nctx, monkeypatch):
    """Launches of >= 64 positions run with skew 3 (IN / PROJ CTAs signal the grid barrier before their epilogue, Z and
    the logits are consumed two ticks after the GEMM, 4 slots): bit-equal to the skew 2 schedule on ragged streams
    (incl. empty and one-position ones) over two launches with carried state, final states equal, and oracle-close."""
    from oracle import lstm_numpy as ln
    n = 37
    wd = synth.weights(77 + width, width=width, depth=depth, voc_size=256, contexts=((200, 10),) * nctx, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 64)
    nat = r._native
    assert nat.info()['tc_enabled'] == 1
    nat.set_option('tc_df', 0)                                  # the grid-barrier kernel (the dataflow kernel has no skew)
    rng = np.random.default_rng(width)
    sizes = rng.integers(2, 200, n)
    sizes[0], sizes[1], sizes[2] = 199, 1, 2                   # the longest decides the schedule; empty and single positions
    texts = [synth.text(9000 + i, int(s)) for i, s in enumerate(sizes)]
    ctx = [[int(x) for x in rng.integers(0, 200, nctx)] for _ in range(n)]
    slots = r._pool.acquire(n)
    results = {}
    for skew in ('2', '3'):
        nat.set_option('tc_skew', int(skew))
        nat.state_reset(slots)
        outs = []
        for piece in (slice(0, 90), slice(89, None)):          # second launch continues from the carried state
            idx = [r._indices(t[piece])[0] if len(t[piece]) else np.zeros(1, dtype=np.int32) for t in texts]
            lens = np.array([max(len(i) - 1, 0) for i in idx])
            offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
            chars = np.concatenate([i[:-1] for i in idx]).astype(np.int32)
            targets = np.concatenate([i[1:] for i in idx]).astype(np.int32)
            logp = np.full(len(chars), np.nan, dtype=np.float32)
            nat.rate(slots, offsets, chars, targets, None, ctx, logp)
            outs.append((idx, offsets, logp))
        results[skew] = (outs, nat.state_to_host(slots))
    nat.set_option('tc_skew', 0)
    (o2, s2), (o3, s3) = results['2'], results['3']
    for (_, _, a), (_, _, b) in zip(o2, o3):
        assert not np.isnan(b).any()
        np.testing.assert_array_equal(a, b)
    np.testing.assert_array_equal(s2, s3)
    # oracle on a few streams (fp64), both launches chained
    m = util.to_model(wd).astype(np.float64)
    for k in (0, 2, 5, 11):
        states = ln.zero_states(m, 1, np.float64)
        for idx, offsets, logp in o3:
            if len(idx[k]) < 2:
                continue
            x = idx[k][:-1][None, :].astype(np.int64)
            z = [np.full_like(x, c) for c in ctx[k]]
            out, states = ln.run_sequence(m, x, z, states, want='target', targets=idx[k][1:][None, :].astype(np.int64))
            assert np.abs(logp[offsets[k]:offsets[k + 1]] - np.log(out[0])).max() <= 1e-3
        ref = np.concatenate(states, axis=0)
        assert np.abs(s3[k] - ref).max() <= 1e-4 * np.abs(ref).max()


@pytest.mark.parametrize('width,depth,nctx,n', [(64, 1, 1, 9), (128, 2, 1, 37), (256, 3, 3, 64), (384, 2, 1, 20), (512, 3, 1, 64)])
def test_dataflow_kernel_matches_barrier_kernel_and_oracle(width, depth, nctx, n):
    """klm_seq_df (a flag per hand-over, no grid barrier, partial sums of Z / logits added by the consumer) against
    klm_seq_tc (one grid barrier per tick) on ragged streams incl. empty and single positions, over two launches with
    carried state: scores and final states agree to rounding, both are oracle-close; the hand-over words survive the
    launches in between (running epochs, no reset)."""
    from oracle import lstm_numpy as ln
    wd = synth.weights(177 + width, width=width, depth=depth, voc_size=256, contexts=((200, 10),) * nctx, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 64)
    nat = r._native
    info = nat.info()
    assert info['tc_enabled'] == 1 and info['tc_dataflow'] == 1
    rng = np.random.default_rng(width + n)
    sizes = rng.integers(2, 300, n)
    sizes[0], sizes[1], sizes[2] = 299, 1, 2
    texts = [synth.text(9500 + i, int(s)) for i, s in enumerate(sizes)]
    ctx = [[int(x) for x in rng.integers(0, 200, nctx)] for _ in range(n)]
    slots = r._pool.acquire(n)
    results = {}
    for df in (1, 0, 1):                                         # dataflow, barrier, dataflow again (epochs keep running)
        nat.set_option('tc_df', 2 if df else 0)                  # (2: the dataflow kernel for sequences, too)
        nat.state_reset(slots)
        outs = []
        for piece in (slice(0, 120), slice(119, None)):          # second launch continues from the carried state
            idx = [r._indices(t[piece])[0] if len(t[piece]) else np.zeros(1, dtype=np.int32) for t in texts]
            lens = np.array([max(len(i) - 1, 0) for i in idx])
            offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
            chars = np.concatenate([i[:-1] for i in idx]).astype(np.int32)
            targets = np.concatenate([i[1:] for i in idx]).astype(np.int32)
            logp = np.full(len(chars), np.nan, dtype=np.float32)
            nat.rate(slots, offsets, chars, targets, None, ctx, logp)
            outs.append((idx, offsets, logp))
        if df in results:                                        # the second dataflow run: bit-identical to the first
            for (_, _, a), (_, _, b) in zip(results[df][0], outs):
                np.testing.assert_array_equal(a, b)
            np.testing.assert_array_equal(results[df][1], nat.state_to_host(slots))
        results[df] = (outs, nat.state_to_host(slots))
    nat.set_option('tc_df', 1)
    (o1, s1), (o0, s0) = results[1], results[0]
    for (_, _, a), (_, _, b) in zip(o1, o0):
        assert not np.isnan(a).any()
        assert np.abs(a.astype(np.float64) - b).max() <= 1e-4     # (two fp32-level evaluation orders)
    assert np.abs(s1 - s0).max() <= 1e-4 * max(np.abs(s0).max(), 1.0)
    m = util.to_model(wd).astype(np.float64)
    for k in (0, 2, 5, min(n - 1, 11)):
        states = ln.zero_states(m, 1, np.float64)
        for idx, offsets, logp in o1:
            if len(idx[k]) < 2:
                continue
            x = idx[k][:-1][None, :].astype(np.int64)
            z = [np.full_like(x, c) for c in ctx[k]]
            out, states = ln.run_sequence(m, x, z, states, want='target', targets=idx[k][1:][None, :].astype(np.int64))
            assert np.abs(logp[offsets[k]:offsets[k + 1]] - np.log(out[0])).max() <= 1e-3
        ref = np.concatenate(states, axis=0)
        assert np.abs(s1[k] - ref).max() <= 1e-4 * np.abs(ref).max()


@pytest.mark.parametrize('n', [9, 64, 65, 117, 200, 256, 300])
def test_dataflow_step_mode_matches_fma_step(n):
    """klm_step with targets and with full rows through the dataflow kernel's step mode (one launch for up to 256
    hypotheses, states through the pool, ragged group sizes, zero and carried input states) against the FP32 FMA kernels
    advancing the same hypotheses eight at a time."""
    wd = synth.weights(61, width=256, depth=3, voc_size=200, contexts=((200, 10),), style='trained')
    maparr = synth.vocabulary()[:200]
    r = product(wd, maparr, 32, stateful=False, incremental=True)
    nat = r._native
    assert nat.info()['tc_dataflow'] == 1
    rng = np.random.default_rng(n)
    V = 200
    prev = r._pool.acquire(n)
    nat.state_reset(prev)
    # a first step from zero states (slot -1 for some), then a second one from the states of the first
    ref_states = None
    for step in range(2):
        chars = rng.integers(0, V, n).astype(np.int32)
        targets = rng.integers(0, V, n).astype(np.int32)
        ctx = rng.integers(0, 200, (n, 1)).astype(np.int32)
        s_in = np.array(prev, dtype=np.int32)
        if step == 0:
            s_in[::3] = -1
        out_a, out_b = r._pool.acquire(n), r._pool.acquire(n)
        got = np.full(n, np.nan, dtype=np.float32)
        nat.step(chars, ctx, s_in, out_a, targets, got)
        full = np.full((n, V), np.nan, dtype=np.float32)
        nat.step(chars, ctx, s_in, out_b, None, full)
        launches = nat.info()['tc_launches']
        want = np.empty(n, dtype=np.float32)
        out_c = r._pool.acquire(n)
        for k in range(0, n, 8):                                 # <= 8 hypotheses per call: the FMA kernels
            sl = slice(k, min(k + 8, n))
            tmp = np.empty(sl.stop - sl.start, dtype=np.float32)
            nat.step(chars[sl], ctx[sl], s_in[sl], out_c[sl], targets[sl], tmp)
            want[sl] = tmp
        assert nat.info()['tc_launches'] == launches
        assert not np.isnan(got).any() and not np.isnan(full).any()
        assert np.abs(got.astype(np.float64) - want).max() <= 1e-4
        assert np.abs(np.log(full[np.arange(n), targets].astype(np.float64)) - want).max() <= 1e-4
        np.testing.assert_allclose(full.sum(axis=1), 1.0, rtol=1e-5)
        sa, sb, sc = nat.state_to_host(out_a), nat.state_to_host(out_b), nat.state_to_host(out_c)
        np.testing.assert_array_equal(sa, sb)
        assert np.abs(sa - sc).max() <= 1e-4 * np.abs(sc).max()
        prev = out_a


@pytest.mark.parametrize('width,depth,n_docs,nctx', [(1024, 3, 128, 1), (1024, 2, 97, 2), (512, 4, 65, 1), (768, 2, 100, 1)])
def test_tw_wide_streamed_kernel_matches_narrow_kernel_and_oracle(width, depth, n_docs, nctx):
    """klm_seq_tw (65..128 streams per launch of a streamed-weights plan: a row of the accumulator is a stream, three MMAs
    per k-step into the same columns) against klm_seq_tcs on the same input and against the oracle: ragged lengths incl.
    one-character texts, several context variables, ONE launch where the narrow kernel needs two."""
    wd = synth.weights(31 + width + depth, width=width, depth=depth, voc_size=256, contexts=((200, 10),) * nctx, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 32)
    nat = r._native
    assert nat.info()['tc_enabled'] == 3
    rng = np.random.default_rng(width + n_docs)
    sizes = rng.integers(2, 40, n_docs)
    sizes[1] = 1
    sizes[n_docs - 1] = 57
    texts = [synth.text(4100 + i, int(s)) for i, s in enumerate(sizes)]
    ctxs = [[int(x) for x in rng.integers(0, 200, nctx)] for _ in texts]
    l0 = nat.info()['tc_launches']
    wide = r.rate_many(texts, ctxs)
    l1 = nat.info()['tc_launches']
    nat.set_option('tc_wide', 0)
    narrow = r.rate_many(texts, ctxs)
    l2 = nat.info()['tc_launches']
    nat.set_option('tc_wide', 1)
    assert l1 - l0 == 1 and l2 - l1 in (1, 2)          # (a remainder of <= 8 streams goes to the FMA kernel)
    assert max(logp_err(a, b) for a, b in zip(wide, narrow)) <= 1e-4
    o = OracleRater(util.to_model(wd), synth.mapping_from_array(maparr), 32, logger=quiet)
    worst = 0.0
    for i in sorted({0, 1, 63, 64, n_docs // 2, n_docs - 1}):
        ref, _ = o.rate2(texts[i], ctxs[i])
        worst = max(worst, logp_err(wide[i], [p for _, p in ref]))
    assert worst <= 1e-3, worst


def test_tw_states_carry_over_launches_like_the_narrow_kernel():
    """Carried state in, final state out of klm_seq_tw: two launches over 100 streams continue from the pool slots; scores
    and states equal those of klm_seq_tcs (which the oracle tests above pin) within the state tolerance."""
    width, depth, n = 1024, 2, 100
    wd = synth.weights(8, width=width, depth=depth, voc_size=256, contexts=((200, 10),), style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 16)
    nat = r._native
    texts = [synth.text(900 + i, 20 + i % 23) for i in range(n)]
    results = {}
    for mode in (1, 0):
        nat.set_option('tc_wide', mode)
        slots = r._pool.acquire(n)
        nat.state_reset(slots)
        logps = []
        for piece in (slice(0, 9), slice(8, None)):          # second piece continues from the carried state
            idx = [r._indices(t[piece])[0] for t in texts]
            lens = np.array([len(i) - 1 for i in idx])
            offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
            chars = np.concatenate([i[:-1] for i in idx]).astype(np.int32)
            targets = np.concatenate([i[1:] for i in idx]).astype(np.int32)
            logp = np.full(len(chars), np.nan, dtype=np.float32)
            nat.rate(slots, offsets, chars, targets, None, [[(3 * k) % 200] for k in range(n)], logp)
            logps.append(logp)
        results[mode] = (np.concatenate(logps), nat.state_to_host(slots))
        r._pool.release_many(slots)
    nat.set_option('tc_wide', 1)
    assert not np.isnan(results[1][0]).any()
    assert np.abs(results[1][0].astype(np.float64) - results[0][0]).max() <= 1e-4
    assert np.abs(results[1][1] - results[0][1]).max() <= 1e-4 * np.abs(results[0][1]).max()


def test_tw_padding_positions_take_context_zero_like_the_narrow_kernel():
    """n_real (window padding of Rater.rate, rating.py:1141-1143: positions >= n_real[s] see context 0) through klm_seq_tw:
    90 streams with padded tails, scores equal to klm_seq_tcs on the same call."""
    width, depth, n = 1024, 2, 90
    wd = synth.weights(19, width=width, depth=depth, voc_size=256, contexts=((200, 10),), style='trained')
    r = product(wd, synth.vocabulary(), 16)
    nat = r._native
    rng = np.random.default_rng(90)
    lens = rng.integers(3, 30, n)
    n_real = np.maximum(lens - rng.integers(0, 4, n), 1).astype(np.int32)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    chars = rng.integers(1, 200, int(lens.sum())).astype(np.int32)
    targets = rng.integers(1, 200, int(lens.sum())).astype(np.int32)
    ctx = [[int(c)] for c in rng.integers(1, 200, n)]
    out = {}
    for mode in (1, 0):
        nat.set_option('tc_wide', mode)
        slots = r._pool.acquire(n)
        nat.state_reset(slots)
        logp = np.full(len(chars), np.nan, dtype=np.float32)
        before = nat.info()['tc_wide_launches']
        nat.rate(slots, offsets, chars, targets, n_real, ctx, logp)
        assert nat.info()['tc_wide_launches'] - before == (1 if mode else 0)
        out[mode] = logp
        r._pool.release_many(slots)
    nat.set_option('tc_wide', 1)
    assert not np.isnan(out[1]).any()
    assert np.abs(out[1].astype(np.float64) - out[0]).max() <= 1e-4


def test_tw_rows_rate_queues_of_texts_in_one_launch():
    """Continuous batching of klm_seq_tw: 700 texts (some empty, some of one character) queued on the 128 rows of ONE launch
    -- a gap tick between two texts of a row swaps the state -- against klm_seq_tcs in 11 launches: scores and the final
    state of every text (each in its own pool slot), plus the oracle on a sample."""
    width, depth, n = 1024, 2, 700
    wd = synth.weights(23, width=width, depth=depth, voc_size=256, contexts=((200, 10),), style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 16)
    nat = r._native
    rng = np.random.default_rng(700)
    lens = rng.integers(0, 25, n)
    lens[[3, 77, 130]] = 0
    lens[[4, 500]] = 1
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    chars = rng.integers(1, 200, int(lens.sum())).astype(np.int32)
    targets = rng.integers(1, 200, int(lens.sum())).astype(np.int32)
    ctx = [[int(c)] for c in rng.integers(0, 200, n)]
    res = {}
    for mode in (1, 0):
        nat.set_option('tc_wide', mode)
        slots = r._pool.acquire(n)
        nat.state_reset(slots)
        logp = np.full(len(chars), np.nan, dtype=np.float32)
        w0, t0 = nat.info()['tc_wide_launches'], nat.info()['tc_launches']
        nat.rate(slots, offsets, chars, targets, None, ctx, logp)
        if mode:
            assert nat.info()['tc_wide_launches'] - w0 == 1 and nat.info()['tc_launches'] - t0 == 1
        res[mode] = (logp, nat.state_to_host(slots))
        r._pool.release_many(slots)
    nat.set_option('tc_wide', 1)
    assert not np.isnan(res[1][0]).any()
    assert np.abs(res[1][0].astype(np.float64) - res[0][0]).max() <= 1e-4
    assert np.abs(res[1][1] - res[0][1]).max() <= 1e-4 * np.abs(res[0][1]).max()
    assert (res[1][1][[3, 77, 130]] == 0).all()                      # empty texts leave their (zero) state alone
    from oracle import lstm_numpy as ln
    w = util.to_model(wd)
    for k in (0, 4, 129, 699):
        if lens[k] == 0:
            continue
        x = chars[offsets[k]:offsets[k + 1]][None, :].astype(np.int64)
        out, _ = ln.run_sequence(w, x, [np.full_like(x, ctx[k][0])], ln.zero_states(w, 1), want='target',
                                 targets=targets[offsets[k]:offsets[k + 1]][None, :].astype(np.int64))
        assert np.abs(res[1][0][offsets[k]:offsets[k + 1]] - np.log(out[0].astype(np.float64))).max() <= 1e-3


@pytest.mark.parametrize('n', [100, 300])
def test_tw_step_full_rows_and_states_match_fma(n):
    """klm_step (Rater.predict: one position, full softmax rows, states handed back through other pool slots) for more
    than 64 hypotheses on a streamed-weights model: one launch of klm_seq_tw (300 hypotheses queue up on its 128 rows)
    against the FMA K-parallel kernel (groups of 8) on the same inputs, two steps deep."""
    wd = synth.weights(83, width=1024, depth=2, voc_size=256, style='trained')
    maparr = synth.vocabulary()
    r = product(wd, maparr, 8, stateful=False, incremental=True)
    nat = r._native
    assert nat.info()['tc_enabled'] == 3 and nat.info()['tc_wide'] == 1
    cands = [chr(97 + (i * 7) % 26) for i in range(n)]
    w0 = nat.info()['tc_wide_launches']
    p1, s1 = r.predict(list(cands), [None] * n, [3])
    p2, s2 = r.predict(list(cands[::-1]), list(s1), [3])
    assert nat.info()['tc_wide_launches'] - w0 == 2
    np.testing.assert_allclose(np.stack(p2).sum(axis=1), 1.0, rtol=1e-5)
    for lo in (0, 64, n - 8):
        q1, t1 = r.predict(cands[lo:lo + 8], [None] * 8, [3])
        q2, t2 = r.predict(cands[::-1][lo:lo + 8], list(t1), [3])
        assert logp_err(np.stack(p1[lo:lo + 8]), np.stack(q1)) <= 1e-3
        assert logp_err(np.stack(p2[lo:lo + 8]), np.stack(q2)) <= 1e-3
        a = np.stack([np.concatenate(list(s), 0) for s in s2[lo:lo + 8]])
        b = np.stack([np.concatenate(list(s), 0) for s in t2])
        assert np.abs(a - b).max() <= 1e-4 * np.abs(b).max()
