"""Race hunt for klm_seq_df: repeat launches on the same input, count positions that differ from the grid-barrier kernel
(tolerance 1e-4: two fp32 evaluation orders) under combinations of the memory-ordering switches."""
import os, sys, logging, itertools
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ocrd_keraslm_b200 import synth
from ocrd_keraslm_b200.lib import Rater
log = logging.getLogger('x'); log.setLevel(logging.CRITICAL)

TOL = 1e-4
REPS = int(os.environ.get('DF_RACE_REPS', '6'))

def run(width, depth, n, T, opts, reps=None):
    reps = reps or REPS
    wd = synth.weights(5 + width, width=width, depth=depth, voc_size=256, style='trained')
    r = Rater.from_weights(wd, synth.vocabulary(), 64, logger=log)
    nat = r._native
    texts = [synth.text(100 + i, T) for i in range(n)]
    ctxs = [[i % 200] for i in range(n)]
    slots = r._pool.acquire(n)
    idx = [r._indices(t)[0] for t in texts]
    lens = np.array([len(i) - 1 for i in idx]); offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    chars = np.concatenate([i[:-1] for i in idx]).astype(np.int32); targets = np.concatenate([i[1:] for i in idx]).astype(np.int32)
    def once():
        out = np.full(len(chars), np.nan, dtype=np.float32)
        nat.state_reset(slots); nat.rate(slots, offsets, chars, targets, None, ctxs, out)
        return out
    nat.set_option('tc_df', 0); ref = once(); nat.set_option('tc_df', 2)
    for k, v in opts.items(): nat.set_option(k, v)
    bad = []
    first = None
    for _ in range(reps):
        o = once()
        d = np.abs(o.astype(np.float64) - ref)
        bad.append(int((d > TOL).sum()) + int(np.isnan(o).sum()))
        if first is None: first = o
    # where is the first bad position (tick) of the last run?
    d = np.abs(o.astype(np.float64) - ref) > TOL
    t_first = [int(np.argmax(d[offsets[s]:offsets[s+1]])) if d[offsets[s]:offsets[s+1]].any() else -1 for s in range(n)]
    tf = [t for t in t_first if t >= 0]
    print('W%d D%d n%d T%d %s: bad positions per run %s of %d; first bad tick min %s, streams hit %d' % (
        width, depth, n, T, opts, bad, len(chars), min(tf) if tf else None, len(tf)), flush=True)

cases = [(512, 3, 64, int(os.environ.get('DF_RACE_T', '400'))), (128, 2, 37, 400), (256, 3, 64, 300)]
modes = [{}] + [dict(kv.split('=') for kv in arg.split(',')) for arg in sys.argv[1:]]
for w, d, n, T in cases:
    for opts in modes:
        run(w, d, n, T, dict((k, int(v)) for k, v in opts.items()))
