"""Where a beam step of the dataflow kernel spends its time: global-timer trace (ns) of one klm_step launch."""
import os, sys, logging
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
os.environ['KLM_TC_TRACE_MODE'] = '5'
import bench
from ocrd_keraslm_b200 import synth
from ocrd_keraslm_b200.lib import Rater
log = logging.getLogger('x'); log.setLevel(logging.CRITICAL)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 117
wd = bench.model_weights('cfg2')
r = Rater.from_weights(wd, synth.vocabulary(), 512, stateful=False, incremental=True, logger=log)
nat = r._native
rng = np.random.default_rng(1)
prev = r._pool.acquire(n); nat.state_reset(prev)
def step():
    chars = rng.integers(1, 256, n).astype(np.int32); targets = rng.integers(1, 256, n).astype(np.int32)
    ctx = np.full((n, 1), 179, dtype=np.int32); out = np.empty(n, dtype=np.float32)
    dst = r._pool.acquire(n)
    nat.step(chars, ctx, prev, dst, targets, out)
    r._pool.release_many(dst)
    return nat.info()['last_kernel_ms']
for _ in range(5): step()
print('kernel ms (untraced): %s' % ['%.3f' % step() for _ in range(5)])
for _ in range(6):                 # (the first launches of the traced instantiation are cold: take the last)
    nat.tc_trace_arm(0)
    ms = step()
tr, roles = nat.tc_trace_read()
tr = tr.astype(np.int64)
t0 = tr[tr > 0].min()
names = {0: 'REC', 1: 'IN', 2: 'PROJ', 3: 'LSE'}
ev = ['flags', 'issue0', 'chunk0', 'chunkN', 'mma_done', 'drained', 'staged', 'published']
print('traced kernel ms %.3f; events in ns after the first event of the launch (min..max over CTAs)' % ms)
for tick in range((n + 63) // 64):
    for kind in (0, 1, 2):
        for layer in range(3):
            sel = [i for i in range(len(roles)) if roles[i, 0] == kind and roles[i, 1] == layer]
            if not sel: continue
            x = tr[sel][:, tick, :]
            print('tick %d %s l%d: ' % (tick, names[kind], layer) + ' '.join(
                '%s %d..%d' % (e, (x[:, k][x[:, k] > 0] - t0).min() if (x[:, k] > 0).any() else -1,
                               (x[:, k][x[:, k] > 0] - t0).max() if (x[:, k] > 0).any() else -1) for k, e in enumerate(ev)))

x = tr[:, 15, :5]
names5 = ['kernel entry', 'setup done (barriers, TMEM, image copies issued)', 'REC prologue done (initial h published)', 'ticks done', 'exit']
for k, nm in enumerate(names5):
    v = x[:, k][x[:, k] > 0]
    if len(v): print('%s: %d..%d ns after the first tick event' % (nm, v.min() - t0, v.max() - t0))
