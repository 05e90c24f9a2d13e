"""Per-CTA timeline of the grid-barrier sequence kernel: which blocks are the stragglers of a tick?"""
import os, sys, logging
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
from ocrd_keraslm_b200 import synth
from ocrd_keraslm_b200.lib import Rater
log = logging.getLogger('t'); log.setLevel(logging.CRITICAL)
T = 3000
r = Rater.from_weights(bench.model_weights('cfg2'), synth.vocabulary(), 512, logger=log)
texts = [synth.text(100 + i, T) for i in range(64)]
ctxs = [[179]] * 64
r.rate_many(texts, ctxs)
r._native.tc_trace_arm(1000)
r.rate_many(texts, ctxs)
tr, roles = r._native.tc_trace_read()
tr = tr.astype(np.int64)
names = {0: 'REC', 1: 'IN', 2: 'PROJ', 3: 'LSE'}
x = tr[:, 2:14, :]
rel = (x - x[:, :, :1]).mean(axis=1)            # [block][event] mean offset from the block's own tick start
period = (x[:, 1:, 0] - x[:, :-1, 0]).mean()
print('period %.0f clk' % period)
ev = ['start', 'issue0', 'chunk0', 'chunkN', 'mma_done', 'drained', 'stored', 'arrive']
for b in range(len(roles)):
    print('block %3d %-4s l%d pairs %2d: ' % (b, names[int(roles[b, 0])], roles[b, 1], roles[b, 2]) + ' '.join('%s %5.0f' % (e, v) for e, v in zip(ev, rel[b])))
