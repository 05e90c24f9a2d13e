"""Timeline of the tcgen05 sequence kernel: per role, mean SM-clock offsets of the per-tick events.
usage: tc_trace.py [cfg2|W,D] [T]"""
import os, sys, logging
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from ocrd_keraslm_b200 import synth, _native
from ocrd_keraslm_b200.lib import Rater

wl = sys.argv[1] if len(sys.argv) > 1 else 'cfg2'
T = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
log = logging.getLogger('t'); log.setLevel(logging.CRITICAL)
if wl in bench.WORKLOADS:
    width, depth, length, contexts, docs, chars, what = bench.WORKLOADS[wl]
    wd = bench.model_weights(wl)
else:
    width, depth = [int(x) for x in wl.split(',')]
    length = 512
    wd = synth.weights(3, width=width, depth=depth, style='keras')
r = Rater.from_weights(wd, synth.vocabulary(), length, logger=log)
NS = int(os.environ.get('TC_TRACE_STREAMS', '64'))          # 128: the wide streamed kernel (klm_seq_tw)
texts = [synth.text(100 + i, T) for i in range(NS)]
ctxs = [[179]] * NS if len(r.context_shapes) == 1 else [[1] * len(r.context_shapes)] * NS
r.rate_many(texts, ctxs)
r._native.tc_trace_arm(1000)
r.rate_many(texts, ctxs)
tr, roles = r._native.tc_trace_read()
info = r._native.info()
print('grid', info['tc_grid'], 'last kernel ms', info['last_kernel_ms'], 'per tick us', 1e3 * info['last_kernel_ms'] / (T + 2 * depth))
names = {0: 'REC', 1: 'IN', 2: 'PROJ', 3: 'LSE'}
tr = tr.astype(np.int64)
ev = ['start', 'issue0', 'chunk0', 'chunkN', 'mma_done', 'drained', 'stored', 'arrive']
if os.environ.get('KLM_TC_TRACE_MODE') == '4':      # epilogue phases of one epilogue thread, relative to the accumulators being complete
    ev = ['mma_done', 'r0_drain', 'r0_xchg', 'r0_fin', 'r1_drain', 'r1_xchg', 'r1_fin', 'arrive']
for kind in (0, 1, 2, 3):
    for layer in range(depth):
        sel = [i for i in range(len(roles)) if roles[i, 0] == kind and roles[i, 1] == layer]
        if not sel:
            continue
        x = tr[sel][:, 2:14, :]                   # [cta][tick][ev]
        period = (x[:, 1:, 0] - x[:, :-1, 0]).mean()
        rel = (x - x[:, :, :1]).mean(axis=(0, 1))
        if kind == 3:
            print('LSE n=%d period %.0f clk | start +0 arrive %+.0f (slowest CTA %+.0f)' % (
                len(sel), period, rel[7], (x[:, :, 7] - x[:, :, 0]).mean(axis=1).max()))
            break
        print('%s l%d n=%d pairs=%s period %.0f clk | ' % (names[kind], layer, len(sel), sorted(set(int(v) for v in roles[sel, 2])), period) +
              ' '.join('%s %+.0f' % (e, v) for e, v in zip(ev, rel)) +
              ' | slowest CTA arrive %+.0f' % (x[:, :, 7] - x[:, :, 0]).mean(axis=1).max())
