import os, sys, logging
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ocrd_keraslm_b200 import synth
from ocrd_keraslm_b200.lib import Rater
from oracle.rater_oracle import OracleRater
from tests import util
log = logging.getLogger('x'); log.setLevel(logging.CRITICAL)
W, D = int(sys.argv[1]), int(sys.argv[2])
wd = synth.weights(11, width=W, depth=D, voc_size=256, contexts=((200, 10),), style='trained')
maparr = synth.vocabulary()
r = Rater.from_weights(wd, maparr, 64, logger=log)
info = r._native.info()
print('tc_enabled', info['tc_enabled'], 'grid', info['tc_grid'])
rng = np.random.default_rng(3)
n = int(sys.argv[3]) if len(sys.argv) > 3 else 100
texts = [synth.text(500 + i, int(k)) for i, k in enumerate(rng.integers(5, 60, n))]
ctxs = [[int(c)] for c in rng.integers(0, 200, n)]
r._native.set_option('tc_wide', 1)
out_w = r.rate_many(texts, ctxs)
l1 = r._native.info()['tc_launches']
r._native.set_option('tc_wide', 0)
out_n = r.rate_many(texts, ctxs)
l2 = r._native.info()['tc_launches']
print('launches wide', l1 - info['tc_launches'], 'narrow', l2 - l1)
d = max(np.abs(np.log(a) - np.log(b)).max() for a, b in zip(out_w, out_n))
print('max |dlogp| wide vs narrow', d)
o = OracleRater(util.to_model(wd), synth.mapping_from_array(maparr), 64, logger=log)
worst = 0
for i in (0, 1, n // 2, n - 1):
    o.reset_states()
    want = o.rate(texts[i], ctxs[i])
    worst = max(worst, np.abs(np.log(out_w[i]) - np.log(want)).max())
print('max |dlogp| wide vs oracle', worst)
