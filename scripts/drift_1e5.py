#!/usr/bin/env python
"""Drift of the tensor-core sequence kernel over 16 streams x 99 999 positions at width 512 / depth 3 with the
weights of bench.py's cfg2 (Keras-default initialisation), against the fp64 oracle windows committed in
tests/golden/long_w512_keras.npz (tests/golden/make_long_reference.py).  One launch, 99 999 ticks.

Run on the GPU box (scripts/gpu_run.sh <tag> drift); prints one JSON object: max |dlogp| per window, relative error of
the final states (and at T = 20 000 from a second, shorter run), `ok` against BASELINE.json's tolerances.
Too long for the -m gpu suite (the 20 000-position test in tests/test_gpu_long.py is the gate); the result is kept
under profiles/.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests import util  # noqa: E402
from tests.golden.make_long_reference import long_inputs  # noqa: E402
from tests.test_gpu_long import LOGP_TOL, STATE_TOL, product, rate_streams  # noqa: E402


def main():
    name = 'keras'
    ref = np.load(os.path.join(util.GOLDEN_DIR, 'long_w512_%s.npz' % name))
    wd, maparr, texts, ctxs = long_inputs(name)
    r = product(wd, maparr, 512)
    info = r._native.info()
    out = {'model': 'width 512 depth 3, %s weights' % name, 'streams': len(texts), 'tc_enabled': info['tc_enabled'],
           'max_abs_dlogp': {}, 'state_rel_err': {}, 'tolerance': {'logp': LOGP_TOL, 'state': STATE_TOL}}
    ok = True
    for T in (20000, 99999):
        if 'states_%d' % T not in ref.files:
            continue
        t0 = time.time()
        logp, states = rate_streams(r, [t[:T + 1] for t in texts], ctxs)
        out['seconds_T=%d' % T] = round(time.time() - t0, 3)
        want = ref['states_%d' % T].astype(np.float64)
        rel = float(np.abs(states - want).max() / np.abs(want).max())
        out['state_rel_err']['T=%d' % T] = rel
        ok = ok and rel <= STATE_TOL
        if T == 99999:
            for a, b in ref['windows']:
                w = ref['logp_%d_%d' % (a, b)].astype(np.float64)
                g = np.stack([lp[a:b] for lp in logp]).astype(np.float64)
                err = float(np.abs(g - w).max()) if not np.isnan(g).any() else float('nan')
                out['max_abs_dlogp']['T=%d..%d' % (a, b)] = err
                ok = ok and err <= LOGP_TOL
            out['logp_sum'] = float(np.concatenate(logp).astype(np.float64).sum())
            out['logp_sum_fp64_oracle'] = float(ref['logp_sum_fp64'])
    out['tc_launches'] = r._native.info()['tc_launches'] - info['tc_launches']
    out['ok'] = bool(ok)
    print(json.dumps(out, indent=1))
    return 0 if ok else 1


if __name__ == '__main__':
    sys.exit(main())
