import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ocrd_keraslm_b200 import _native
rng = np.random.default_rng(1)
for N, K in [(32, 16), (64, 64), (160, 128), (256, 64)]:
    A = rng.integers(-4, 5, (128, K)).astype(np.float16)
    B = rng.integers(-4, 5, (N, K)).astype(np.float16)
    ref = A.astype(np.float32) @ B.astype(np.float32).T
    try:
        raw = _native.selftest_umma2_raw(A, B)
    except Exception as e:
        print(N, K, 'ERR', e); continue
    D = np.concatenate([np.concatenate([raw[r, :64], raw[r, 64:]], axis=1) for r in range(2)], axis=0)
    ok = np.array_equal(D, ref)
    print('N=%d K=%d layout-2x2 match: %s' % (N, K, ok))
    if not ok:
        # find where each reference row shows up
        for r in range(2):
            for lane in (0, 1, 16, 32, 63, 64, 65, 96, 127):
                row = raw[r, lane]
                hits = [(m, n0) for m in range(128) for n0 in range(0, N, N // 2) if np.array_equal(ref[m, n0:n0 + N // 2], row)]
                print('  cta', r, 'lane', lane, '->', hits[:3], 'nonzero' if row.any() else 'zero')
