#!/usr/bin/env python
"""fp64 oracle reference of LONG runs at BASELINE config 2's model shape (width 512, depth 3), for the drift tests.

The fp64 oracle (oracle/lstm_torch_cpu.py, checked against oracle/lstm_numpy.py in tests/test_oracle_golden.py)
needs minutes for 16 streams x 100 000 positions, so it runs HERE (build container) and the GPU tests compare
against the committed windows:

    tests/golden/long_w512_trained.npz   16 streams x 20 000 positions, "trained-like" weights (synth.weights(543, style='trained'))
    tests/golden/long_w512_keras.npz     16 streams x 99 999 positions, the weights of bench.py's cfg2 (synth.weights(9, style='keras'))

Each file: `windows` int64 [k,2] (position ranges), `logp_<a>_<b>` float32 [16, b-a] = log p of the target characters in
fp64, rounded once to float32; `states_<T>` float32 [16, 2*depth, 512] = (h1,c1,h2,c2,h3,c3) after consuming T positions.
Inputs are regenerated from seeds by the tests (`long_inputs` below is imported by them), so nothing else is stored.

This is an ORACLE-generated fixture (the arithmetic of the reference lives in Keras/TF and cannot run here);
the reference-generated fixtures are golden.json / golden.npz (make_golden.py).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from ocrd_keraslm_b200 import synth  # noqa: E402

N_STREAMS = 16
CONFIGS = {
    #           weights seed, style, positions, state checkpoints, windows
    'trained': (543, 'trained', 20000, (20000,), ((0, 1000), (9000, 10000), (19000, 20000))),
    'keras': (9, 'keras', 99999, (20000, 99999), ((0, 1000), (9000, 10000), (19000, 20000), (49000, 50000), (98999, 99999))),
}


def long_inputs(name):
    """(weights dict, mapping array, texts, contexts) of a configuration; texts have positions + 1 characters."""
    seed, style, positions, _, _ = CONFIGS[name]
    wd = synth.weights(seed, width=512, depth=3, voc_size=256, contexts=((200, 10),), style=style)
    texts = [synth.text(77000 + 100 * seed + i, positions + 1) for i in range(N_STREAMS)]
    ctxs = [[(37 * i + 5) % 200] for i in range(N_STREAMS)]
    ctxs[3] = [0]                                            # one underspecified stream
    return wd, synth.vocabulary(), texts, ctxs


def main():
    from tests import util
    from oracle.lstm_torch_cpu import TorchCpuModel
    for name in sys.argv[1:] or sorted(CONFIGS):
        seed, style, positions, checkpoints, windows = CONFIGS[name]
        wd, maparr, texts, ctxs = long_inputs(name)
        c_i = synth.mapping_from_array(maparr)[0]
        idx = np.stack([np.fromiter((c_i.get(ch, 0) for ch in t), dtype=np.int64, count=len(t)) for t in texts])
        zs = [np.array([c[0] for c in ctxs], dtype=np.int64)]
        model = TorchCpuModel(util.to_model(wd), dtype=np.float64)
        out = {'windows': np.array(windows, dtype=np.int64)}
        logp = np.empty((N_STREAMS, positions), dtype=np.float64)
        states, done = None, 0
        t0 = time.time()
        for cp in checkpoints:
            lp, states = model.rate(idx[:, done:cp], zs, idx[:, done + 1:cp + 1], states)
            logp[:, done:cp] = lp
            out['states_%d' % cp] = np.stack([s.numpy() for s in states], axis=1).astype(np.float32)
            done = cp
            print('%s: %d positions, %.0f s' % (name, done, time.time() - t0), flush=True)
        assert done == positions
        for a, b in windows:
            out['logp_%d_%d' % (a, b)] = logp[:, a:b].astype(np.float32)
        out['logp_sum_fp64'] = np.array(logp.sum())
        np.savez_compressed(os.path.join(HERE, 'long_w512_%s.npz' % name), **out)


if __name__ == '__main__':
    main()
