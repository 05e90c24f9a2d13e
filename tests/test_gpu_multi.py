"""Multi-GPU path: units that rate independently are sharded over the ranks, one process per GPU, no data-path
collective; the gathered result must equal what one GPU computes, bit for bit.  Skipped with fewer than 2 GPUs."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent('''
    import os, sys, logging
    sys.path.insert(0, %(root)r)
    import numpy as np
    import torch
    import torch.distributed as dist
    from ocrd_keraslm_b200 import shard, synth
    from ocrd_keraslm_b200.lib import Rater
    rank = int(sys.argv[1])
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', init_method='tcp://127.0.0.1:%(port)d', rank=rank, world_size=2,
                            device_id=torch.device('cuda', rank))
    log = logging.getLogger('w'); log.setLevel(logging.CRITICAL)
    wd = synth.weights(12, width=256, depth=2, voc_size=256, style='trained')
    r = Rater.from_weights(wd, synth.vocabulary(), 64, logger=log, device=rank)
    rng = np.random.default_rng(3)
    texts = [synth.text(500 + i, int(n)) for i, n in enumerate(rng.integers(0, 90, 300))]
    texts[7] = 'a\\u0308bc' + texts[7]                     # decomposed: NFC makes it shorter
    ctxs = [[int(c)] for c in rng.integers(0, 200, len(texts))]
    got = shard.rate_sharded(r, texts, ctxs, gather=True)
    part = shard.rate_sharded(r, texts, ctxs, gather=False)
    whole = r.rate_many(texts, ctxs)                       # every rank also computes everything on its own GPU
    mine = shard.partition([len(__import__('unicodedata').normalize('NFC', t)) for t in texts], 2)[rank]
    assert len(got) == len(whole)
    for i, (a, b) in enumerate(zip(got, whole)):
        assert a.shape == b.shape and np.array_equal(a, b), i
        assert (part[i] is not None) == (i in mine)
    dist.barrier()
    dist.destroy_process_group()
    print('rank', rank, 'ok')
''')


def test_rate_sharded_two_gpus_equals_one_gpu(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    port = 29600 + (os.getpid() % 2000)
    script = tmp_path / 'worker.py'
    script.write_text(WORKER % {'root': ROOT, 'port': port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
             for r in range(2)]
    outs = [p.communicate(timeout=600)[0].decode() for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert 'rank %d ok' % r in o
