"""CPU-only tests: the C-ABI library loads and exports every declared symbol, host logic of the
drop-in Rater that needs no device (model file round trip, vocabulary mapping, CuDNN weight
conversion, Node semantics), and the multi-rank sharding logic over gloo (world_size 2)."""
import ctypes
import os
import re
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from ocrd_keraslm_b200 import _native, h5lite, shard, synth
from ocrd_keraslm_b200.lib import rating

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ------------------------------------------------------------------ C ABI
def declared_symbols():
    with open(os.path.join(ROOT, 'include', 'keraslm_b200.h')) as f:
        src = f.read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(klm_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    lib = _native.load()                      # builds in-tree if needed; dlopen needs no GPU
    names = declared_symbols()
    assert 'klm_rate' in names and 'klm_step' in names and 'klm_create' in names
    for name in names:
        assert hasattr(lib, name), name
    bound = set(n for n, _, _ in _native.SYMBOLS)
    assert set(names) <= bound, set(names) - bound


def test_ctypes_structs_have_the_size_the_library_was_compiled_with():
    lib = _native.load()
    assert lib.klm_abi_sizeof(0) == ctypes.sizeof(_native.KlmConfig)
    assert lib.klm_abi_sizeof(1) == ctypes.sizeof(_native.KlmInfo)
    assert lib.klm_abi_sizeof(7) == -1


def test_no_cpu_fallback_without_device():
    """The product path fails loudly when there is no CUDA device (no silent CPU route)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    wd = synth.weights(1, width=16, depth=1, voc_size=256)
    r = rating.Rater()
    r.width, r.depth, r.length = 16, 1, 8
    r.set_mapping(synth.vocabulary())
    r.configure()
    with pytest.raises(Exception):
        r.set_weights(wd)


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, 'ocrd_keraslm_b200')):
        for fn in files:
            if fn.endswith('.py'):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), fn


# ------------------------------------------------------------------ model file format
def test_model_file_round_trip(tmp_path):
    wd = synth.weights(3, width=32, depth=2, voc_size=256, contexts=((200, 10),))
    maparr = synth.vocabulary()
    mapping = synth.mapping_from_array(maparr)
    fn = str(tmp_path / 'model.h5')
    rating.write_model_file(fn, wd, 32, 2, 64, True, False, mapping, 256, {'loss': [1.0, 0.5]})
    r = rating.Rater()
    r.load_config(fn)
    assert (r.width, r.depth, r.length, r.stateful, r.variable_length) == (32, 2, 64, True, False)
    assert r.voc_size == 256 and r.mapping[0]['a'] == mapping[0]['a'] and r.history == {'loss': [1.0, 0.5]}
    assert r.context_shapes == [(200, 10)]
    with h5lite.File(fn, 'r') as f:
        back = rating.read_keras_weights(f, 2)
    for k, v in wd.items():
        if isinstance(v, tuple):
            for a, b in zip(v, back[k]):
                np.testing.assert_array_equal(a, b)
        else:
            np.testing.assert_array_equal(v, back[k])


def test_cudnn_weight_conversion():
    """bias of 8W marks CuDNNLSTM-saved weights: per-gate transposed kernels, two biases added."""
    rng = np.random.default_rng(0)
    W, Din = 4, 6
    K = rng.standard_normal((Din, 4 * W)).astype(np.float32)
    R = rng.standard_normal((W, 4 * W)).astype(np.float32)
    b = rng.standard_normal(8 * W).astype(np.float32)
    Kc, Rc, bc = rating.convert_cudnn_weights(K, R, b)
    for g in range(4):
        blk = K[:, g * W:(g + 1) * W]
        np.testing.assert_array_equal(Kc[:, g * W:(g + 1) * W], blk.T.reshape(blk.shape, order='F'))
        np.testing.assert_array_equal(Rc[:, g * W:(g + 1) * W], R[:, g * W:(g + 1) * W].T)
    np.testing.assert_allclose(bc, b[:4 * W] + b[4 * W:])
    K2, R2, b2 = rating.convert_cudnn_weights(K, R, b[:4 * W])
    assert K2 is K and R2 is R


def test_vocabulary_lookup_table():
    r = rating.Rater()
    r.set_mapping(synth.vocabulary())
    idx, known = r._indices('aZ\n' + synth.UNMAPPED_CHAR + '\U0001F600')
    assert idx[0] == r.mapping[0]['a'] and idx[2] == r.mapping[0]['\n']
    assert list(known) == [True, True, True, False, False] and idx[3] == 0 and idx[4] == 0


def test_node_ordering_is_by_projected_cost():
    a = rating.Node(None, 'x', 1.0)
    b = rating.Node(None, 'y', 1.0)
    assert a == b and a in [b]
    with pytest.raises(TypeError):
        hash(a)
    alt = synth.Alt('abcd', 0.9, 0)
    c = rating.Node(None, 'ab', 1.0, parent=a, extras=(None, alt))
    assert c.cum_cost == 2.0 and c.pro_cost() == 3.0 and c.length == 2


# ------------------------------------------------------------------ sharding
def test_partition_is_balanced_and_deterministic():
    rng = np.random.default_rng(5)
    lengths = rng.integers(20, 121, 1000)
    parts = shard.partition(lengths, 8)
    assert sorted(i for p in parts for i in p) == list(range(1000))
    loads = [int(sum(lengths[i] for i in p)) for p in parts]
    assert max(loads) - min(loads) <= 121
    assert parts == shard.partition(lengths, 8)
    assert shard.partition([], 2) == [[], []]
    assert shard.partition([5], 3) == [[0], [], []]


WORKER = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, %(root)r)
    import numpy as np
    import torch.distributed as dist
    from ocrd_keraslm_b200 import shard

    class FakeRater(object):
        """Stand-in with rate_many semantics: score of position t of a text depends on the text only."""
        calls = 0
        fail_on_rank = None
        def rate_many(self, texts, contexts=None):
            import unicodedata
            if FakeRater.fail_on_rank == dist.get_rank():
                raise RuntimeError('device lost')
            FakeRater.calls += len(texts)
            out = []
            for t, c in zip(texts, contexts):
                t = unicodedata.normalize('NFC', t)          # like Rater.rate_many: one score per NFC character
                v = np.array([((ord(ch) * 31 + k + (c[0] if c else 0)) %% 97) / 97.0 for k, ch in enumerate(t)], dtype=np.float32)
                if len(v):
                    v[0] = 1.0
                out.append(v)
            return out

    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%(port)d', rank=int(sys.argv[1]), world_size=2)
    rng = np.random.default_rng(11)
    texts = [''.join(chr(97 + int(x)) for x in rng.integers(0, 26, int(n))) for n in rng.integers(0, 40, 23)]
    texts[3] = 'a\u0308bc' + texts[3]            # decomposed sequences: NFC makes these texts shorter
    texts[17] = 'e\u0301' * 9
    ctxs = [[int(rng.integers(0, 200))] for _ in texts]
    full = FakeRater().rate_many(texts, ctxs)
    assert len(full[17]) == 9
    FakeRater.calls = 0
    r = FakeRater()
    got = shard.rate_sharded(r, texts, ctxs, gather=True)
    import unicodedata
    mine = shard.partition([len(unicodedata.normalize('NFC', t)) for t in texts], 2)[dist.get_rank()]
    assert FakeRater.calls == len(mine), (FakeRater.calls, len(mine))
    assert len(got) == len(full)
    for a, b in zip(got, full):
        assert a is not None and a.shape == b.shape and np.array_equal(a, b)
    part = shard.rate_sharded(r, texts, ctxs, gather=False)
    for i, a in enumerate(part):
        assert (a is not None) == (i in mine)
    # a rank whose rating fails must not leave the other one hanging in the gather: both raise
    FakeRater.fail_on_rank = 1
    try:
        shard.rate_sharded(r, texts, ctxs, gather=True)
        raise SystemExit('no exception on rank %%d' %% dist.get_rank())
    except RuntimeError as e:
        assert ('device lost' in str(e)) == (dist.get_rank() == 1), str(e)
    FakeRater.fail_on_rank = None
    dist.barrier()
    dist.destroy_process_group()
    print('rank', sys.argv[1], 'ok')
''')


def test_rate_sharded_world_size_2_gloo(tmp_path):
    port = 29500 + (os.getpid() % 2000)
    script = tmp_path / 'worker.py'
    script.write_text(WORKER % {'root': ROOT, 'port': port})
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
             for r in range(2)]
    outs = [p.communicate(timeout=240)[0].decode() for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, o
        assert 'rank %d ok' % r in o


# ------------------------------------------------------------------ h5lite against a file written by libhdf5
def test_h5lite_reads_a_libhdf5_written_file():
    """The only third-party HDF5 file on the box: scipy's MATLAB 7.4 test file -- written by libhdf5 1.6, superblock v0
    behind a 512-byte user block (base address 512), symbol-table group, v1 object headers, layout message v2.  Not a
    Keras file (none is available: h5lite's writer -> reader round trip stays the only check of those objects), but it
    pins the reader's superblock / group / object-header / dataspace / datatype code against the real library."""
    scipy = pytest.importorskip('scipy')
    path = os.path.join(os.path.dirname(scipy.__file__), 'io', 'matlab', 'tests', 'data', 'testhdf5_7.4_GLNX86.mat')
    if not os.path.exists(path):
        pytest.skip('scipy test data not installed')
    from ocrd_keraslm_b200 import h5lite
    with h5lite.File(path, 'r') as f:
        assert list(f.keys()) == ['testdouble']
        d = f['testdouble']
        assert d.shape == (9, 1) and d.dtype == np.float64
        assert d.attrs['MATLAB_class'] == b'double'
        np.testing.assert_allclose(d[()].ravel(), np.arange(9) * np.pi / 4, rtol=1e-15)
    with open(path, 'rb') as fh:
        data = fh.read()
    with pytest.raises(h5lite.H5Error):                         # user block cut off: base address no longer matches
        h5lite._Reader(data[512:])
    with pytest.raises(h5lite.H5Error):
        h5lite._Reader(b'MATLAB 7.0 MAT-file' + b'\0' * 2000)


# ------------------------------------------------------------------ the golden fixtures are what the reference produces
@pytest.mark.skipif(not os.path.exists('/root/reference/ocrd_keraslm/lib/rating.py'), reason='needs /root/reference (build container only)')
def test_golden_fixtures_are_reproduced_by_the_reference(tmp_path):
    """Re-runs tests/golden/make_golden.py (the reference's own rating.py on the Keras shim) in a scratch copy and checks
    that the committed fixtures come out again, bit for bit."""
    import json
    import shutil
    work = tmp_path / 'repo'
    (work / 'tests').mkdir(parents=True)
    shutil.copytree(os.path.join(ROOT, 'tests', 'golden'), work / 'tests' / 'golden')
    for pkg in ('oracle', 'ocrd_keraslm_b200'):
        os.symlink(os.path.join(ROOT, pkg), work / pkg)
    for name in ('golden.json', 'golden.npz'):
        os.remove(work / 'tests' / 'golden' / name)
    res = subprocess.run([sys.executable, str(work / 'tests' / 'golden' / 'make_golden.py')], stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, timeout=600)
    assert res.returncode == 0, res.stdout.decode()
    with open(work / 'tests' / 'golden' / 'golden.json', encoding='utf-8') as f:
        new = json.load(f)
    with open(os.path.join(ROOT, 'tests', 'golden', 'golden.json'), encoding='utf-8') as f:
        old = json.load(f)
    assert new == old
    a, b = np.load(work / 'tests' / 'golden' / 'golden.npz'), np.load(os.path.join(ROOT, 'tests', 'golden', 'golden.npz'))
    assert sorted(a.files) == sorted(b.files)
    for k in a.files:
        np.testing.assert_array_equal(a[k], b[k], err_msg=k)


def test_native_partition_follows_the_python_rule():
    """shard.partition hands large inputs to klm_lpt_partition (host helper of the library): the same lists as the
    Python statement of the rule, incl. ties, empty units and more ranks than units."""
    rng = np.random.default_rng(11)
    for n, world in ((shard.NATIVE_PARTITION_MIN, 2), (3000, 8), (10000, 3), (777, 1), (600, 16)):
        lengths = rng.integers(0, 121, n)
        assert shard.partition(lengths, world) == shard._partition_py(np.asarray(lengths, dtype=np.int64), world)
    few = np.concatenate([np.full(600, 7), np.zeros(5, dtype=np.int64)])
    assert shard.partition(few, 700) == shard._partition_py(few.astype(np.int64), 700)
