"""ctypes binding of libkeraslm_b200.so (include/keraslm_b200.h) and its in-tree build.

There is no CPU fallback: if the library cannot be loaded or no CUDA device is present the
product raises, it never routes through another implementation.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB_PATH = os.path.join(HERE, 'libkeraslm_b200.so')
SOURCES = ['klm_api.cu', 'klm_seq_fma.cu', 'klm_seq_tc.cu', 'klm_seq_df.cu', 'klm_seq_rg.cu', 'klm_seq_tw.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-shared', '-Xcompiler', '-fPIC']

KLM_MAX_CTX = 4
MEM_HOST, MEM_DEVICE = 0, 1
NOALLOC_SCRATCH_BYTES = 65536      # KLM_NOALLOC_SCRATCH_BYTES (include/keraslm_b200.h)


class NativeError(RuntimeError):
    pass


def _sources():
    out = [os.path.join(CSRC, s) for s in SOURCES]
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.h'))]
    hdrs.append(os.path.join(os.path.dirname(HERE), 'include', 'keraslm_b200.h'))
    return out, hdrs


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    srcs, hdrs = _sources()
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(f) > t for f in srcs + hdrs)


def build(force=False, verbose=False):
    """Compile the CUDA sources for sm_100a into ocrd_keraslm_b200/libkeraslm_b200.so (in-tree).

    Every source is compiled to an object on its own (in parallel; only the ones whose source or a header is newer
    than the object), then linked."""
    if not force and not needs_build():
        return LIB_PATH
    # one builder at a time: the ranks of a torchrun job that all find the library stale would otherwise compile into
    # the same object files and link over each other (seen on an 8-GPU box: "file too short" / undefined symbols);
    # the others wait for the lock, find the library fresh and return
    import fcntl
    os.makedirs(os.path.join(HERE, 'build'), exist_ok=True)
    with open(os.path.join(HERE, 'build', '.lock'), 'w') as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not needs_build():
                return LIB_PATH
            return _build_locked(force, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(force, verbose):
    from concurrent.futures import ThreadPoolExecutor
    nvcc = os.environ.get('NVCC') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(nvcc):
        nvcc = 'nvcc'
    srcs, hdrs = _sources()
    objdir = os.path.join(HERE, 'build')
    os.makedirs(objdir, exist_ok=True)
    newest_hdr = max(os.path.getmtime(h) for h in hdrs)

    def compile_one(src):
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + '.o')
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(src), newest_hdr):
            return obj, 0, ''
        cmd = [nvcc] + [f for f in NVCC_FLAGS if f != '-shared'] + ['-c', '-o', obj, src]
        if verbose:
            print(' '.join(cmd), file=sys.stderr)
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, universal_newlines=True)
        return obj, res.returncode, res.stdout

    with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
        results = list(ex.map(compile_one, srcs))
    for obj, rc, out in results:
        if rc != 0:
            raise NativeError('nvcc failed:\n' + out)
    tmp = LIB_PATH + '.tmp%d' % os.getpid()          # linked beside the library and renamed: a reader never sees half a file
    cmd = [nvcc, '-gencode', 'arch=compute_100a,code=sm_100a', '-shared', '-o', tmp] + [r[0] for r in results]
    if verbose:
        print(' '.join(cmd), file=sys.stderr)
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, universal_newlines=True)
    if res.returncode != 0:
        raise NativeError('nvcc (link) failed:\n' + res.stdout)
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


class KlmConfig(ctypes.Structure):
    _fields_ = [('width', ctypes.c_int32), ('depth', ctypes.c_int32), ('voc_size', ctypes.c_int32),
                ('n_ctx', ctypes.c_int32), ('ctx_rows', ctypes.c_int32 * KLM_MAX_CTX),
                ('ctx_dim', ctypes.c_int32 * KLM_MAX_CTX)]


class KlmInfo(ctypes.Structure):
    _fields_ = [('sm_count', ctypes.c_int32), ('grid', ctypes.c_int32), ('weights_in_smem', ctypes.c_int32),
                ('smem_bytes', ctypes.c_int32), ('kernel_launches', ctypes.c_int64),
                ('last_kernel_ms', ctypes.c_double), ('max_streams', ctypes.c_int32), ('tc_enabled', ctypes.c_int32),
                ('tc_grid', ctypes.c_int32), ('tc_smem_bytes', ctypes.c_int32), ('tc_launches', ctypes.c_int64),
                ('cl_launches', ctypes.c_int64), ('cl_cluster', ctypes.c_int32), ('tc_dataflow', ctypes.c_int32),
                ('tc_wide', ctypes.c_int32), ('tc_wide_launches', ctypes.c_int64)]


# every symbol include/keraslm_b200.h declares: (name, restype, argtypes)
_P = ctypes.c_void_p
_I = ctypes.c_int32
SYMBOLS = [
    ('klm_create', ctypes.c_int, [ctypes.POINTER(KlmConfig), ctypes.POINTER(_P), ctypes.c_int, ctypes.POINTER(_P)]),
    ('klm_destroy', ctypes.c_int, [_P]),
    ('klm_last_error', ctypes.c_char_p, []),
    ('klm_abi_sizeof', ctypes.c_int, [ctypes.c_int]),
    ('klm_set_stream', ctypes.c_int, [_P, _P]),
    ('klm_sync', ctypes.c_int, [_P]),
    ('klm_set_option', ctypes.c_int, [_P, ctypes.c_char_p, ctypes.c_int64]),
    ('klm_state_pool_resize', ctypes.c_int, [_P, _I]),
    ('klm_state_reset', ctypes.c_int, [_P, _P, _I]),
    ('klm_state_copy', ctypes.c_int, [_P, _P, _P, _I]),
    ('klm_state_to_host', ctypes.c_int, [_P, _P, _I, _P]),
    ('klm_state_from_host', ctypes.c_int, [_P, _P, _I, _P]),
    ('klm_state_l2_distance', ctypes.c_int, [_P, _P, _P, _I, _I, _P]),
    ('klm_state_l2_distance_noalloc', ctypes.c_int, [_P, _P, _P, _I, _I, _P]),
    ('klm_rate', ctypes.c_int, [_P, _I, _P, _P, _P, _P, _P, _P, _P, ctypes.c_int]),
    ('klm_step', ctypes.c_int, [_P, _I, _P, _P, _P, _P, _P, _P, ctypes.c_int]),
    ('klm_step_topk', ctypes.c_int, [_P, _I, _P, _P, _P, _P, _I, _P, _P]),
    ('klm_get_info', ctypes.c_int, [_P, ctypes.POINTER(KlmInfo)]),
    ('klm_lpt_partition', ctypes.c_int, [_P, ctypes.c_int64, _I, _P]),
    ('klm_measure_fma_peak', ctypes.c_int, [ctypes.c_int, ctypes.POINTER(ctypes.c_double)]),
    ('klm_selftest_umma', ctypes.c_int, [ctypes.c_int, _I, _I, _P, _P, _P]),
    ('klm_selftest_umma2', ctypes.c_int, [ctypes.c_int, _I, _I, _P, _P, _P]),
    ('klm_tc_trace', ctypes.c_int, [_P, _I, _P, _P]),
    ('klm_beam_lattice', ctypes.c_int, [_P, _I, _P, _P, _P, _I, _P, _P, _P, _P, _P, ctypes.c_double, _I, _I, _I, _P, _I,
                                        _P, _P, _P, _P, _P, _P, _P]),
    ('klm_beam_edge', ctypes.c_int, [_P, _I, _P, _P, _P, _I, _P, _P, _P, _P, ctypes.c_double, _I, _I, _P, _I,
                                     _P, _P, _P, _P, _P, _P, _P]),
]

_lib = None


def load():
    """dlopen the library (building it first if the sources are newer) and bind all symbols."""
    global _lib
    if _lib is not None:
        return _lib
    if needs_build():
        build()
    lib = ctypes.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _ptr(a):
    """Host numpy array / int address / None -> void pointer value."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags['C_CONTIGUOUS']
        return a.ctypes.data
    return int(a)


def pinned_empty(n, dtype):
    """Page-locked host array (through torch, which is only the buffer carrier here)."""
    import torch
    tdt = {np.dtype(np.int32): torch.int32, np.dtype(np.float32): torch.float32}[np.dtype(dtype)]
    return torch.empty(int(n), dtype=tdt, pin_memory=True).numpy()


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def operand_image(mat):
    """fp16 [rows, K] -> the kernels' shared-memory operand layout [K/8][rows][8] (as uint16)."""
    rows, K = mat.shape
    assert K % 8 == 0
    return np.ascontiguousarray(mat.astype(np.float16).reshape(rows, K // 8, 8).transpose(1, 0, 2)).view(np.uint16)


def lpt_partition(lengths, world):
    """rank of every unit under longest-processing-time-first assignment (klm_lpt_partition; host only)."""
    lib = load()
    lengths = np.ascontiguousarray(lengths, dtype=np.int64)
    rank_of = np.empty(len(lengths), dtype=np.int32)
    rc = lib.klm_lpt_partition(_ptr(lengths), len(lengths), world, _ptr(rank_of))
    if rc != 0:
        raise NativeError('libkeraslm_b200 error %d: %s' % (rc, lib.klm_last_error().decode('utf-8', 'replace')))
    return rank_of


def measure_fma_peak(device=0):
    """Measured FP32 FMA throughput of the device in TFLOP/s (klm_measure_fma_peak)."""
    lib = load()
    v = ctypes.c_double(0)
    rc = lib.klm_measure_fma_peak(device, ctypes.byref(v))
    if rc != 0:
        raise NativeError('libkeraslm_b200 error %d: %s' % (rc, lib.klm_last_error().decode('utf-8', 'replace')))
    return v.value


def selftest_umma(A, B, device=0):
    """D = A . B^T on the tensor cores through the library's self-test kernel; A [128, K], B [N, K] (fp16 values)."""
    lib = load()
    A = np.asarray(A, dtype=np.float16)
    B = np.asarray(B, dtype=np.float16)
    assert A.shape[0] == 128 and A.shape[1] == B.shape[1]
    D = np.empty((128, B.shape[0]), dtype=np.float32)
    ai, bi = operand_image(A), operand_image(B)
    rc = lib.klm_selftest_umma(device, B.shape[0], A.shape[1], ai.ctypes.data, bi.ctypes.data, D.ctypes.data)
    if rc != 0:
        raise NativeError('klm_selftest_umma: %s' % lib.klm_last_error().decode())
    return D


def selftest_umma2_raw(A, B, device=0):
    """CTA-pair MMA (cta_group::2): A [128, K], B [N, K]; returns the raw TMEM dump [2][128 lanes][N/2 columns]."""
    lib = load()
    A = np.asarray(A, dtype=np.float16)
    B = np.asarray(B, dtype=np.float16)
    N, K = B.shape
    ai = np.concatenate([operand_image(A[:64]).ravel(), operand_image(A[64:]).ravel()])
    bi = np.concatenate([operand_image(B[:N // 2]).ravel(), operand_image(B[N // 2:]).ravel()])
    D = np.empty((2, 128, N // 2), dtype=np.float32)
    rc = lib.klm_selftest_umma2(device, N, K, ai.ctypes.data, bi.ctypes.data, D.ctypes.data)
    if rc != 0:
        raise NativeError('klm_selftest_umma2: %s' % lib.klm_last_error().decode())
    return D


def selftest_umma2(A, B, device=0):
    """D = A . B^T through the CTA-pair MMA, reassembled from the 2x2 accumulator layout:
    CTA r holds rows 64r..64r+63; lanes 0..63 carry columns [0, N/2), lanes 64..127 columns [N/2, N)."""
    raw = selftest_umma2_raw(A, B, device)
    return np.concatenate([np.concatenate([raw[r, :64], raw[r, 64:]], axis=1) for r in range(2)], axis=0)


class Model(object):
    """Owner of one klm_model handle."""

    def __init__(self, width, depth, voc_size, ctx_shapes, weights, device=0):
        lib = load()
        cfg = KlmConfig()
        cfg.width, cfg.depth, cfg.voc_size, cfg.n_ctx = width, depth, voc_size, len(ctx_shapes)
        if len(ctx_shapes) > KLM_MAX_CTX:
            raise ValueError('at most %d context variables are supported' % KLM_MAX_CTX)
        for k, (rows, dim) in enumerate(ctx_shapes):
            cfg.ctx_rows[k], cfg.ctx_dim[k] = rows, dim
        self._keep = [np.ascontiguousarray(w, dtype=np.float32) for w in weights]
        assert len(self._keep) == 1 + len(ctx_shapes) + 3 * depth
        arr = (_P * len(self._keep))(*[w.ctypes.data for w in self._keep])
        handle = _P()
        self._lib = lib
        self._h = None
        self.check(lib.klm_create(ctypes.byref(cfg), arr, device, ctypes.byref(handle)))
        self._h = handle
        self._keep = None
        self.width, self.depth, self.voc_size, self.n_ctx = width, depth, voc_size, len(ctx_shapes)
        self.device = device
        self.n_slots = 0

    def check(self, rc):
        if rc != 0:
            msg = self._lib.klm_last_error().decode('utf-8', 'replace')
            if rc == -1:
                raise ValueError(msg)
            raise NativeError('libkeraslm_b200 error %d: %s' % (rc, msg))

    def close(self):
        if self._h is not None:
            self._lib.klm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        self.check(self._lib.klm_set_stream(self._h, _P(cuda_stream) if cuda_stream else None))

    def sync(self):
        self.check(self._lib.klm_sync(self._h))

    def set_option(self, name, value):
        """Run-time switch of the library (see klm_set_option)."""
        self.check(self._lib.klm_set_option(self._h, name.encode('ascii'), int(value)))

    def info(self):
        inf = KlmInfo()
        self.check(self._lib.klm_get_info(self._h, ctypes.byref(inf)))
        return dict((f, getattr(inf, f)) for f, _ in KlmInfo._fields_)

    def beam_edge(self, hyp_slot, hyp_last_char, hyp_cum_cost, alt_offsets, alt_chars, alt_conf, ctx, lm_weight,
                  batch_size, scratch_slots, max_out):
        """See klm_beam_edge.  Returns (hyp, alt, cum_cost, slot) arrays of the finished candidates in beam order
        and (device steps, hypothesis steps)."""
        hs, hl = _i32(hyp_slot), _i32(hyp_last_char)
        hc = np.ascontiguousarray(hyp_cum_cost, dtype=np.float64)
        ao, ac = _i32(alt_offsets), _i32(alt_chars)
        af = np.ascontiguousarray(alt_conf, dtype=np.float64)
        cx = None if ctx is None or len(ctx) == 0 else _i32(ctx)
        sc = _i32(scratch_slots)
        oh, oa = np.empty(max_out, np.int32), np.empty(max_out, np.int32)
        oc, os_ = np.empty(max_out, np.float64), np.empty(max_out, np.int32)
        n_out = ctypes.c_int32(0)
        st, hst = ctypes.c_int64(0), ctypes.c_int64(0)
        self.check(self._lib.klm_beam_edge(self._h, len(hs), _ptr(hs), _ptr(hl), _ptr(hc), len(af), _ptr(ao), _ptr(ac),
                                           _ptr(af), _ptr(cx), float(lm_weight), int(batch_size), len(sc), _ptr(sc),
                                           int(max_out), _ptr(oh), _ptr(oa), _ptr(oc), _ptr(os_), ctypes.byref(n_out),
                                           ctypes.byref(st), ctypes.byref(hst)))
        k = n_out.value
        return oh[:k], oa[:k], oc[:k], os_[:k], st.value, hst.value

    def beam_lattice(self, hyp_slot, hyp_last_char, hyp_cum_cost, edge_alt_first, alt_offsets, alt_chars, alt_conf, ctx,
                     lm_weight, batch_size, beam_width, scratch_slots):
        """See klm_beam_lattice.  Returns (out_first, out_hyp, out_alt, out_cum_cost, final_slot, steps, hypothesis steps)."""
        hs, hl = _i32(hyp_slot), _i32(hyp_last_char)
        hc = np.ascontiguousarray(hyp_cum_cost, dtype=np.float64)
        ef, ao, ac = _i32(edge_alt_first), _i32(alt_offsets), _i32(alt_chars)
        af = np.ascontiguousarray(alt_conf, dtype=np.float64)
        cx = None if ctx is None or len(ctx) == 0 else _i32(ctx)
        sc = _i32(scratch_slots)
        n_edges = len(ef) - 1
        max_out = n_edges * int(beam_width)
        of = np.zeros(n_edges + 1, np.int32)
        oh, oa = np.empty(max(max_out, 1), np.int32), np.empty(max(max_out, 1), np.int32)
        oc = np.empty(max(max_out, 1), np.float64)
        fs = np.empty(int(beam_width), np.int32)
        st, hst = ctypes.c_int64(0), ctypes.c_int64(0)
        self.check(self._lib.klm_beam_lattice(self._h, len(hs), _ptr(hs), _ptr(hl), _ptr(hc), n_edges, _ptr(ef), _ptr(ao),
                                              _ptr(ac), _ptr(af), _ptr(cx), float(lm_weight), int(batch_size),
                                              int(beam_width), len(sc), _ptr(sc), max_out, _ptr(of), _ptr(oh), _ptr(oa),
                                              _ptr(oc), _ptr(fs), ctypes.byref(st), ctypes.byref(hst)))
        return of, oh, oa, oc, fs, st.value, hst.value

    def tc_trace_arm(self, first_tick):
        self.check(self._lib.klm_tc_trace(self._h, first_tick, None, None))

    def tc_trace_read(self):
        grid = self.info()['tc_grid']
        out = np.zeros((grid, 16, 8), dtype=np.uint64)
        roles = np.zeros((grid, 4), dtype=np.int32)
        self.check(self._lib.klm_tc_trace(self._h, 0, _ptr(out), _ptr(roles)))
        return out, roles

    # ---- state pool ----
    def pool_resize(self, n):
        if n > self.n_slots:
            self.check(self._lib.klm_state_pool_resize(self._h, n))
            self.n_slots = n

    def state_reset(self, slots):
        s = _i32(slots)
        self.check(self._lib.klm_state_reset(self._h, _ptr(s), len(s)))

    def state_copy(self, src, dst):
        a, b = _i32(src), _i32(dst)
        assert len(a) == len(b)
        self.check(self._lib.klm_state_copy(self._h, _ptr(a), _ptr(b), len(a)))

    def state_to_host(self, slots):
        s = _i32(slots)
        out = np.empty((len(s), 2 * self.depth, self.width), dtype=np.float32)
        self.check(self._lib.klm_state_to_host(self._h, _ptr(s), len(s), _ptr(out)))
        return out

    def state_from_host(self, slots, values):
        s = _i32(slots)
        v = np.ascontiguousarray(values, dtype=np.float32).reshape(len(s), 2 * self.depth, self.width)
        self.check(self._lib.klm_state_from_host(self._h, _ptr(s), len(s), _ptr(v)))

    def state_l2_distance(self, a, b, n_arrays):
        a, b = _i32(a), _i32(b)
        out = np.empty((len(a), n_arrays), dtype=np.float32)
        self.check(self._lib.klm_state_l2_distance(self._h, _ptr(a), _ptr(b), len(a), n_arrays, _ptr(out)))
        return out

    def state_l2_distance_noalloc(self, a, b, n_arrays, out=None):
        """klm_state_l2_distance_noalloc: the same distances through the static device scratch (no device allocation;
        at most NOALLOC_SCRATCH_BYTES; calls that overlap on one GPU race)."""
        a, b = _i32(a), _i32(b)
        if out is None:
            out = np.empty((len(a), n_arrays), dtype=np.float32)
        self.check(self._lib.klm_state_l2_distance_noalloc(self._h, _ptr(a), _ptr(b), len(a), n_arrays, _ptr(out)))
        return out

    # ---- the path ----
    def rate(self, slots, offsets, chars, targets, n_real, ctx, out, mem=MEM_HOST):
        """See klm_rate.  chars/targets/out: numpy arrays (MEM_HOST) or device addresses (MEM_DEVICE)."""
        slots = _i32(slots)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        n_real = None if n_real is None else _i32(n_real)
        ctx = None if ctx is None else _i32(ctx)
        self.check(self._lib.klm_rate(self._h, len(slots), _ptr(slots), _ptr(offsets), _ptr(chars), _ptr(targets),
                                      _ptr(n_real), _ptr(ctx), _ptr(out), mem))

    def step(self, chars, ctx, slots_in, slots_out, targets, out, mem=MEM_HOST):
        chars = _i32(chars)
        ctx = None if ctx is None else _i32(ctx)
        si, so = _i32(slots_in), _i32(slots_out)
        targets = None if targets is None else _i32(targets)
        self.check(self._lib.klm_step(self._h, len(chars), _ptr(chars), _ptr(ctx), _ptr(si), _ptr(so), _ptr(targets),
                                      _ptr(out), mem))

    def step_topk(self, chars, ctx, slots_in, slots_out, k):
        """See klm_step_topk.  Returns (indices [n, k] int32, probabilities [n, k] float32), ascending by probability."""
        chars = _i32(chars)
        ctx = None if ctx is None else _i32(ctx)
        si, so = _i32(slots_in), _i32(slots_out)
        idx = np.empty((len(chars), k), dtype=np.int32)
        prob = np.empty((len(chars), k), dtype=np.float32)
        self.check(self._lib.klm_step_topk(self._h, len(chars), _ptr(chars), _ptr(ctx), _ptr(si), _ptr(so), int(k), _ptr(idx),
                                           _ptr(prob)))
        return idx, prob
