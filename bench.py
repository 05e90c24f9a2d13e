#!/usr/bin/env python
"""Benchmark of the stateful rating path (BASELINE.json metric: chars/sec rated).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload cfg2|cfg1|cfg4|cfg5] [--docs D] [--chars C]

One "step" = one pass of the hot path over one batch of synthetic documents: every document
is rated from a zero state, all documents of the batch advancing concurrently in the
persistent kernel (cfg2: 64 documents x 100 000 characters, width 512, depth 3, length 512).

Our arm prints ONE JSON line with
  value        chars/s, inputs resident in HBM when the timed region starts (C ABI, KLM_MEM_DEVICE),
               CUDA events on the launching stream, L2 flushed between timed steps, max over ranks
  e2e          the same metric through the public API (`Rater.rate_many` on Python strings):
               tokenisation, pinned host buffers, H2D, kernel, D2H inside the timed region
  roofline     the persistent sequence kernel against the measured tensor peak (MEASURED_PEAKS.json)
  cpu_baseline the CPU oracle (oracle/, a port: the reference's Keras CPU path cannot be installed
               here) timed on the host cores on a bounded sample (rank 0, N=1 only)
  cpu_baseline_torch  a second, stronger CPU line: all documents batched through torch CPU
               (oracle/lstm_torch_cpu.py), so that the speed-up is not inflated by a slow baseline
`--impl reference` times that CPU arm alone, on the same config, with every host core.
Multi-GPU (torchrun): documents are independent, every rank rates its own batch, no data-path
collective; NCCL only carries the timing reduction and a checksum.
"""
import argparse
import json
import logging
import math
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from ocrd_keraslm_b200 import synth  # noqa: E402

WORKLOADS = {
    #        width depth length contexts            docs   chars   what
    'cfg1': (128, 2, 256, ((200, 10),), 1, 100000, 'stateful rating width=128 depth=2 length=256, 1 document x 100000 chars'),
    'cfg2': (512, 3, 512, ((200, 10),), 64, 100000, 'stateful rating width=512 depth=3 length=512, 64 documents x 100000 chars'),
    'cfg3': (512, 3, 512, ((200, 10),), 4, 200, 'rate_best beam search, beam=100, 5 alternatives/word, pages of 200 words chained by traceback, width=512 depth=3'),
    'cfg4': (512, 3, 512, ((200, 16),) * 3, 64, 100000, 'context-conditioned (3 context variables, dim 16) width=512 depth=3, 64 documents x 100000 chars'),
    'cfg5': (1024, 3, 512, ((200, 10),), 10000, 70, 'page-sharded rating width=1024 depth=3, 10000 lines of 20..120 chars'),
}
CFG_INDEX = {'cfg1': 1, 'cfg2': 2, 'cfg3': 3, 'cfg4': 4, 'cfg5': 5}
FALLBACK_PEAKS = {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0}

quiet = logging.getLogger('bench')
quiet.setLevel(logging.CRITICAL)


def flops_per_char(width, depth, contexts, voc_size=256):
    """F = sum_l 2*4W*(D_l+W) + 2*V*W."""
    d1 = width + sum(d for _, d in contexts)
    return sum(2 * 4 * width * ((d1 if l == 0 else width) + width) for l in range(depth)) + 2 * voc_size * width


def make_docs(workload, n_docs, n_chars, first_doc):
    """Seeded synthetic documents and their context rows."""
    width, depth, length, contexts, _, _, _ = WORKLOADS[workload]
    cfg = CFG_INDEX[workload]
    texts, ctxs = [], []
    for d in range(first_doc, first_doc + n_docs):
        seed = 20261104 + 1000 * cfg + d
        rng = np.random.default_rng(seed)
        size = n_chars if workload != 'cfg5' else int(rng.integers(20, 121))
        texts.append(synth.text(seed, size))
        if len(contexts) == 1:
            ctxs.append([179])
        else:
            row = [int(x) for x in rng.integers(0, 200, len(contexts))]
            if rng.random() < 0.1:
                row = [0] * len(contexts)
            ctxs.append(row)
    return texts, ctxs


def model_weights(workload):
    width, depth, length, contexts, _, _, _ = WORKLOADS[workload]
    return synth.weights(7 + CFG_INDEX[workload], width=width, depth=depth, voc_size=256, contexts=contexts, style='keras')


def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            p = json.load(f)
        return p, 'measured (MEASURED_PEAKS.json)'
    except Exception:
        return dict(FALLBACK_PEAKS), 'fallback (B200_PROFILING.md)'


# ------------------------------------------------------------------------------------------ CPU arm
_CPU = {}


def _cpu_worker(job):
    """Rate one sample document with the oracle, reference call structure (one model call per window,
    batch 1: rating.py:516), single BLAS thread; documents run in parallel across processes."""
    from threadpoolctl import threadpool_limits
    from oracle.rater_oracle import OracleRater
    doc, n_chars = job
    with threadpool_limits(limits=1):
        o = OracleRater(_CPU['model'], _CPU['mapping'], _CPU['length'], logger=quiet)
        t0 = time.perf_counter()
        probs = o.rate(_CPU['texts'][doc][:n_chars], _CPU['ctxs'][doc])
        dt = time.perf_counter() - t0
    return len(probs), dt, float(np.log(np.maximum(np.asarray(probs[1:], dtype=np.float64), 1e-300)).sum())


class CpuArm(object):
    """The oracle port on the host cores: one process per core, one document each."""

    def __init__(self, workload, procs=None):
        import multiprocessing as mp
        from tests import util
        width, depth, length, contexts, _, _, _ = WORKLOADS[workload]
        self.procs = procs or len(os.sched_getaffinity(0))
        wd = model_weights(workload)
        _CPU['model'] = util.to_model(wd)
        _CPU['mapping'] = synth.mapping_from_array(synth.vocabulary())
        _CPU['length'] = length
        self.workload = workload
        self.pool = None
        self.mp = mp.get_context('fork')

    def prepare(self, n_chars):
        texts, ctxs = make_docs(self.workload, self.procs, max(n_chars, 2), 0)
        _CPU['texts'], _CPU['ctxs'] = texts, ctxs
        self.pool = self.mp.Pool(self.procs)          # fork AFTER the globals are in place

    def step(self, n_chars):
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_worker, [(d, n_chars) for d in range(self.procs)], chunksize=1)
        dt = time.perf_counter() - t0
        return sum(r[0] for r in res), dt

    def close(self):
        if self.pool is not None:
            self.pool.close()
            self.pool.join()
            self.pool = None


def torch_cpu_line(workload, rater_indices, seconds=8.0, n_docs=0, n_chars=0):
    """Second, stronger CPU line: ALL documents of the workload advance together through
    oracle/lstm_torch_cpu.py (one addmm per layer and tick, torch CPU on all host cores) -- not the reference's
    call structure (batch 1), so it is reported beside `cpu_baseline`, not in its place.  Bounded sample: the first
    characters of every document, sized from a short calibration run to about `seconds` of CPU work."""
    import torch
    from tests import util
    from oracle.lstm_torch_cpu import TorchCpuModel
    width, depth, length, contexts, docs, chars, what = WORKLOADS[workload]
    docs, chars = n_docs or docs, n_chars or chars
    if workload in ('cfg3', 'cfg5') or chars < 66:
        return None                                   # beams / ragged lines: the batched form has no fixed shape
    model = TorchCpuModel(util.to_model(model_weights(workload)))
    cal = 24
    n_max = min(chars - 1, 4096)
    texts, ctxs = make_docs(workload, docs, chars, 0)      # the GPU arm's documents (their prefixes are what is compared)
    idx = np.stack([rater_indices(t[:n_max + 1]) for t in texts]).astype(np.int64)
    zs = [np.array([c[k] for c in ctxs], dtype=np.int64) for k in range(len(contexts))]
    model.rate(idx[:, :cal], zs, idx[:, 1:cal + 1])
    t0 = time.perf_counter()
    model.rate(idx[:, :cal], zs, idx[:, 1:cal + 1])
    per_tick = (time.perf_counter() - t0) / cal
    n = int(max(64, min(n_max, seconds / per_tick)))
    t0 = time.perf_counter()
    logp, _ = model.rate(idx[:, :n], zs, idx[:, 1:n + 1])
    dt = time.perf_counter() - t0
    line = {'value': docs * n / dt, 'unit': 'chars/s', 'cores': torch.get_num_threads(),
            'kind': 'port (batched: all %d documents per addmm, torch CPU / MKL -- not the reference call structure)' % docs,
            'sample': '%d documents x first %d chars together (of %s), %.1f s' % (docs, n, what, dt),
            'checksum_logp': float(logp.astype(np.float64).sum())}
    # the same positions once more in fp64: the truth the GPU scores of this very run are compared with (`parity`)
    t0 = time.perf_counter()
    ref64, _ = TorchCpuModel(util.to_model(model_weights(workload)), dtype=np.float64).rate(idx[:, :n], zs, idx[:, 1:n + 1])
    line['_ref64'] = ref64
    line['_ref64_s'] = time.perf_counter() - t0
    line['fp32_vs_fp64_max_abs_dlogp'] = float(np.abs(logp.astype(np.float64) - ref64).max())
    return line


def cpu_sample_chars(workload):
    """Characters per process and step: sized for ~5-8 s of CPU work per step."""
    width, depth = WORKLOADS[workload][0], WORKLOADS[workload][1]
    per_char = 2.0e-3 * (width / 512.0) ** 2 * (depth / 3.0)      # measured ~1.6 ms/char at W512 D3 on one core
    if workload == 'cfg5':
        return 120
    return int(min(WORKLOADS[workload][5], max(513, 6.0 / per_char)))


def run_reference(args):
    """`--impl reference`: the CPU arm alone (rank 0 only)."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    wl = args.workload
    arm = CpuArm(wl)
    n_chars = cpu_sample_chars(wl)
    arm.prepare(n_chars)
    for _ in range(args.warmup):
        arm.step(min(n_chars, 257))                   # warm-up: page in BLAS, fork pool, short sample
    total_chars, total_t = 0, 0.0
    for _ in range(args.steps):
        c, dt = arm.step(n_chars)
        total_chars += c
        total_t += dt
    arm.close()
    value = total_chars / total_t
    line = base_line(args, value, 1000.0 * total_t / max(args.steps, 1))
    sample = '%d processes x 1 document x %d chars per step (of %s)' % (arm.procs, n_chars, WORKLOADS[wl][6])
    line.update({
        'impl': 'reference',
        'dtype': 'f32',
        'cpu_baseline': {'value': value, 'unit': 'chars/s', 'cores': arm.procs, 'kind': 'port', 'sample': sample,
                         'note': 'reference Keras/TF CPU path is not installable here (no keras/tensorflow/h5py wheels, '
                                 'no network): oracle/ port of its algorithm with the reference call structure '
                                 '(batch 1, one model call per window), documents spread over all host cores'},
        'e2e': {'value': value, 'unit': 'chars/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    })
    print(json.dumps(line))



# ------------------------------------------------------------------------------------------ beam search (cfg3)
def beam_lattices(n_pages, n_words, first_page=0):
    """The synthetic input of cfg3: one lattice per page (built OUTSIDE the timed region, like the texts of the other
    workloads -- generating a 200-word graph costs ~50 ms of Python that is not part of the rating path)."""
    return [synth.lattice(20261104 + 3000 + pg, n_words=n_words, n_alts=5) for pg in range(first_page, first_page + n_pages)]


def beam_pages(rater, n_pages, n_words, first_page=0, lattices=None):
    """rate_best over `n_pages` synthetic lattices chained through the traceback (wrapper/rate.py:186-290)."""
    tb = None
    chosen = []
    for graph, start, end in (lattices if lattices is not None else beam_lattices(n_pages, n_words, first_page)):
        path, entropy, tb = rater.rate_best(graph, start, end, start_traceback=tb, context=[179], lm_weight=0.5,
                                            beam_width=100, beam_clustering_dist=0)
        chosen.append([alt.Unicode for _, alt, _ in path])
    path, entropy, _ = rater.next_path(tb[0], ([], tb[1]))
    chosen.append([alt.Unicode for _, alt, _ in path])
    return chosen


def run_beam(args):
    """`--workload cfg3`: beam steps/sec of rate_best (one step = one batched advance of all live hypotheses
    by one character).  Our arm: Rater.rate_best with the device step kernel; reference arm: the oracle port."""
    width, depth, length, contexts, pages, words, what = WORKLOADS['cfg3']
    pages = args.docs or pages
    words = args.chars or words
    wd = model_weights('cfg2')                      # cfg3 uses the model of cfg2
    if args.impl == 'reference':
        if int(os.environ.get('RANK', '0')) != 0:
            return
        from tests import util
        from oracle.rater_oracle import OracleRater
        o = OracleRater(util.to_model(wd), synth.mapping_from_array(synth.vocabulary()), length, logger=quiet)
        words_cpu = min(words, 12)
        beam_pages(o, 1, 4)
        lats = [beam_lattices(1, words_cpu) for _ in range(args.steps)]
        t0 = time.perf_counter()
        c0 = o.device_calls
        for k in range(args.steps):
            beam_pages(o, 1, words_cpu, lattices=lats[k])
        dt = time.perf_counter() - t0
        steps = o.device_calls - c0
        line = base_line(args, steps / dt, 1000.0 * dt / args.steps)
        line.update({'metric': 'beam steps/sec (rate_best)', 'unit': 'steps/s', 'impl': 'reference', 'gpu_launches': 0,
                     'cpu_baseline': {'value': steps / dt, 'unit': 'steps/s', 'cores': len(os.sched_getaffinity(0)), 'kind': 'port',
                                      'sample': '%d page(s) of %d words per step, numpy BLAS threads' % (1, words_cpu)},
                     'e2e': {'value': steps / dt, 'unit': 'steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}})
        print(json.dumps(line))
        return
    import torch
    from ocrd_keraslm_b200.lib import Rater
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    r = Rater.from_weights(wd, synth.vocabulary(), length, stateful=False, incremental=True, logger=quiet, device=local)
    for _ in range(min(args.warmup, 2)):
        beam_pages(r, 1, 10)
    r._native.sync()
    l0 = r._native.info()['kernel_launches']
    s0, h0 = r.device_steps, r.hypothesis_steps
    lats = [beam_lattices(pages, words) for _ in range(args.steps)]          # (fresh graph objects per step, built before the clock)
    t0 = time.perf_counter()
    for k in range(args.steps):
        beam_pages(r, pages, words, lattices=lats[k])
    r._native.sync()
    dt = time.perf_counter() - t0
    steps, hyps = r.device_steps - s0, r.hypothesis_steps - h0
    line = base_line(args, steps / dt, 1000.0 * dt / args.steps)
    F = flops_per_char(width, depth, contexts)
    pk, pk_src = peaks()
    line.update({'metric': 'beam steps/sec (rate_best)', 'unit': 'steps/s',
                 'hypothesis_steps_per_s': hyps / dt, 'mean_hypotheses_per_step': hyps / max(steps, 1),
                 'gpu_launches': r._native.info()['kernel_launches'] - l0,
                 'e2e': {'value': steps / dt, 'unit': 'steps/s', 'h2d_bytes_per_step': int(12 * hyps / args.steps),
                         'd2h_bytes_per_step': int(4 * hyps / args.steps), 'api': 'Rater.rate_best(graph, ...)'},
                 'roofline': {'bound': 'tensor', 'kernel': 'klm_seq_df (step mode: a resident instance per rate_best call, steps posted through mapped memory)',
                              'achieved': hyps / dt * F / 1e12,
                              'peak': float(pk.get('bf16_tflops_sustained', FALLBACK_PEAKS['bf16_tflops_sustained'])), 'unit': 'TFLOP/s',
                              'frac': hyps / dt * F / 1e12 / float(pk.get('bf16_tflops_sustained', FALLBACK_PEAKS['bf16_tflops_sustained'])),
                              'traffic': None, 'peak_source': pk_src + ', bf16 sustained',
                              'note': 'achieved = algorithmic FLOPs of all hypothesis steps / WALL CLOCK of the whole search (host-side '
                                      'beam logic included); a beam step is a latency chain (REC -> IN -> REC -> ... -> PROJ -> LSE for one '
                                      'position), not a throughput problem: KLM_STEP_TIMING=1 prints the device share per step'}})
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and clock-event reasons of one GPU every 100 ms while the timed region runs."""

    def __init__(self, device_index):
        threading.Thread.__init__(self, daemon=True)
        self.stop_ev = threading.Event()
        self.sm = []
        self.reasons = set()
        self.sm_max = None
        self.power = []
        self.err = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get('CUDA_VISIBLE_DEVICES')
            idx = device_index
            if vis:
                ids = [v.strip() for v in vis.split(',') if v.strip()]
                if device_index < len(ids) and ids[device_index].isdigit():
                    idx = int(ids[device_index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:                         # noqa: BLE001
            self.nv = None
            self.err = repr(e)

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksEventReasonHwSlowdown: 'hw_slowdown',
            nv.nvmlClocksEventReasonHwThermalSlowdown: 'hw_thermal_slowdown',
            nv.nvmlClocksEventReasonSwThermalSlowdown: 'sw_thermal_slowdown',
            nv.nvmlClocksEventReasonSwPowerCap: 'sw_power_cap',
            nv.nvmlClocksEventReasonHwPowerBrakeSlowdown: 'hw_power_brake_slowdown',
            nv.nvmlClocksEventReasonApplicationsClocksSetting: 'applications_clocks_setting',
        }
        while not self.stop_ev.is_set():
            try:
                self.sm.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception as e:                     # noqa: BLE001
                self.err = repr(e)
                break
            self.stop_ev.wait(0.1)

    def finish(self):
        self.stop_ev.set()
        self.join(timeout=2.0)
        out = {'sm_mhz': float(np.median(self.sm)) if self.sm else None, 'sm_max_mhz': self.sm_max,
               'reasons': sorted(self.reasons), 'samples': len(self.sm)}
        if self.power:
            out['power_w_max'] = round(max(self.power), 1)
        if self.err:
            out['error'] = self.err
        return out


# ------------------------------------------------------------------------------------------ our arm
def base_line(args, value, ms_per_step):
    width, depth, length, contexts, docs, chars, what = WORKLOADS[args.workload]
    docs = args.docs or docs
    chars = args.chars or chars
    return {
        'metric': 'chars/sec rated (stateful LSTM LM)',
        'value': value,
        'unit': 'chars/s',
        'n_gpus': args.gpus,
        'steps': args.steps,
        'warmup': args.warmup,
        'ms_per_step': ms_per_step,
        'higher_is_better': True,
        'scaling': 'weak',
        'vs_baseline': None,
        'dtype': 'f32',
        'data': 'synthetic (seeded Zipf-like text, random-init Keras-default weights; SURVEY.md 8d)',
        'config': {'workload': '%s: %s' % (args.workload, what), 'width': width, 'depth': depth, 'length': length,
                   'voc_size': 256, 'n_ctx': len(contexts), 'ctx_dim': contexts[0][1],
                   'docs_per_gpu': docs, 'chars_per_doc': chars if args.workload != 'cfg5' else 'U{20..120}',
                   'flops_per_char': flops_per_char(width, depth, contexts),
                   'l2': 'flushed (512 MiB write) between timed steps'},
    }


def run_sharded(args):
    """`--sharded`: the page-sharded form BASELINE.json words for cfg5 -- every rank holds the SAME list of units (10 000
    lines), `shard.rate_sharded` gives each rank its longest-processing-time-first share, rates it with `Rater.rate_many`
    and gathers all scores on every rank (NCCL all_gather) -- all inside the timed region, through the public API, from
    Python strings.  STRONG scaling: the work is fixed, value = characters of the whole list / wall clock of the slowest
    rank."""
    import torch
    import torch.distributed as dist
    from ocrd_keraslm_b200 import shard
    from ocrd_keraslm_b200.lib import Rater
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    args.gpus = world
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    wl = args.workload
    width, depth, length, contexts, docs, chars, what = WORKLOADS[wl]
    docs = args.docs or docs
    chars = args.chars or chars
    rater = Rater.from_weights(model_weights(wl), synth.vocabulary(), length, logger=quiet, device=local)
    texts, ctxs = make_docs(wl, docs, chars, 0)                       # the same units on every rank
    total_chars = sum(max(len(t) - 1, 0) for t in texts)

    def barrier():
        if world > 1:
            dist.barrier()

    for _ in range(args.warmup):
        res = shard.rate_sharded(rater, texts, ctxs, gather=True)
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    l0 = rater._native.info()['kernel_launches']
    barrier()
    sampler.start()
    phases = {'rate_many_s': 0.0, 'total_s': 0.0}
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = shard.rate_sharded(rater, texts, ctxs, gather=True)
    torch.cuda.synchronize()
    dt_local = time.perf_counter() - t0
    barrier()
    clocks = sampler.finish()
    launches = rater._native.info()['kernel_launches'] - l0
    # how the slowest rank's time splits: its own rating (host tokenisation + launches + copies) vs the rest (gather, unpack)
    t1 = time.perf_counter()
    mine = shard.partition([len(t) for t in texts], world)[rank]
    rater.rate_many([texts[i] for i in mine], [ctxs[i] for i in mine])
    rate_only = time.perf_counter() - t1
    dt = dt_local
    if world > 1:
        t = torch.tensor([dt_local, rate_only, float(launches)], dtype=torch.float64, device='cuda')
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dt, rate_only, launches = float(tm[0].item()), float(tm[1].item()), int(t[2].item())
    checksum = float(sum(np.log(np.maximum(r[1:].astype(np.float64), 1e-300)).sum() for r in res))
    value = total_chars * args.steps / dt
    line = base_line(args, value, 1000.0 * dt / args.steps)
    line['scaling'] = 'strong'
    line['config']['units'] = '%d lines, the same list on every rank; LPT partition over %d rank(s); scores gathered on every rank' % (docs, world)
    line['config'].pop('docs_per_gpu', None)
    line.update({
        'e2e': {'value': value, 'unit': 'chars/s', 'h2d_bytes_per_step': 8 * total_chars, 'd2h_bytes_per_step': 4 * total_chars,
                'api': 'shard.rate_sharded(rater, texts, contexts, gather=True) -> Rater.rate_many on the rank\'s share + all_gather'},
        'gpu_launches': launches, 'clocks': clocks, 'checksum_logp': checksum,
        'per_step_ms': {'whole': 1000.0 * dt / args.steps, 'rating_of_the_largest_share': 1000.0 * rate_only,
                        'gather_and_unpack': 1000.0 * (dt / args.steps - rate_only)},
        'note': 'no device-resident leg in this mode: value is the end-to-end number',
    })
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_ours(args):
    import torch
    import torch.distributed as dist
    from ocrd_keraslm_b200 import _native
    from ocrd_keraslm_b200.lib import Rater

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit('bench.py --gpus %d must be launched with torchrun (one rank per GPU)' % args.gpus)
        args.gpus = world
    # ---- CPU baseline beside it (rank 0, N=1 only); runs first, before CUDA is initialised (it forks) ----
    cpu_line = torch_line = None
    if world == 1 and not args.no_cpu:
        arm = CpuArm(args.workload)
        n_chars = cpu_sample_chars(args.workload)
        arm.prepare(n_chars)
        arm.step(min(n_chars, 129))
        c, dt = arm.step(n_chars)
        arm.close()
        vocab_map = synth.mapping_from_array(synth.vocabulary())[0]
        torch_line = torch_cpu_line(args.workload, lambda t: np.array([vocab_map.get(ch, 0) for ch in t], dtype=np.int64),
                                    n_docs=args.docs, n_chars=args.chars)
        cpu_line = {
            'value': c / dt, 'unit': 'chars/s', 'cores': arm.procs, 'kind': 'port',
            'sample': '%d processes x 1 document x %d chars (of %s), %.1f s' % (arm.procs, n_chars,
                                                                               WORKLOADS[args.workload][6], dt)}
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)')
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    wl = args.workload
    width, depth, length, contexts, docs, chars, what = WORKLOADS[wl]
    docs = args.docs or docs
    chars = args.chars or chars
    F = flops_per_char(width, depth, contexts)

    wd = model_weights(wl)
    rater = Rater.from_weights(wd, synth.vocabulary(), length, logger=quiet, device=local)
    native = rater._native
    texts, ctxs = make_docs(wl, docs, chars, rank * docs)

    # ---- device-resident form of the batch (tokenised once, outside the timed region) ----
    lens = np.array([max(len(t) - 1, 0) for t in texts], dtype=np.int64)
    order = np.argsort(-lens, kind='stable')
    offsets = np.zeros(docs + 1, dtype=np.int64)
    offsets[1:] = np.cumsum(lens[order])
    total = int(offsets[-1])
    h_chars = np.empty(total, dtype=np.int32)
    h_targets = np.empty(total, dtype=np.int32)
    for j, i in enumerate(order):
        idx, _ = rater._indices(texts[i])
        h_chars[offsets[j]:offsets[j + 1]] = idx[:-1]
        h_targets[offsets[j]:offsets[j + 1]] = idx[1:]
    ctx_rows = np.array([ctxs[i] for i in order], dtype=np.int32)
    d_chars = torch.from_numpy(h_chars).cuda()
    d_targets = torch.from_numpy(h_targets).cuda()
    d_out = torch.empty(total, dtype=torch.float32, device='cuda')
    flush = torch.empty(512 << 20, dtype=torch.uint8, device='cuda')
    stream = torch.cuda.Stream()              # a real (non-NULL) stream: NULL would mean "the library's own"
    torch.cuda.set_stream(stream)
    native.set_stream(stream.cuda_stream)
    slots = rater._pool.acquire(docs)
    chars_per_step = int(lens.sum())                 # characters scored (the first of each document is free)

    def device_step():
        native.state_reset(slots)
        native.rate(slots, offsets, d_chars.data_ptr(), d_targets.data_ptr(), None, ctx_rows, d_out.data_ptr(),
                    mem=_native.MEM_DEVICE)

    def barrier():
        if world > 1:
            dist.barrier()

    for _ in range(args.warmup):
        device_step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    launches0 = native.info()['kernel_launches']
    barrier()
    torch.cuda.synchronize()
    sampler.start()
    step_ms, kernel_ms = [], []
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()                                # untimed: evict the batch from L2
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        device_step()
        e1.record(stream)
        e1.synchronize()
        step_ms.append(e0.elapsed_time(e1))
        kernel_ms.append(native.info()['last_kernel_ms'])
    torch.cuda.synchronize()
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.finish()
    launches = native.info()['kernel_launches'] - launches0
    checksum = float(d_out.double().sum().item())

    total_ms = float(sum(step_ms))
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        cs = torch.tensor([checksum, float(launches)], dtype=torch.float64, device='cuda')
        dist.all_reduce(cs, op=dist.ReduceOp.SUM)
        checksum, launches = float(cs[0].item()), int(cs[1].item())
    value = world * chars_per_step * args.steps / (total_ms / 1000.0)

    # ---- end to end through the public API: strings in, probabilities out ----
    for _ in range(min(args.warmup, 2)):
        rater.rate_many(texts, ctxs)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = rater.rate_many(texts, ctxs)
    native.sync()
    e2e_t = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_t], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_t = float(t.item())
    e2e_value = world * chars_per_step * args.steps / e2e_t
    e2e_checksum = float(sum(np.log(np.maximum(r[1:].astype(np.float64), 1e-300)).sum() for r in res))
    h2d = 2 * 4 * total + docs * (4 * 2 + 8 + 4 * len(contexts))
    d2h = 4 * total

    # ---- roofline of the dominant kernel (the persistent sequence kernel; one launch per step) ----
    pk, pk_src = peaks()
    info = native.info()
    k_ms = float(np.mean(kernel_ms))
    achieved = chars_per_step * F / (k_ms / 1000.0) / 1e12
    peak = float(pk.get('bf16_tflops_sustained', FALLBACK_PEAKS['bf16_tflops_sustained']))
    traffic = None
    try:
        with open(os.path.join(ROOT, 'profiles', 'traffic.json')) as f:
            tj = json.load(f)
        if tj.get(wl, {}).get('dram_bytes_per_launch_measured') is not None:
            fixed = tj[wl]['dram_bytes_per_launch_measured'] - 8 * tj[wl]['positions_in_measured_launch']
            traffic = fixed + 8 * chars_per_step          # measured one-time part + the per-position streams
    except Exception:
        pass
    roofline = {
        'bound': 'tensor', 'kernel': (('klm_seq_tw' if info.get('tc_wide_launches', 0) > 0 else 'klm_seq_tcs') if info.get('tc_enabled') == 3 else 'klm_seq_tc') if info.get('tc_launches', 0) > 0 else 'klm_seq_fma', 'achieved': achieved, 'peak': peak,
        'unit': 'TFLOP/s', 'frac': achieved / peak, 'traffic': traffic, 'peak_source': pk_src + ', bf16 sustained',
        'kernel_ms': k_ms, 'algorithmic_flops_per_launch': chars_per_step * F,
        'algorithmic_hbm_bytes_per_launch': 8 * chars_per_step,
        'hbm_frac': (8 * chars_per_step / (k_ms / 1000.0) / 1e9) / float(pk.get('hbm_gbs', FALLBACK_PEAKS['hbm_gbs'])),
        'grid': info['tc_grid'] if info.get('tc_launches', 0) > 0 else info['grid'],
        'smem_bytes': info['tc_smem_bytes'] if info.get('tc_launches', 0) > 0 else info['smem_bytes'],
        'weights_in_smem': (0 if info.get('tc_enabled') == 3 else 1) if info.get('tc_launches', 0) > 0 else info['weights_in_smem'],
    }

    if info.get('tc_launches', 0) == 0:
        # <= 8 streams: the FP32 kernels (klm_seq_rg: one cluster, weights in registers; klm_seq_fma_ll) -- the roofline is the
        # measured FP32 FMA peak, and what actually bounds them is the hand-over latency of the recurrence
        fma_peak = _native.measure_fma_peak(local)
        ticks = chars + depth + 1
        roofline.update({
            'bound': 'fp32-fma', 'kernel': 'klm_seq_rg' if info.get('cl_launches', 0) > 0 else 'klm_seq_fma_ll',
            'peak': fma_peak, 'frac': achieved / fma_peak, 'peak_source': 'measured here (klm_measure_fma_peak: FFMA chains on all SMs)',
            'grid': info['cl_cluster'] if info.get('cl_launches', 0) > 0 else info['grid'],
            'latency_floor': {
                'us_per_char_measured': 1e3 * k_ms / max(chars_per_step // max(docs, 1), 1),
                'us_per_char_floor': (215 + 250) / 1965.0,
                'note': 'one stream: every character is one trip of the hidden vector between the SMs that hold the model -- '
                        'distributed shared memory ~215 clk (B300_MICROARCH.md) + the dependent chain of one LSTM cell (~250 clk '
                        'of MUFU latency) at %d MHz; the kernel does not touch the FMA roofline (%d FLOP per character)' % (1965, F)},
        })
        del ticks

    line = base_line(args, value, total_ms / args.steps)
    line.update({
        'e2e': {'value': e2e_value, 'unit': 'chars/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                'api': 'Rater.rate_many(texts, contexts) -> klm_rate(KLM_MEM_HOST)', 'ms_per_step': 1000 * e2e_t / args.steps},
        'gpu_launches': launches,
        'clocks': clocks,
        'roofline': roofline,
        'checksum_logp': checksum,
        'e2e_checksum_logp': e2e_checksum,
        'wall_s_timed_region': wall,
    })

    if cpu_line is not None:
        line['cpu_baseline'] = cpu_line
        if torch_line is not None:
            # parity at the bench configuration itself: the scores this run produced (device-resident leg and the
            # end-to-end leg) against the fp64 CPU oracle on the first n positions of every document
            ref64 = torch_line.pop('_ref64')
            n_par = ref64.shape[1]
            inv = np.empty(docs, dtype=np.int64)
            inv[order] = np.arange(docs)
            h_out = d_out.cpu().numpy()
            dev = np.stack([h_out[offsets[inv[i]]:offsets[inv[i]] + n_par] for i in range(docs)]).astype(np.float64)
            e2e = np.stack([np.log(np.maximum(res[i][1:n_par + 1].astype(np.float64), 1e-300)) for i in range(docs)])
            line['parity'] = {
                'positions': int(ref64.size), 'max_abs_dlogp': float(np.abs(dev - ref64).max()),
                'max_abs_dlogp_e2e': float(np.abs(e2e - ref64).max()), 'tolerance': 1e-3,
                'reference': 'oracle/lstm_torch_cpu.py in fp64, first %d positions of all %d documents (%.1f s)' % (
                    n_par, docs, torch_line.pop('_ref64_s')),
                'checksum_logp_gpu_same_positions': float(dev.sum()), 'checksum_logp_ref': float(ref64.sum())}
            if not line['parity']['max_abs_dlogp'] <= 1e-3 or not line['parity']['max_abs_dlogp_e2e'] <= 1e-3:
                line['parity']['FAILED'] = True
            line['cpu_baseline_torch'] = torch_line
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='cfg2', choices=sorted(WORKLOADS))
    ap.add_argument('--docs', type=int, default=0, help='documents per GPU (default: the workload\'s)')
    ap.add_argument('--chars', type=int, default=0, help='characters per document (default: the workload\'s)')
    ap.add_argument('--no-cpu', action='store_true', help='skip the CPU baseline leg')
    ap.add_argument('--sharded', action='store_true', help='strong scaling: the same units on every rank, rate_sharded + gather timed')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    # stdout carries exactly ONE JSON line: libraries that write to file descriptor 1 on their own (NCCL prints its
    # version there when the process group comes up) are sent to stderr, the JSON goes to the original descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, 'w', buffering=1)
    try:
        if args.workload == 'cfg3':
            run_beam(args)
        elif args.impl == 'reference':
            run_reference(args)
        elif args.sharded:
            run_sharded(args)
        else:
            run_ours(args)
    finally:
        sys.stdout.flush()


if __name__ == '__main__':
    main()
