"""Sharding of independent rating units (documents, pages, lines) over the GPUs of a node.

The reference has no multi-GPU code at all; units that the reference would
rate from a reset state share nothing but the read-only weights, so the B200 build runs one
process per GPU (``torchrun``), every rank rates its own units with its own ``Rater`` and no
collective touches the data path.  The only communication is the optional gather of the
per-character scores at the end (NCCL all-gather over NVLink, or gloo on CPU for the tests).
"""
import numpy as np


def partition(lengths, world_size):
    """Longest-processing-time-first assignment of units to ranks.

    Returns a list of ``world_size`` index lists; deterministic (ties broken by index), so every
    rank computes the same partition without communicating.  Large inputs go through the library's
    host helper ``klm_lpt_partition`` (10 000 lines: 0.4 ms instead of 5.8 ms); `_partition_py` states the rule.
    """
    lengths = np.asarray(lengths, dtype=np.int64)
    if len(lengths) < NATIVE_PARTITION_MIN:
        return _partition_py(lengths, world_size)
    from . import _native
    rank_of = _native.lpt_partition(lengths, world_size)
    order = np.argsort(rank_of, kind='stable')                    # indices grouped by rank, ascending within a rank
    counts = np.bincount(rank_of, minlength=world_size)
    edges = [0] + np.cumsum(counts).tolist()
    flat = order.tolist()
    return [flat[a:b] for a, b in zip(edges, edges[1:])]


NATIVE_PARTITION_MIN = 512


def _partition_py(lengths, world_size):
    import heapq
    order = np.lexsort((np.arange(len(lengths)), -lengths))       # longest first, ties by index
    heap = [(0, r) for r in range(world_size)]                    # (load, rank): least loaded first, ties by rank
    parts = [[] for _ in range(world_size)]
    costs = (lengths + 1).tolist()                                # +1: an empty unit still costs a slot
    for i in order.tolist():
        load, r = heap[0]
        parts[r].append(i)
        heapq.heapreplace(heap, (load + costs[i], r))
    for p in parts:
        p.sort()
    return parts


def rate_sharded(rater, texts, contexts=None, gather=True, group=None):
    """Rate `texts` (independent units, each from a zero state) across the ranks of the default
    process group: rank r rates ``partition(...)[r]`` with ``rater.rate_many``.

    gather=True  : every rank returns the full list (one float32 array of probabilities per text)
    gather=False : every rank returns a list with its own units filled in and None elsewhere
    Without an initialised process group this is just ``rater.rate_many``.

    Texts are NFC-normalised HERE, once, so that the partition, the buffer sizes and the unpacking offsets on every
    rank are derived from the lengths ``rate_many`` will return (it normalises, too: a decomposed sequence shrinks).
    A rank whose local rating fails still takes part in the collectives: an error flag is reduced first, and every
    rank raises if any rank failed -- nobody is left waiting in the gather.
    """
    import unicodedata
    import torch
    import torch.distributed as dist
    texts = [t if t.isascii() else unicodedata.normalize('NFC', t) for t in texts]
    n = len(texts)
    if contexts is None:
        contexts = [None] * n
    if not (dist.is_available() and dist.is_initialized()):
        return rater.rate_many(texts, contexts)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lens = np.fromiter(map(len, texts), dtype=np.int64, count=n)
    parts = partition(lens, world)
    mine = parts[rank]
    out = [None] * n
    failure = None
    flat_mine = None                                       # this rank's scores as ONE array, if rate_many has them that way
    try:
        res = rater.rate_many([texts[i] for i in mine], [contexts[i] for i in mine]) if mine else []
        if len(res) != len(mine):
            raise ValueError('rate_many returned %d results for %d texts' % (len(res), len(mine)))
        got = np.fromiter(map(len, res), dtype=np.int64, count=len(res))
        wrong = np.flatnonzero(got != lens[mine]) if mine else got
        if wrong.size:
            j = int(wrong[0])
            raise ValueError('rate_many returned %d scores for a text of %d characters' % (got[j], lens[mine[j]]))
        for i, r in zip(mine, res):
            out[i] = r if type(r) is np.ndarray and r.dtype == np.float32 else np.asarray(r, dtype=np.float32)
        flat_mine = getattr(res, 'flat', None)
    except Exception as e:                                 # pylint: disable=broad-except
        failure = e
    if world == 1:
        if failure is not None:
            raise failure
        return out
    backend = dist.get_backend(group)
    dev = torch.device('cuda', torch.cuda.current_device()) if backend == 'nccl' else torch.device('cpu')
    flag = torch.tensor([1 if failure is not None else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=group)
    if failure is not None:
        raise failure
    if int(flag.item()):
        raise RuntimeError('rate_sharded: rating failed on another rank')
    if not gather:
        return out
    # one padded all-gather of the flat score arrays; sizes are known to everybody
    sizes = [int(lens[p].sum()) if p else 0 for p in parts]
    cap = max(max(sizes), 1)
    flat = np.zeros(cap, dtype=np.float32)
    if mine and sizes[rank]:
        if flat_mine is not None and flat_mine.shape == (sizes[rank],):
            flat[:sizes[rank]] = flat_mine
        else:
            np.concatenate([out[i] for i in mine], out=flat[:sizes[rank]])
    send = torch.from_numpy(flat).to(dev)
    recv = torch.empty(world * cap, dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(recv, send, group=group)
    everything = recv.cpu().numpy().reshape(world, cap)
    for r, p in enumerate(parts):
        if r == rank or not p:
            continue
        # views into the gathered row, cut at the cumulative lengths of that rank's units (no copy per text)
        row = everything[r]
        edges = [0] + np.cumsum(lens[p]).tolist()
        for i, a, b in zip(p, edges, edges[1:]):
            out[i] = row[a:b]
    return out
