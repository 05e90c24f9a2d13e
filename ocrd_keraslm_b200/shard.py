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
    rank computes the same partition without communicating.
    """
    lengths = np.asarray(lengths, dtype=np.int64)
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    loads = [0] * world_size
    parts = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (loads[k], k))
        parts[r].append(i)
        loads[r] += int(lengths[i]) + 1            # +1: an empty unit still costs a slot
    for p in parts:
        p.sort()
    return parts


def rate_sharded(rater, texts, contexts=None, gather=True, group=None):
    """Rate `texts` (independent units, each from a zero state) across the ranks of the default
    process group: rank r rates ``partition(...)[r]`` with ``rater.rate_many``.

    gather=True  : every rank returns the full list (one float32 array of probabilities per text)
    gather=False : every rank returns a list with its own units filled in and None elsewhere
    Without an initialised process group this is just ``rater.rate_many``.
    """
    import torch
    import torch.distributed as dist
    n = len(texts)
    if contexts is None:
        contexts = [None] * n
    if not (dist.is_available() and dist.is_initialized()):
        return rater.rate_many(texts, contexts)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lengths = [len(t) for t in texts]
    parts = partition(lengths, world)
    mine = parts[rank]
    res = rater.rate_many([texts[i] for i in mine], [contexts[i] for i in mine]) if mine else []
    out = [None] * n
    for i, r in zip(mine, res):
        out[i] = np.asarray(r, dtype=np.float32)
    if not gather or world == 1:
        return out
    # one padded all-gather of the flat score arrays; sizes are known to everybody
    sizes = [sum(lengths[i] for i in p) for p in parts]
    cap = max(max(sizes), 1)
    backend = dist.get_backend(group)
    dev = torch.device('cuda', torch.cuda.current_device()) if backend == 'nccl' else torch.device('cpu')
    flat = np.zeros(cap, dtype=np.float32)
    pos = 0
    for i in mine:
        flat[pos:pos + lengths[i]] = out[i]
        pos += lengths[i]
    send = torch.from_numpy(flat).to(dev)
    recv = [torch.empty(cap, dtype=torch.float32, device=dev) for _ in range(world)]
    dist.all_gather(recv, send, group=group)
    for r, p in enumerate(parts):
        if r == rank:
            continue
        buf = recv[r].cpu().numpy()
        pos = 0
        for i in p:
            out[i] = buf[pos:pos + lengths[i]].copy()
            pos += lengths[i]
    return out
