"""Partitioning of the two hot paths over the GPUs of one box.

Nothing here touches samples: acquisition shards by PRN slab (whole PRNs per
rank, so the per-PRN argmax stays local) or by IF stream, tracking by channel.
The only exchange is an all-gather of the per-shard result tuples
(prn, acquired, freqi, codei, nintg, acqfreq, peakr, cn0, maxp, maxp2) -
``torch.distributed`` over NCCL on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import List, Sequence

TUPLE_FIELDS = ("prn", "acquired", "freqi", "codei", "nintg", "acqfreq", "peakr", "cn0", "maxp", "maxp2")


def slab(n: int, rank: int, world: int) -> range:
    """Contiguous balanced slab of ``n`` units for ``rank``: the first n % world
    ranks hold one more unit (12 channels over 8 ranks -> 2,2,2,2,1,1,1,1)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


def prn_slab(prns: Sequence[int], rank: int, world: int) -> List[int]:
    return [prns[i] for i in slab(len(prns), rank, world)]


def pack_tuples(prns: Sequence[int], results: Sequence[dict], width: int):
    """results of one shard as a fixed-width float64 tensor (rows beyond the shard
    are marked with prn = -1 so that every rank contributes the same shape)."""
    import torch
    t = torch.full((width, len(TUPLE_FIELDS)), -1.0, dtype=torch.float64)
    for i, (p, r) in enumerate(zip(prns, results)):
        t[i] = torch.tensor([float(p)] + [float(r[k]) for k in TUPLE_FIELDS[1:]], dtype=torch.float64)
    return t


def gather_tuples(prns: Sequence[int], results: Sequence[dict], width: int, device=None):
    """all-gather the shard's tuples; returns the merged list of dicts ordered by rank
    then slab position (= the order of an unsharded run over the same PRN list)."""
    import torch
    import torch.distributed as dist
    t = pack_tuples(prns, results, width)
    if device is not None:
        t = t.to(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        parts = [torch.empty_like(t) for _ in range(dist.get_world_size())]
        dist.all_gather(parts, t)
    else:
        parts = [t]
    out = []
    for part in parts:
        for row in part.cpu().tolist():
            if row[0] < 0:
                continue
            d = dict(zip(TUPLE_FIELDS, row))
            for k in ("prn", "acquired", "freqi", "codei", "nintg"):
                d[k] = int(d[k])
            out.append(d)
    return out
