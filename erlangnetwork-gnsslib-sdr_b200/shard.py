"""Partitioning of the two hot paths over the GPUs of one box.

Nothing here touches samples: acquisition shards by PRN slab (whole PRNs per
rank, so the per-PRN argmax stays local), by Doppler slab (bins of the grid per
rank, when there are fewer PRNs than GPUs) or by IF stream, tracking by channel.
The only exchange is ONE all-gather of the per-shard result tuples
(prn, acquired, freqi, codei, nintg, acqfreq, peakr, cn0, maxp, maxp2) per search -
``torch.distributed`` over NCCL on the GPU box, gloo in the CPU tests.

Doppler slabs.  Every (PRN, bin) row of pcorrelator is independent
(src/sdrcmn.c:759-768); what couples the bins of one PRN is checkacquisition's
argmax over the whole grid (src/sdracq.c:76-88) and the early exit of
sdracquisition after every integration (src/sdracq.c:29-43).  A slab plan
therefore runs every step without the early exit and keeps the tuple of its
own best row per step; ``merge_doppler_slabs`` replays the reference's decision
over the gathered histories: per step the slab holding the global maximum
(first flat index wins, as maxvd does) supplies the statistics of the winning
row, and the search of a PRN ends at the first step whose peak ratio exceeds
the threshold.
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


def doppler_slab(nfreq: int, rank: int, world: int) -> range:
    """bins of the Doppler grid held by ``rank`` (contiguous, balanced)."""
    return slab(nfreq, rank, world)


def merge_doppler_slabs(histories: Sequence[Sequence[Sequence[dict]]], bin_offsets: Sequence[int],
                        nsamp: int, threshold: float) -> List[dict]:
    """histories[r][step][p]: tuple of slab r after integration step ``step`` for PRN index p
    (freqi local to the slab); bin_offsets[r]: first grid bin of slab r.  Returns the tuples
    the unsharded search would report: src/sdracq.c:29-43 (early exit), 76-88 (argmax)."""
    nslab = len(histories)
    nstep = len(histories[0])
    nprn = len(histories[0][0])
    out = []
    for p in range(nprn):
        res = None
        for step in range(nstep):
            best = None
            for r in range(nslab):
                t = histories[r][step][p]
                flat = (t["freqi"] + bin_offsets[r]) * nsamp + t["codei"]
                key = (t["maxp"], -flat)                      # larger power, then the first flat index
                if best is None or key > best[0]:
                    best = (key, r, t)
            _, r, t = best
            res = dict(t)
            res["freqi"] = t["freqi"] + bin_offsets[r]
            res["nintg"] = step + 1
            res["acquired"] = 1 if t["peakr"] > threshold else 0
            if res["acquired"]:
                break
        out.append(res)
    return out


HIST_FIELDS = ("acquired", "freqi", "codei", "nintg", "acqfreq", "peakr", "cn0", "maxp", "maxp2", "meanp")


def gather_histories(history: Sequence[Sequence[dict]], device=None):
    """all-gather the per-step tuples of this rank's Doppler slab: history[step][p] -> list over
    ranks of the same structure (ONE collective per search)."""
    import torch
    import torch.distributed as dist
    nstep, nprn = len(history), len(history[0])
    t = torch.tensor([[[float(history[s_][p].get(k, 0.0)) for k in HIST_FIELDS] for p in range(nprn)]
                      for s_ in range(nstep)], dtype=torch.float64)
    if device is not None:
        t = t.to(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        parts = [torch.empty_like(t) for _ in range(dist.get_world_size())]
        dist.all_gather(parts, t)
    else:
        parts = [t]
    out = []
    for part in parts:
        rows = part.cpu().tolist()
        hist = []
        for s_ in range(nstep):
            step = []
            for p in range(nprn):
                d = dict(zip(HIST_FIELDS, rows[s_][p]))
                for k in ("acquired", "freqi", "codei", "nintg"):
                    d[k] = int(d[k])
                step.append(d)
            hist.append(step)
        out.append(hist)
    return out
