"""Multi-GPU host logic on CPU: world_size-2 gloo run of the slab partitioning and
the tuple all-gather that bench.py / a sharded acquisition use.  No
compute: the per-shard "results" are synthetic tuples keyed by PRN."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import sdrb200_pkg


def _fake(p):
    return dict(acquired=int(p % 3 == 0), freqi=p % 71, codei=(p * 977) % 16368, nintg=1,
                acqfreq=4.092e6 + 200.0 * (p % 71 - 35), peakr=1.0 + p / 7.0, cn0=30.0 + p,
                maxp=p * 1e-3, maxp2=p * 1e-4)


def _worker(rank, world, port, prns, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sdr = sdrb200_pkg.load()
    mine = sdr.shard.prn_slab(prns, rank, world)
    width = -(-len(prns) // world)
    merged = sdr.shard.gather_tuples(mine, [_fake(p) for p in mine], width)
    # tracking: channels by slab, no collective on the data path; only the counts are summed
    chans = list(sdr.shard.slab(12, rank, world))
    t = torch.tensor([float(len(chans))], dtype=torch.float64)
    dist.all_reduce(t)
    # max-over-ranks timing reduction as in bench.py
    tm = torch.tensor([0.5 + rank], dtype=torch.float64)
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    q.put((rank, mine, merged, chans, float(t[0]), float(tm[0])))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


@pytest.mark.parametrize("nprn", [32, 5])
def test_world2_slabs_and_gather(nprn):
    world = 2
    prns = list(range(1, nprn + 1))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, prns, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    slabs = [o[1] for o in outs]
    assert sum(slabs, []) == prns                                   # disjoint, ordered cover
    assert abs(len(slabs[0]) - len(slabs[1])) <= 1
    want = [dict(prn=p, **_fake(p)) for p in prns]
    for o in outs:
        assert o[2] == want                                         # every rank holds the merged table
        assert o[4] == 12.0 and o[5] == 1.5
    assert sorted(outs[0][3] + outs[1][3]) == list(range(12))


def test_slab_shapes():
    sdr = sdrb200_pkg.load()
    assert [len(sdr.shard.slab(12, r, 8)) for r in range(8)] == [2, 2, 2, 2, 1, 1, 1, 1]
    assert [len(sdr.shard.slab(3, r, 8)) for r in range(8)] == [1, 1, 1, 0, 0, 0, 0, 0]
    assert sum((list(sdr.shard.slab(64, r, 8)) for r in range(8)), []) == list(range(64))
    with pytest.raises(ValueError):
        sdr.shard.slab(4, 2, 2)


# ---------------------------------------------------------------- Doppler slabs
def _grid_case(seed, nprn=3, nfreq=11, nsamp=64, nstep=4):
    """synthetic accumulated power matrices per step and the tuples an unsharded search and every
    Doppler slab would report (argmax with first-index-wins, peak ratio against the second maximum
    of the winning row) - pure numpy, no GPU."""
    import numpy as np
    rng = np.random.default_rng(seed)
    P = np.cumsum(rng.random((nstep, nprn, nfreq, nsamp)).astype(np.float32), axis=0)
    P[:, 0, 7, 13] += np.arange(1, nstep + 1) ** 2 * 0.6       # PRN 0: a peak that outgrows the noise over the steps
    P[:, 1, 2, 5] = P[:, 1, 9, 5] = 50.0                      # PRN 1: an exact tie in two slabs
    return P


def _tuple_of(Prow_grid, f0, nsamp):
    """tuple of the best row of a (slab of a) grid [nf][nsamp]"""
    import numpy as np
    flat = int(np.argmax(Prow_grid))                          # first index wins
    fi, ci = divmod(flat, nsamp)
    row = Prow_grid[fi]
    maxp = float(row[ci])
    rest = np.delete(row, ci)
    maxp2 = float(rest.max())
    return dict(acquired=0, freqi=fi, codei=ci, nintg=0, acqfreq=1000.0 * (fi + f0), peakr=maxp / maxp2,
                cn0=10.0, maxp=maxp, maxp2=maxp2, meanp=float(rest.mean()))


def _unsharded(P, threshold):
    nstep, nprn, nfreq, nsamp = P.shape
    out = []
    for p in range(nprn):
        for s_ in range(nstep):
            t = _tuple_of(P[s_, p], 0, nsamp)
            t["nintg"] = s_ + 1
            t["acquired"] = int(t["peakr"] > threshold)
            if t["acquired"]:
                break
        out.append(t)
    return out


def test_merge_doppler_slabs_equals_unsharded():
    sdr = sdrb200_pkg.load()
    for seed in range(5):
        P = _grid_case(seed)
        nstep, nprn, nfreq, nsamp = P.shape
        for world in (1, 2, 3, 4):
            offs, hists = [], []
            for r in range(world):
                b = sdr.shard.doppler_slab(nfreq, r, world)
                offs.append(b.start)
                hists.append([[_tuple_of(P[s_, p, b.start:b.stop], b.start, nsamp) for p in range(nprn)] for s_ in range(nstep)])
            merged = sdr.shard.merge_doppler_slabs(hists, offs, nsamp, threshold=1.8)
            want = _unsharded(P, 1.8)
            for m, w in zip(merged, want):
                for k in ("acquired", "freqi", "codei", "nintg", "acqfreq", "peakr", "maxp", "maxp2"):
                    assert m[k] == w[k], (seed, world, k, m, w)
    assert any(w["acquired"] and w["nintg"] > 1 for w in _unsharded(_grid_case(0), 1.8))   # early exit exercised


def _worker_doppler(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sdr = sdrb200_pkg.load()
    P = _grid_case(3)
    nstep, nprn, nfreq, nsamp = P.shape
    b = sdr.shard.doppler_slab(nfreq, rank, world)
    mine = [[_tuple_of(P[s_, p, b.start:b.stop], b.start, nsamp) for p in range(nprn)] for s_ in range(nstep)]
    hists = sdr.shard.gather_histories(mine)                 # the one collective of the search
    offs = [sdr.shard.doppler_slab(nfreq, r, world).start for r in range(world)]
    merged = sdr.shard.merge_doppler_slabs(hists, offs, nsamp, threshold=1.8)
    q.put((rank, merged))
    dist.barrier()
    dist.destroy_process_group()


def test_world2_doppler_slabs():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_doppler, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = _unsharded(_grid_case(3), 1.8)
    for _, merged in outs:
        for m, w in zip(merged, want):
            for k in ("acquired", "freqi", "codei", "nintg", "acqfreq", "peakr", "maxp", "maxp2"):
                assert m[k] == w[k], (k, m, w)
