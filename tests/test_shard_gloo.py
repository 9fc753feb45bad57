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
