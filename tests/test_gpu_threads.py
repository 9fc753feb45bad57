"""GPU: the sdr.h-level entry points are called from one pthread per channel in the
reference (src/sdrmain.c:255-280, up to 32 threads, bin/gnss-sdrcli.ini:5).  The engine
keeps per-thread workspaces and streams; concurrent callers must get exactly the answers
of sequential callers (SURVEY 8b "Threading")."""
import threading

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

NTHREADS = 8


def test_concurrent_correlator_and_pcorrelator(sdr, oracle):
    rng = np.random.default_rng(2024)
    cfg, prns = util.scene_c1(sdr, nsats=3)
    data = sdr.ifgen_host(cfg, 0, 2 * 16368 + 64)
    jobs = []
    for t in range(NTHREADS):
        code = util.l1ca_code(sdr, prns[t % len(prns)])
        args = []
        for _ in range(12):
            n = 16368 + int(rng.integers(-1, 2))
            off = int(rng.integers(0, 32))
            args.append((data[off:off + n], 1, 1 / util.FS1, n, util.IF1 + float(rng.uniform(-5e3, 5e3)),
                         float(rng.uniform(0, 6.28)), 1.023e6 + float(rng.uniform(-4, 4)),
                         float(rng.uniform(0, 1023)), [3, 6, 9, 12, 15, 18], code))
        jobs.append(args)
    # acquisition-side primitive with a per-thread Doppler grid (m = 2n = 4096, I/Q)
    cfg5, vis5 = util.scene_c5(sdr, nsats=3)
    d5 = sdr.ifgen_host(cfg5, 0, 4096)
    x5, _, _ = oracle.codefft(vis5[0], 1, util.FS5)
    grids = [np.array([(i - 5) * 200.0 + 400.0 * t for i in range(11)]) for t in range(NTHREADS)]

    seq_c = [[sdr.correlator(*a) for a in args] for args in jobs]
    seq_p = [sdr.pcorrelator(d5, 2, 1 / util.FS5, 2048, g, 1.023e6, 4096, x5) for g in grids]

    out_c, out_p, errs = [None] * NTHREADS, [None] * NTHREADS, []

    def work(t):
        try:
            res = []
            for rep in range(3):
                res = [sdr.correlator(*a) for a in jobs[t]]
                out_p[t] = sdr.pcorrelator(d5, 2, 1 / util.FS5, 2048, grids[t], 1.023e6, 4096, x5)
            out_c[t] = res
        except Exception as e:                                  # noqa: BLE001
            errs.append((t, repr(e)))

    th = [threading.Thread(target=work, args=(t,)) for t in range(NTHREADS)]
    [x.start() for x in th]
    [x.join() for x in th]
    assert not errs, errs
    for t in range(NTHREADS):
        for r, g in zip(seq_c[t], out_c[t]):
            assert np.array_equal(r[0], g[0]) and np.array_equal(r[1], g[1]) and r[2] == g[2] and r[3] == g[3]
        assert np.array_equal(seq_p[t], out_p[t]), t
    # and the sequential answers are the oracle's
    r = oracle.correlator(*jobs[0][0])
    assert np.array_equal(r[0], seq_c[0][0][0]) and r[3] == seq_c[0][0][3]


def test_concurrent_batched_engines(sdr):
    """several acquisition plans and tracking engines driven from different host threads
    (config 5 style: independent IF streams) give the single-thread answers."""
    streams, plans, ref = [], [], []
    freq = np.array([(i - 35) * 200.0 for i in range(71)])
    for t in range(4):
        cfg, vis = util.scene_c5(sdr, nsats=4, seed=0xB2005000 + t)
        total = 65536 + 300 * 2048
        st = sdr.Stream(2, total)
        st.push_ifgen(cfg, total)
        codes = np.stack([sdr.gencode(p)[0] for p in vis])
        acq = sdr.Acquisition(util.FS5, 2, 2048, 2, 1e-3, freq, codes, 1.023e6, intg=2)
        streams.append(st); plans.append((acq, vis))

    def chain(t):
        acq, vis = plans[t]
        b0 = 65536 - 3 * 2048
        got = acq.run(streams[t], b0)
        chans = [sdr.make_trk_chan(util.FS5, 0.0, util.CF, 2, p, sdr.gencode(p)[0], b0 + g["codei"], g["acqfreq"],
                                   4, 1, 1, (5.0, 2.0), (30.0, 20.0), (200.0, 50.0))
                 for p, g in zip(vis, got) if g["acquired"]]
        rec, nrec = sdr.Tracking(chans).run(streams[t], 250)
        return got, rec.tobytes(), list(nrec)

    ref = [chain(t) for t in range(4)]
    out, errs = [None] * 4, []

    def work(t):
        try:
            out[t] = chain(t)
        except Exception as e:                                  # noqa: BLE001
            errs.append((t, repr(e)))

    th = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    [x.start() for x in th]
    [x.join() for x in th]
    assert not errs, errs
    for t in range(4):
        assert out[t][0] == ref[t][0] and out[t][2] == ref[t][2] and out[t][1] == ref[t][1], t
        assert sum(g["acquired"] for g in ref[t][0]) >= 3
