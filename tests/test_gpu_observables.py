"""GPU parity of SURVEY 8f-4: what sdrsync.c and sdrout.c consume, produced by the batched engine.

* the 80-deep interpolation stocks setobsdata shifts (src/sdrtrk.c:160-209: L, D, S, remcout, codei,
  cntout, codeisum) kept on the device and fetched newest-first, against the reference's own arrays
  after the same run: codei / cntout / codeisum equal, L / D / S / remcout within the observable
  tolerances of DESIGN.md (measured: <= 1e-10);
* syncthread's interpolation of carrier phase and Doppler at a reference sample (src/sdrsync.c:96-121,
  the reference's own uint64todouble + interp1) fed once from the GPU's stocks and once from the
  reference's: equal within 1e-9 cycles / 1e-9 Hz; pseudorange inputs (codei, remcout) as above;
* the per-period tracking log (writelog, src/sdrout.c:386-436) written from sdrb200_trk_rec_t against
  the reference's own log file of the same channel: identical text, line by line.
Needs the reference build of the oracle (the port has no sdrsync/sdrout)."""
import os
import tempfile

import numpy as np
import pytest

import oracle_bind as ob
import util

pytestmark = pytest.mark.gpu


def test_stocks_interpolation_and_log_against_reference(sdr):
    if not ob.have_ref():
        pytest.skip("reference build of the oracle not present")
    o = ob.ref()
    cfg, vis = util.scene_c5(sdr, nsats=3, seed=0xB2000081)
    prns = vis
    nper, ns = 6500, 2048
    avail = 65536
    total = avail + (nper + 8) * ns
    st = sdr.Stream(2, total)
    st.push_ifgen(cfg, total)
    host = st.read(0, total)
    cfgo = ob.cfg_rtlsdr()
    log = tempfile.NamedTemporaryFile(suffix=".csv", delete=False)
    log.close()
    try:
        o.set_logfile(log.name, 1)                                   # channel number 1 = prns[0]
        acq_o, rec_o, ntr_o = o.acq_track(cfgo, host, avail, prns, nper)
        o.set_logfile(None)
        ref_log = open(log.name).read()
    finally:
        os.unlink(log.name)
    assert all(a.acquired for a in acq_o) and min(ntr_o) == nper
    stocks_o = [o.last_stocks(i) for i in range(len(prns))]
    b0 = avail - 11 * ns
    chans = [sdr.make_trk_chan(util.FS5, 0.0, util.CF, 2, p, sdr.gencode(p)[0], a.buffloc, a.acqfreq,
                               4, 1, 1, (5.0, 2.0), (30.0, 20.0), (200.0, 50.0)) for p, a in zip(prns, acq_o)]
    trk = sdr.Tracking(chans)
    rec, nrec = trk.run(st, nper)
    assert nrec == [nper] * len(prns)
    for i in range(len(prns)):
        g, r = trk.fetch_obs(i), stocks_o[i]
        assert r["cntout"][0] > 0 and r["codei"][79] > 0             # the stocks are full: 80 observation epochs
        for k in ("codei", "cntout", "codeisum"):
            assert np.array_equal(g[k], r[k]), (i, k)
        assert np.abs(g["L"] - r["L"]).max() <= 2e-2 and np.abs(g["D"] - r["D"]).max() <= 5e-3
        assert np.abs(g["remcout"] - r["remcout"]).max() <= 2e-5 and np.abs(g["S"] - r["S"]).max() <= 1e-6
        # syncthread: reference sample between the stocks' ends, base as src/sdrsync.c:107-108
        sampbase = int(r["codei"][79]) - 10 * ns
        for frac in (0.1, 0.5, 0.93):
            sampref = int(r["codei"][79] + frac * (r["codei"][0] - r["codei"][79]))
            Lg, Dg = o.interp_obs(g, sampbase, float(sampref - sampbase))
            Lr, Dr = o.interp_obs(r, sampbase, float(sampref - sampbase))
            assert abs(Lg - Lr) <= 1e-9 * max(1.0, abs(Lr)) and abs(Dg - Dr) <= 1e-9 * max(1.0, abs(Dr)), (i, frac, Lg, Lr, Dg, Dr)
            # pseudorange term of src/sdrsync.c:113-114 from both stocks
            Pg = (float(int(g["codei"][0]) - sampref) - g["remcout"][0])
            Pr = (float(int(r["codei"][0]) - sampref) - r["remcout"][0])
            assert abs(Pg - Pr) <= 2e-5
    # the tracking log of channel 1, text against text
    got_log = sdr.trk_log(rec[0], 4, 1)
    gl, rl = got_log.splitlines(), ref_log.splitlines()
    assert len(gl) == len(rl) == nper + 1 and gl[0] == rl[0]
    diff = [k for k in range(1, len(rl)) if gl[k] != rl[k]]
    assert not diff, (len(diff), gl[diff[0]], rl[diff[0]])
