"""GPU parity: batched tracking engine vs the oracle's replay of
sdrtracking + cumsumcorr + pll + dll + sdrnavigation (src/sdrmain.c:264-315).

Integer decisions (sample counts, buffer positions, loop scheduling, bit-sync
index, nav bits) must be identical.  I/Q sums are exact integers/32 as long as
the NCO index sequences agree; the loop filters use CUDA libm (atan2/atan/sqrt)
instead of glibc, so NCO frequencies may differ in the last bits and a sample
index can flip: tolerance = 2 flipped samples per tap (2*2*127*32/32 = 508)."""
import numpy as np
import pytest

import oracle_bind as ob
import util

pytestmark = pytest.mark.gpu

IQ_TOL = 508.0
FREQ_RTOL = 1e-9


def _compare(sdr, cfgo, acq, rec_o, ntr_o, rec_g, ntr_g, ntap):
    for ch, a in enumerate(acq):
        n = min(ntr_o[ch], ntr_g[ch])
        assert n > 0
        o, g = rec_o[ch][:n], rec_g[ch][:n]
        for f in ("cnt", "buffloc", "currnsamp", "flagsync", "flagloopfilter", "swsync", "swloop",
                  "synci", "biti"):
            assert np.array_equal(o[f], g[f]), (a.prn, f, int(np.argmax(o[f] != g[f])))
        m = o["swsync"] == 1
        assert np.array_equal(o["bit"][m], g["bit"][m]), (a.prn, "nav bits")
        dI = np.abs(o["II"][:, :ntap] - g["II"][:, :ntap]).max()
        dQ = np.abs(o["QQ"][:, :ntap] - g["QQ"][:, :ntap]).max()
        assert dI <= IQ_TOL and dQ <= IQ_TOL, (a.prn, dI, dQ)
        for f in ("carrfreq", "codefreq"):
            assert np.allclose(o[f], g[f], rtol=FREQ_RTOL, atol=0), (a.prn, f)
        for f in ("remcode", "remcarr"):
            assert np.abs(o[f] - g[f]).max() < 1e-6, (a.prn, f)
        # observables (setobsdata, src/sdrtrk.c:160-209): scheduling and integer slots identical;
        # float slots follow carrfreq/codefreq (rel 1e-9 of 4 MHz = 4 mHz), L0 integrates D0
        for f in ("flagobs", "obs_codei", "obs_cnt"):
            assert np.array_equal(o[f], g[f]), (a.prn, f)
        assert o["flagobs"].max() == 2
        assert np.abs(o["D0"] - g["D0"]).max() <= 5e-3, (a.prn, "D0")
        assert np.abs(o["remcout0"] - g["remcout0"]).max() <= 2e-5, (a.prn, "remcout0")
        assert np.abs(o["L0"] - g["L0"]).max() <= 2e-2, (a.prn, "L0", np.abs(o["L0"] - g["L0"]).max())
        assert np.abs(o["S0"] - g["S0"]).max() <= 1e-6 and np.abs(o["Isum"] - g["Isum"]).max() <= IQ_TOL, (a.prn, "S0")
        print(f"prn {a.prn}: obs max|dD0|={np.abs(o['D0'] - g['D0']).max():.2e} Hz, max|dL0|="
              f"{np.abs(o['L0'] - g['L0']).max():.2e} cyc, max|dS0|={np.abs(o['S0'] - g['S0']).max():.2e}")
        exact = np.array_equal(o["II"][:, :ntap], g["II"][:, :ntap]) and np.array_equal(o["QQ"][:, :ntap], g["QQ"][:, :ntap])
        print(f"prn {a.prn}: {n} periods, I/Q bit-exact={exact}, max|dI|={dI}, "
              f"max|dcarrfreq|={np.abs(o['carrfreq'] - g['carrfreq']).max():.3e}")


def _run(sdr, oracle, cfg, prns, cfgo, nper, fs, dtype, ctype=1):
    nsamp = int(fs * 1e-3)
    avail = 65536 * ((3 * nsamp) // 65536 + 1)
    total = avail + nsamp * (nper + 8)
    st = sdr.Stream(dtype, total)
    st.push_ifgen(cfg, total)
    host = st.read(0, total)
    acq, rec_o, ntr_o = oracle.acq_track(cfgo, host, avail, prns, nper, ctypes=[ctype] * len(prns))
    assert all(a.acquired for a in acq)
    chans = [sdr.make_trk_chan(cfgo.f_sf, cfgo.f_if, cfgo.f_cf, dtype, a.prn, sdr.gencode(a.prn, ctype)[0],
                               a.buffloc, a.acqfreq, cfgo.corrn, cfgo.corrd, cfgo.corrp,
                               list(cfgo.dllb), list(cfgo.pllb), list(cfgo.fllb), ctype=ctype) for a in acq]
    eng = sdr.Tracking(chans)
    rec_g, ntr_g = eng.run(st, nper)
    _compare(sdr, cfgo, acq, rec_o, ntr_o, rec_g, ntr_g, 1 + 2 * cfgo.corrn)
    # a second launch continues where the first stopped; reset restores the hand-over state
    eng.reset()
    rec_a, na = eng.run(st, nper // 2)
    rec_b, nb = eng.run(st, nper - nper // 2)
    for ch in range(len(chans)):
        whole = rec_g[ch][:ntr_g[ch]]
        parts = np.concatenate([rec_a[ch][:na[ch]], rec_b[ch][:nb[ch]]])
        assert whole.tobytes() == parts[:len(whole)].tobytes()


def test_tracking_real_if(sdr, oracle):
    cfg, prns = util.scene_c1(sdr, nsats=3)
    _run(sdr, oracle, cfg, prns, ob.cfg_iffile(acq_intg=1), 4600, util.FS1, 1)


def test_tracking_iq_rtlsdr_style(sdr, oracle):
    cfg, prns = util.scene_c5(sdr, nsats=4)
    _run(sdr, oracle, cfg, prns, ob.cfg_rtlsdr(acq_intg=2), 4600, util.FS5, 2)


def test_tracking_sbas_l1(sdr, oracle):
    """SBAS L1 channels (CTYPE_L1SBAS: the C/A-family codes 120-158, 500 symbols/s = 2 code periods per
    symbol, loop interval 2 ms, src/sdr.h:154,165; bit sync takes the 'keyed on PRN > 5' branch of
    checksync, src/sdrnav.c:203-215): the same closed-loop parity as the L1CA tests, against the reference
    build.  Frame synchronisation of SBAS (Viterbi) is the host's."""
    rng = np.random.default_rng(0x5BA5)
    prns = [122, 133]
    sats = [dict(prn=p, cn0=47.0, doppler=float(rng.uniform(-2500, 2500)), code_phase=float(rng.uniform(0, 1023 * 2)),
                 carr_phase=float(rng.uniform(0, 1)), rate=0.2, nav_on=1, nav_seed=500 + p) for p in prns]
    cfg = sdr.make_ifgen(util.FS5, 0.0, util.CF, 2, sats, sigma=20.0, seed=0xB2005BA5, bitlen=2)
    _run(sdr, oracle, cfg, prns, ob.cfg_rtlsdr(acq_intg=3), 4000, util.FS5, 2, ctype=27)


def test_tracking_frame_sync_sbas(sdr):
    """SBAS L1 frame synchronisation on the device: the K = 7 Viterbi decoder of predecodefec() over the 1512-symbol
    register at every symbol, two consecutive preambles, CRC-24Q (src/sdrnav.c:300-321, 347-356, 383-411,
    lib/ka9q-fec/viterbi27_port.c).  IF carrying convolutionally encoded type-12 messages with valid CRC
    (tests/testkit/ifgen.h nav_on = 4), one satellite sent inverted.  Oracle: the reference build itself (this fork
    runs SBAS channels).  The symbol at which the preambles are accepted, the polarity, flagtow and the hand-over
    (firstsf, firstsfcnt) must be the reference's."""
    if not ob.have_ref():
        pytest.skip("reference build not present (the port has no SBAS frame sync)")
    oracle = ob.ref()
    rng = np.random.default_rng(0x5BA6)
    prns = [122, 133, 138]
    sats = [dict(prn=p, cn0=49.0, doppler=float(rng.uniform(-2500, 2500)), code_phase=float(rng.uniform(0, 1023 * 2)),
                 carr_phase=float(rng.uniform(0, 1)), rate=0.2, nav_on=4, nav_seed=(900 + p) | ((i & 1) << 16))
            for i, p in enumerate(prns)]
    cfg = sdr.make_ifgen(util.FS5, 0.0, util.CF, 2, sats, sigma=20.0, seed=0xB2005BA6, bitlen=2)
    nper = 10500
    avail = 65536
    total = avail + 2048 * (nper + 8)
    st = sdr.Stream(2, total)
    st.push_ifgen(cfg, total)
    host = st.read(0, total)
    cfgo = ob.cfg_rtlsdr(acq_intg=3)
    acq, rec_o, ntr_o = oracle.acq_track(cfgo, host, avail, prns, nper, ctypes=[27] * len(prns))
    assert all(a.acquired for a in acq)
    chans = [sdr.make_trk_chan(cfgo.f_sf, cfgo.f_if, cfgo.f_cf, 2, a.prn, sdr.gencode(a.prn, 27)[0],
                               a.buffloc, a.acqfreq, cfgo.corrn, cfgo.corrd, cfgo.corrp,
                               list(cfgo.dllb), list(cfgo.pllb), list(cfgo.fllb), ctype=27) for a in acq]
    rec_g, ntr_g = sdr.Tracking(chans).run(st, nper)
    synced = 0
    for ch, a in enumerate(acq):
        assert ntr_o[ch] == nper and ntr_g[ch] == nper
        o, g = rec_o[ch], rec_g[ch]
        for f in ("currnsamp", "flagsync", "synci", "flagsyncf", "flagtow", "polarity"):
            assert np.array_equal(o[f], g[f]), (a.prn, f, int(np.argmax(o[f] != g[f])))
        if o["flagsyncf"].max() == 1:                       # the reference's 2-period bit sync can lock across symbols
            synced += 1
            k = int(np.argmax(o["flagsyncf"] > 0))
            assert int(g["firstsfcnt"][-1]) == int(o["cnt"][k]) and int(g["firstsf"][-1]) == int(o["buffloc"][k])
            print(f"prn {a.prn}: frame sync at period {k}, polarity {int(g['polarity'][-1])}")
    assert synced >= 2, "the reference did not frame-sync on at least two of the three satellites"


def test_tracking_frame_sync_lnav(sdr, oracle):
    """Frame synchronisation on the device (findpreamble + paritycheck for GPS L1CA,
    src/sdrnav.c:43-58, 324-411; paritycheck_l1ca, src/sdrnav_gps.c:141-166): IF carrying LNAV
    subframes (preamble, TOW, valid parity; csrc/ifgen.h nav_on = 2).  The period at which the
    preamble is accepted, the polarity it fixes and the +0.5 cycle term it adds to the carrier
    phase observable (src/sdrtrk.c:186-195) must be the reference's."""
    rng = np.random.default_rng(5)
    prns = [2, 5, 7]
    sats = [dict(prn=p, cn0=48.0, doppler=float(b * 200.0 + rng.uniform(-60, 60)),
                 code_phase=float(rng.uniform(0, 1023 * 20)), carr_phase=float(rng.uniform(0, 1)), rate=0.3,
                 nav_on=2, nav_seed=100 + p) for p, b in zip(prns, (3, -7, 11))]
    cfg = sdr.make_ifgen(util.FS5, 0.0, util.CF, 2, sats, sigma=20.0, seed=0xB2000071)
    nper = 13000
    avail = 65536
    total = avail + 2048 * (nper + 8)
    st = sdr.Stream(2, total)
    st.push_ifgen(cfg, total)
    host = st.read(0, total)
    cfgo = ob.cfg_rtlsdr(acq_intg=3)
    acq, rec_o, ntr_o = oracle.acq_track(cfgo, host, avail, prns, nper)
    assert all(a.acquired for a in acq)
    chans = [sdr.make_trk_chan(cfgo.f_sf, cfgo.f_if, cfgo.f_cf, 2, a.prn, sdr.gencode(a.prn)[0],
                               a.buffloc, a.acqfreq, cfgo.corrn, cfgo.corrd, cfgo.corrp,
                               list(cfgo.dllb), list(cfgo.pllb), list(cfgo.fllb)) for a in acq]
    rec_g, ntr_g = sdr.Tracking(chans).run(st, nper)
    for ch, a in enumerate(acq):
        n = min(ntr_o[ch], ntr_g[ch])
        assert n == nper
        o, g = rec_o[ch][:n], rec_g[ch][:n]
        assert o["flagsyncf"].max() == 1, (a.prn, "the reference did not frame-sync")
        for f in ("flagsync", "synci", "flagsyncf", "flagtow", "polarity", "flagobs"):
            assert np.array_equal(o[f], g[f]), (a.prn, f, int(np.argmax(o[f] != g[f])))
        k = int(np.argmax(o["flagsyncf"] > 0))
        assert int(g["firstsfcnt"][-1]) == int(o["cnt"][k]) and int(g["firstsf"][-1]) == int(o["buffloc"][k])
        assert np.abs(o["L0"] - g["L0"]).max() <= 2e-2, (a.prn, "L0")
    assert sorted(int(r["polarity"][-1]) for r in rec_g) == [-1, -1, 1]      # the +0.5 term is exercised


def test_tracking_string_sync_glonass_g1(sdr):
    """GLONASS G1 string synchronisation on the device (findpreamble for CTYPE_G1, src/sdrnav.c:394-411:
    the 30-unit time mark at the end of the 200-unit register, 10 code periods per unit, loop interval
    10 ms): IF carrying strings with the time mark (tests/testkit/ifgen.h nav_on = 3), one satellite sent
    inverted.  This fork cannot dispatch a G1 channel (src/sdrcode.c:523-539), so the oracle is the port,
    whose string sync is pinned to the reference's findpreamble() in tests/test_oracle_port.py and whose
    tracking loop is pinned to the reference on L1CA.  The period at which the mark is accepted, the
    polarity and the hand-over (firstsf, firstsfcnt) must be equal."""
    if not ob.have_port():
        pytest.skip("oracle port not built")
    oracle = ob.port()
    fs, nsamp = 5.11e6, 5110                                   # 10 samples per chip, I/Q
    rng = np.random.default_rng(0xB20000C1)
    ks = [-1, 2]                                               # frequency numbers
    sats = [dict(prn=1, cn0=49.0, doppler=float(rng.uniform(-2500, 2500)), carr_offset=k * 562500.0,
                 code_phase=float(rng.uniform(0, 511 * 10)), carr_phase=float(rng.uniform(0, 1)), rate=0.2,
                 nav_on=3, nav_seed=(7 + i) | (i << 16)) for i, k in enumerate(ks)]
    cfg = sdr.make_ifgen(fs, 0.0, 1.602e9, 2, sats, sigma=3.0, seed=0xB20000C1, crate=0.511e6, clen=511, bitlen=10)
    nper = 8200
    avail = 65536 * 2
    total = avail + nsamp * (nper + 8)
    st = sdr.Stream(2, total)
    st.push_ifgen(cfg, total)
    host = st.read(0, total)
    cfgo = ob.OhCfg(f_sf=fs, f_if=0.0, f_cf=1.602e9, dtype=2, corrn=4, corrd=1, corrp=1, acq_intg=4)
    cfgo.dllb[:] = [5.0, 2.0]; cfgo.pllb[:] = [30.0, 20.0]; cfgo.fllb[:] = [200.0, 50.0]
    acq, rec_o, ntr_o = oracle.acq_track(cfgo, host, avail, ks, nper, ctypes=[20] * len(ks))
    assert all(a.acquired for a in acq)
    code = sdr.gencode(1, 20)[0]
    chans = [sdr.make_trk_chan(fs, 0.0, 1.602e9 + 562500.0 * k, 2, k, code, a.buffloc, a.acqfreq, cfgo.corrn, cfgo.corrd,
                               cfgo.corrp, list(cfgo.dllb), list(cfgo.pllb), list(cfgo.fllb), ctype=20,
                               foffset=562500.0 * k) for k, a in zip(ks, acq)]
    rec_g, ntr_g = sdr.Tracking(chans).run(st, nper)
    pols = []
    for ch, k in enumerate(ks):
        assert ntr_o[ch] == nper and ntr_g[ch] == nper
        o, g = rec_o[ch], rec_g[ch]
        assert o["flagsyncf"].max() == 1, (k, "the oracle did not find the time mark")
        for f in ("currnsamp", "flagsync", "synci", "flagsyncf", "flagtow", "polarity"):
            assert np.array_equal(o[f], g[f]), (k, f, int(np.argmax(o[f] != g[f])))
        i0 = int(np.argmax(o["flagsyncf"] > 0))
        assert int(g["firstsfcnt"][-1]) == int(o["cnt"][i0]) and int(g["firstsf"][-1]) == int(o["buffloc"][i0])
        pols.append(int(g["polarity"][-1]))
    assert all(p in (-1, 1) for p in pols)                      # which sign the PLL locked to is the oracle's call (equal above)
