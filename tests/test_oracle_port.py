"""CPU: pins the oracle.  (1) the CPU restatement oracle/port reproduces, bit for
bit, the golden vectors that tests/golden/make_golden.py recorded from the
REFERENCE build (the reference's own C sources + oracle FFT shim); (2) where the
reference build is present (this container) the two are compared directly on
fresh inputs.  Both use oracle/offt.c for the FFT, so equality is exact."""
import os

import numpy as np
import pytest

import oracle_bind as ob
import util

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REC_FIELDS = ["cnt", "buffloc", "currnsamp", "flagsync", "flagloopfilter", "swsync", "swloop", "synci",
              "biti", "bit", "II", "QQ", "sumI0", "sumQ0", "codefreq", "codeErr", "codeNco", "carrfreq",
              "carrErr", "carrNco", "freqErr", "remcode", "remcarr",
              "L0", "D0", "S0", "remcout0", "Isum", "obs_codei", "obs_cnt", "flagobs"]


@pytest.fixture(scope="module")
def port():
    if not ob.have_port():
        pytest.skip("oracle/liboracle_port.so not built")
    return ob.port()


def _acq_rows(res):
    return np.array([[r.acquired, r.freqi, r.codei, r.nintg_used, r.acqfreq, r.peakr, r.cn0, r.maxp, r.buffloc]
                     for r in res])


def test_port_primitives_golden(port):
    g = np.load(os.path.join(G, "primitives.npz"))
    for i, prm in enumerate(g["params"]):
        dtype, n = int(prm[0]), int(prm[1])
        ti, freq, phi0, crate, coff = prm[2:]
        d = g["d_real"][:n] if dtype == 1 else g["d_iq"][:2 * n]
        s = g[f"s{i}"]
        II, QQ, remc, remp = port.correlator(d, dtype, ti, n, freq, phi0, crate, coff, s, g["code7"])
        assert np.array_equal(II, g[f"II{i}"]) and np.array_equal(QQ, g[f"QQ{i}"])
        mI, mQ, mrem = port.mixcarr(d, dtype, ti, n, freq, phi0)
        assert np.array_equal(mI, g[f"mI{i}"]) and np.array_equal(mQ, g[f"mQ{i}"])
        rc, rrem = port.rescode(g["code7"], coff, int(s[-1]), ti * crate, n)
        assert np.array_equal(rc, g[f"rc{i}"])
        assert np.array_equal(np.array([remc, remp, mrem, rrem]), g[f"rem{i}"])


def test_port_codes_golden(port):
    g = np.load(os.path.join(G, "codes.npz"))
    bits = np.unpackbits(g["l1ca"], axis=1)[:, :1023]
    for prn in range(1, 211):
        code, crate = port.gencode(prn, 1)
        assert crate == 1.023e6 and np.array_equal(code > 0, bits[prn - 1].astype(bool)), prn
        assert set(np.unique(code)) == {-1, 1}
    code, crate = port.gencode(3, 20)                    # G1: 511-chip m-sequence, balance -1
    assert len(code) == 511 and crate == 0.511e6 and abs(int(code.sum())) == 1
    ac = np.array([np.dot(code, np.roll(code, k)) for k in range(1, 511)])
    assert np.all(ac == -1)                              # two-valued autocorrelation of an m-sequence
    # the reference's static gencode_G1G2 (src/sdrcode.c:426-444), reached via oracle/g1_wrap.c
    assert np.array_equal(code, g["g1"].astype(np.int16))


def test_port_fft_primitives_golden(port):
    g = np.load(os.path.join(G, "fftprims.npz"))
    for n in (12, 62, 1000):
        a, b = g[f"a{n}"], g[f"b{n}"]
        assert np.array_equal(port.cpxfft(a), g[f"fft{n}"])
        assert np.array_equal(port.cpxfft(a, inverse=True), g[f"ifft{n}"])
        assert np.array_equal(port.cpxconv(a, b, n // 2), g[f"conv{n}"])
        assert np.array_equal(port.cpxpspec(a), g[f"pspec{n}"])
        # against numpy's double-precision FFT: float rounding only
        assert np.abs(g[f"fft{n}"] - np.fft.fft(a.astype(np.complex128))).max() < 1e-5 * np.sqrt(n) * 4
    x, nsamp, nfft = port.codefft(3, 1, util.FS5)
    assert (nsamp, nfft) == (2048, 4096) and np.array_equal(x, g["codefft_prn3_fs2048"])


def test_port_acquisition_golden(port):
    g = np.load(os.path.join(G, "acq_c1.npz"))
    prns = [int(p) for p in g["prns"]]
    res, pw = port.acquire(ob.cfg_iffile(acq_intg=2), g["if_data"], 65536, prns, want_power=True)
    assert np.array_equal(_acq_rows(res), g["result"])
    rows = np.stack([pw[i, res[i].freqi] for i in range(len(prns))]).astype(np.float32)
    assert np.array_equal(rows, g["power_row"])
    assert [r.acquired for r in res] == [1, 1, 1, 1, 0, 0]
    assert {r.nintg_used for r in res} == {1, 2}         # early exit and full depth both covered


def test_port_tracking_golden(port):
    g = np.load(os.path.join(G, "trk_c5.npz"))
    prns = [int(p) for p in g["prns"]]
    nper = int(g["nper"])
    acq, rec, ntr = port.acq_track(ob.cfg_rtlsdr(acq_intg=2), g["if_data"], 65536, prns, nper)
    assert np.array_equal(_acq_rows(acq), g["acq"])
    assert np.array_equal(np.array(ntr), g["ntracked"]) and ntr[0] == nper
    for f in REC_FIELDS:
        assert np.array_equal(np.ascontiguousarray(rec[f]), g["rec_" + f]), f


def test_port_equals_reference_build(port):
    """fresh random inputs, reference build vs port, everything bit-exact."""
    if not ob.have_ref():
        pytest.skip("reference build absent (GPU box): covered by the golden vectors")
    ref = ob.ref()
    rng = np.random.default_rng(77)
    code = ref.gencode(19, 1)[0]
    assert np.array_equal(code, port.gencode(19, 1)[0])
    for dtype, n0, fs, s in ((1, 16368, util.FS1, [3, 6, 9, 12, 15, 18]), (2, 2048, util.FS5, [1, 2, 3, 4])):
        for _ in range(4):
            n = n0 + int(rng.integers(-2, 3))
            d = rng.integers(-128, 128, n * dtype, dtype=np.int8)
            a = (d, dtype, 1 / fs, n, float(rng.uniform(-5e3, 4.1e6)), float(rng.uniform(0, 6.28)),
                 1.023e6 + float(rng.uniform(-5, 5)), float(rng.uniform(0, 1023)), s, code)
            r, p = ref.correlator(*a), port.correlator(*a)
            assert all(np.array_equal(np.asarray(x), np.asarray(y)) for x, y in zip(r, p))
    # a short closed-loop run through acquisition, bit sync pre-phase and the loop filters
    import sdrb200_pkg
    import testkit
    sdr = testkit.attach(sdrb200_pkg.load())
    cfg, vis = util.scene_c5(sdr, nsats=2, seed=0xB2000099)
    nper = 4800                                   # past bit sync: setobsdata runs (flagobs)
    host = sdr.ifgen_host(cfg, 0, 65536 + (nper + 4) * 2048)
    a1, r1, n1 = ref.acq_track(ob.cfg_rtlsdr(acq_intg=1), host, 65536, vis, nper)
    a2, r2, n2 = port.acq_track(ob.cfg_rtlsdr(acq_intg=1), host, 65536, vis, nper)
    assert np.array_equal(_acq_rows(a1), _acq_rows(a2)) and n1 == n2
    for f in REC_FIELDS:
        assert np.array_equal(np.ascontiguousarray(r1[f]), np.ascontiguousarray(r2[f])), f
    assert int(r1["flagobs"].max()) == 2


def test_port_observables_golden(port):
    """setobsdata (src/sdrtrk.c:160-209) after bit sync: the port against the reference's
    records; the IF is regenerated from the seeded generator and pinned by its CRC."""
    import zlib
    import sdrb200_pkg
    import testkit
    sdr = testkit.attach(sdrb200_pkg.load())
    g = np.load(os.path.join(G, "obs_c5.npz"))
    prns, nper, k0 = [int(p) for p in g["prns"]], int(g["nper"]), int(g["k0"])
    cfg, vis = util.scene_c5(sdr, nsats=2, seed=0xB2000026)
    assert vis == prns
    host = sdr.ifgen_host(cfg, 0, 65536 + (nper + 4) * 2048)
    assert np.uint32(zlib.crc32(host.tobytes())) == g["if_crc"]
    acq, rec, ntr = port.acq_track(ob.cfg_rtlsdr(acq_intg=2), host, 65536, prns, nper)
    assert np.array_equal(np.array(ntr), g["ntracked"])
    for f in [k[4:] for k in g.keys() if k.startswith("rec_")]:
        assert np.array_equal(np.ascontiguousarray(rec[f][:, k0:]), g["rec_" + f]), f
    assert int(g["rec_flagobs"].max()) == 2


def test_port_ingest_golden(port):
    """front-end byte unpack (src/rcv/rtlsdr/rtlsdr.c:136-142, src/rcv/stereo/stereo.c:163-204,
    src/rcv/gn3s/gn3s.cpp:143-177): every byte value, and the reference's own routines"""
    g = np.load(os.path.join(G, "ingest.npz"))
    raw = np.arange(256, dtype=np.uint8)
    for f in ("rtlsdr", "stereo1", "stereo2", "gn3s2", "gn3s4"):
        assert np.array_equal(port.unpack(f, raw), g[f]), f
        if ob.have_ref():
            big = np.random.default_rng(3).integers(0, 256, 5000).astype(np.uint8)
            assert np.array_equal(port.unpack(f, big), ob.ref().unpack(f, big)), f


def test_port_frame_sync_equals_reference_build(port):
    """findpreamble + paritycheck + paritycheck_l1ca (src/sdrnav.c:324-411, src/sdrnav_gps.c:141-166)
    and the polarity term of setobsdata (src/sdrtrk.c:186-195): IF with LNAV subframes
    (tests/testkit/ifgen.h nav_on = 2), port against the reference build, bit for bit."""
    if not ob.have_ref():
        pytest.skip("reference build absent (GPU box): the GPU test compares with the port there")
    import sdrb200_pkg
    import testkit
    sdr = testkit.attach(sdrb200_pkg.load())
    ref = ob.ref()
    rng = np.random.default_rng(5)
    prns = [2, 5]
    sats = [dict(prn=p, cn0=48.0, doppler=float(b * 200.0 + rng.uniform(-60, 60)),
                 code_phase=float(rng.uniform(0, 1023 * 20)), carr_phase=float(rng.uniform(0, 1)), rate=0.3,
                 nav_on=2, nav_seed=100 + p) for p, b in zip(prns, (3, -7))]
    cfg = sdr.make_ifgen(util.FS5, 0.0, util.CF, 2, sats, sigma=20.0, seed=0xB2000071)
    nper = 12500
    host = sdr.ifgen_host(cfg, 0, 65536 + (nper + 4) * 2048)
    a1, r1, n1 = ref.acq_track(ob.cfg_rtlsdr(acq_intg=3), host, 65536, prns, nper)
    a2, r2, n2 = port.acq_track(ob.cfg_rtlsdr(acq_intg=3), host, 65536, prns, nper)
    assert n1 == n2 and int(r1["flagsyncf"].max()) == 1
    for f in REC_FIELDS + ["flagsyncf", "flagtow", "polarity"]:
        assert np.array_equal(np.ascontiguousarray(r1[f]), np.ascontiguousarray(r2[f])), f


def test_port_glonass_string_sync_against_reference(port):
    """GLONASS G1 string synchronisation (src/sdrnav.c:394-411 with the parameters of
    src/sdrinit.c:546-559): the port's restatement against the reference's own findpreamble() on
    the same 200-unit registers - the time mark at the end of the register (both polarities), one
    unit off, one unit wrong, and random registers."""
    if not ob.have_ref():
        pytest.skip("reference build not present")
    ref = ob.ref()
    pre = np.array([-1, -1, -1, -1, -1, 1, 1, 1, -1, -1, 1, -1, -1, -1, 1, -1, 1, -1, 1, 1,
                    1, 1, -1, 1, 1, -1, 1, -1, -1, 1], np.int32)
    rng = np.random.default_rng(0x61)
    cases = []
    for pol in (1, -1):
        reg = rng.choice(np.array([-1, 1], np.int32), 200)
        reg[170:] = pol * pre
        cases.append(reg.copy())
        cases.append(np.roll(reg, -1))                   # the mark one unit early
        bad = reg.copy(); bad[170 + int(rng.integers(0, 30))] *= -1
        cases.append(bad)                                # one unit of the mark wrong
    cases += [rng.choice(np.array([-1, 1], np.int32), 200) for _ in range(200)]
    cases.append(np.zeros(200, np.int32))                # the register before any bit has arrived
    found = 0
    for reg in cases:
        r, p = ref.framesync(20, 1, reg), port.framesync(20, 1, reg)
        assert r[0] == p[0], (r, p)
        if r[0]:
            assert r[1] == p[1]
            found += 1
    assert found == 2
