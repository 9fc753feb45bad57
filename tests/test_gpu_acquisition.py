"""GPU parity: batched acquisition vs the oracle's sdracquisition/pcorrelator/
checkacquisition (src/sdracq.c:14-95, src/sdrcmn.c:738-773).

Bit-exact: acquired PRN set, Doppler bin, integer code phase, integrations used.
Tolerance: power values / peak ratio / C/N0 - float32 FFTs of different
structure (the oracle uses the double-precision FFT shim): |dP| <= 2e-4*max(P),
peak ratio and C/N0 within 1e-3 relative / 0.01 dB."""
import numpy as np
import pytest

import oracle_bind as ob
import util

pytestmark = pytest.mark.gpu


def _grid(cfgo, f_if):
    hband, step = (7000.0, 200.0) if cfgo.acq_step <= 0 else (cfgo.acq_hband, cfgo.acq_step)
    nfreq = 2 * int(hband / step) + 1
    return np.array([f_if + (i - (nfreq - 1) // 2) * step for i in range(nfreq)])


def _check(sdr, oracle, cfg, fs, f_if, dtype, prns, cfgo, intg, power_prn=None):
    nsamp = int(fs * 1e-3)
    need = (intg + 1) * nsamp
    avail = 65536 * (need // 65536 + 1)
    st = sdr.Stream(dtype, avail)
    st.push_ifgen(cfg, avail)
    host = st.read(0, avail)
    ref, pw = oracle.acquire(cfgo, host, avail, prns, want_power=power_prn is not None)
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    freq = _grid(cfgo, f_if)
    acq = sdr.Acquisition(fs, dtype, nsamp, int(nsamp / 1023), 1e-3, freq, codes, 1.023e6, intg=intg,
                          keep_power=power_prn is not None)
    buffloc0 = avail - need                                     # src/sdracq.c:24-26
    got = acq.run(st, buffloc0)
    for r, g, p in zip(ref, got, prns):
        assert r.acquired == g["acquired"], (p, r.peakr, g["peakr"])
        assert (r.freqi, r.codei) == (g["freqi"], g["codei"]), (p, r.freqi, r.codei, g)
        assert r.nintg_used == g["nintg"], (p, r.nintg_used, g["nintg"])
        assert abs(r.peakr - g["peakr"]) <= 1e-3 * r.peakr, (p, r.peakr, g["peakr"])
        assert abs(r.cn0 - g["cn0"]) <= 0.01, (p, r.cn0, g["cn0"])
        assert abs(r.maxp - g["maxp"]) <= 2e-4 * r.maxp
        assert r.acqfreq == g["acqfreq"]
        if r.acquired:
            assert r.buffloc == buffloc0 + g["codei"]           # src/sdracq.c:53
    if power_prn is not None:
        i = prns.index(power_prn)
        P = acq.power(i)
        assert np.abs(P - pw[i]).max() <= 2e-4 * pw[i].max()
    # the host-buffer entry point gives the same answer (end-to-end path)
    got2 = acq.run_host(host[buffloc0 * dtype:], need)
    assert got2 == got
    return ref, got


def test_acq_real_if_single_period(sdr, oracle):
    """config 1 style: 16.368 MS/s real, 71 bins, intg forced to 1 (M = 32768 = 4 x 8192)."""
    cfg, vis = util.scene_c1(sdr, nsats=5)
    prns = vis + [2, 4, 32]
    ref, got = _check(sdr, oracle, cfg, util.FS1, util.IF1, 1, prns, ob.cfg_iffile(acq_intg=1), 1, power_prn=vis[1])
    assert [g["acquired"] for g in got[:5]] == [1] * 5
    assert sum(g["acquired"] for g in got[5:]) == 0


def test_acq_real_if_noncoherent_early_exit(sdr, oracle):
    """weaker signals: accumulation over several periods and data-dependent early exit."""
    cfg, vis = util.scene_c1(sdr, nsats=4, seed=0xB2000011)
    for i in range(cfg.nsat):
        cfg.sat[i].cn0_dbhz = 40.0 + 1.5 * i
    prns = vis + [8]
    ref, got = _check(sdr, oracle, cfg, util.FS1, util.IF1, 1, prns, ob.cfg_iffile(acq_intg=4), 4)
    assert len({g["nintg"] for g in got}) > 1                    # early exit exercised


def test_acq_iq_rtlsdr_style(sdr, oracle):
    """config 5 style: 2.048 MS/s I/Q, zero IF, default intg = 10 (M = 4096, fused path)."""
    cfg, vis = util.scene_c5(sdr, nsats=5)
    prns = vis + [1, 11, 21]
    _check(sdr, oracle, cfg, util.FS5, 0.0, 2, prns, ob.cfg_rtlsdr(), 10, power_prn=vis[0])


def test_acq_prn_slabs_equal_unsharded(sdr):
    """PRN slabs are independent - two half plans give the unsharded answer."""
    cfg, vis = util.scene_c1(sdr, nsats=4)
    prns = vis + [2, 4, 5, 7]
    nsamp, avail = 16368, 65536
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    b0 = avail - 2 * nsamp

    def run(sub):
        codes = np.stack([sdr.gencode(p)[0] for p in sub])
        return sdr.Acquisition(util.FS1, 1, nsamp, 16, 1e-3, freq, codes, 1.023e6, intg=1).run(st, b0)

    whole = run(prns)
    parts = sum((run(sdr.shard.prn_slab(prns, r, 2)) for r in range(2)), [])
    assert whole == parts
    assert [g["acquired"] for g in whole] == [1, 1, 1, 1, 0, 0, 0, 0]


def test_acq_coherent_blocks_scratch_path(sdr, oracle):
    """config 4 style: coherent blocks of several code periods (here 5 ms, M = 2^18, 32
    sub-transforms -> the scratch path k_row + k_comb + k_final) against the reference's
    pcorrelator run on a 5-period replica."""
    ncoh, nsamp, fs = 5, 16368, util.FS1
    n = ncoh * nsamp
    cfg, vis = util.scene_c1(sdr, nsats=3, seed=0xB2000044)
    avail = 2 * n + 4096
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    host = st.read(0, avail)
    b0 = avail - 2 * n
    dop = np.round(cfg.sat[0].doppler_hz / 100) * 100
    freq = np.array([util.IF1 + dop + (i - 4) * 100.0 for i in range(9)])
    prns = [vis[0], 4]
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    acq = sdr.Acquisition(fs, 1, nsamp, 16, 1e-3, freq, codes, 1.023e6, intg=1, ncoh=ncoh, keep_power=True)
    got = acq.run(st, b0)
    ci = 1.023e6 / fs
    for i, p in enumerate(prns):
        rep, _ = oracle.rescode(codes[i].astype(np.int16), 0.0, 0, ci, n)     # src/sdrinit.c:650
        x = np.zeros(2 * n, np.complex64)
        x[:n] = rep
        X = oracle.cpxfft(x)                                                  # src/sdrinit.c:652-653
        ref = oracle.pcorrelator(host[b0:b0 + 2 * n], 1, 1 / fs, n, freq, 1.023e6, 2 * n, X)
        P = acq.power(i).ravel()
        assert np.abs(P - ref).max() <= 2e-4 * ref.max(), p
        flat = int(np.argmax(ref))
        assert (got[i]["freqi"], got[i]["codei"]) == (flat // n, flat % n), p
        row, c0, w = ref.reshape(len(freq), n)[flat // n], flat % n, 2 * 16
        keep = np.ones(n, bool)
        keep[np.arange(c0 - w, c0 + w + 1) % n] = False                       # src/sdracq.c:81-82
        peakr = row[c0] / max(row[keep].max(), row[0])
        assert abs(got[i]["peakr"] - peakr) <= 1e-3 * peakr
        assert got[i]["acquired"] == int(peakr > 3.0)
    # a multi-period replica repeats its peak every code period, so the peak RATIO is ~1.2 by
    # construction; what must hold is the Doppler bin of the visible PRN
    assert got[0]["freqi"] == 4 and got[0]["maxp"] > 5 * got[1]["maxp"]


def test_acq_glonass_g1_fdma(sdr):
    """config 3 style: GLONASS G1 (511 chips, FDMA channels k * 562.5 kHz), I/Q at 20 MS/s,
    10 non-coherent integrations; M = 65536 = 8 x 8192 (cluster of 8 CTAs of 512 threads).
    This fork's gencode() has no CTYPE_G1 case (src/sdrcode.c:523-539), so the reference
    build cannot run the channel; the oracle here is the port, whose G1 code is pinned to
    the reference's static gencode_G1G2 (tests/test_oracle_port.py) and whose acquisition
    is pinned bit-for-bit to the reference on L1CA."""
    if not ob.have_port():
        pytest.skip("oracle port not built")
    oracle = ob.port()
    fs, nsamp = 20.0e6, 20000
    rng = np.random.default_rng(0xB2000003)
    chans = [-1, 0, 2]                                   # frequency numbers present
    sats = [dict(prn=1, cn0=float(rng.uniform(44, 47)), doppler=k * 562500.0 + float(rng.uniform(-3000, 3000)),
                 code_phase=float(rng.uniform(0, 511 * 10)), carr_phase=float(rng.uniform(0, 1)), rate=0.0,
                 nav_on=0) for k in chans]
    cfg = sdr.make_ifgen(fs, 0.0, 1.602e9, 2, sats, sigma=2.5, seed=0xB2000003, crate=0.511e6, clen=511, bitlen=10)
    intg = 10
    need = (intg + 1) * nsamp
    avail = 65536 * (need // 65536 + 1)                 # whole front-end blocks, src/sdrrcv.c:469-495
    st = sdr.Stream(2, avail)
    st.push_ifgen(cfg, avail)
    host = st.read(0, avail)
    cfgo = ob.OhCfg(f_sf=fs, f_if=0.0, f_cf=1.602e9, dtype=2, corrn=4, corrd=1, corrp=1)
    cfgo.dllb[:] = [5.0, 2.0]; cfgo.pllb[:] = [30.0, 20.0]; cfgo.fllb[:] = [200.0, 50.0]
    ks = chans + [4]                                     # one absent channel
    ref, _ = oracle.acquire(cfgo, host, avail, ks, ctypes=[20] * len(ks))
    code, crate = sdr.gencode(1, 20)
    assert len(code) == 511 and crate == 0.511e6
    b0 = avail - need
    for k, r in zip(ks, ref):
        freq = np.array([0.0 + (i - 35) * 200.0 + 562500.0 * k for i in range(71)])      # src/sdrinit.c:633-635
        acq = sdr.Acquisition(fs, 2, nsamp, int(nsamp / 511), 1e-3, freq, code[None, :], crate, intg=intg)
        g = acq.run(st, b0)[0]
        assert (r.acquired, r.freqi, r.codei, r.nintg_used) == (g["acquired"], g["freqi"], g["codei"], g["nintg"]), (k, g)
        assert abs(r.peakr - g["peakr"]) <= 1e-3 * r.peakr and abs(r.cn0 - g["cn0"]) <= 0.01
        assert r.acqfreq == g["acqfreq"]
    assert [r.acquired for r in ref] == [1, 1, 1, 0]


def test_acq_doppler_slabs_equal_unsharded(sdr):
    """SURVEY 8e, fewer PRNs than GPUs: the Doppler grid is cut into slabs, every slab plan runs all
    integration steps and keeps its per-step tuples, and shard.merge_doppler_slabs replays
    checkacquisition's argmax over the whole grid (src/sdracq.c:76-88) and sdracquisition's early exit
    (src/sdracq.c:29-43).  Must equal the unsharded search exactly (every field), for 2, 3 and 4
    slabs, with signals that are acquired after different numbers of integrations."""
    cfg, vis = util.scene_c1(sdr, nsats=4, seed=0xB2000011)
    for i in range(cfg.nsat):
        cfg.sat[i].cn0_dbhz = 40.0 + 1.5 * i
    prns = vis + [8]
    nsamp, intg = 16368, 4
    avail = 65536 * 2
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    b0 = avail - (intg + 1) * nsamp
    whole = sdr.Acquisition(util.FS1, 1, nsamp, 16, 1e-3, freq, codes, 1.023e6, intg=intg)
    want = whole.run(st, b0)
    assert len({w["nintg"] for w in want}) > 1                     # the early exit is exercised
    for world in (2, 3, 4):
        hists, offs = [], []
        for r in range(world):
            b = sdr.shard.doppler_slab(len(freq), r, world)
            plan = sdr.Acquisition(util.FS1, 1, nsamp, 16, 1e-3, freq[b.start:b.stop], codes, 1.023e6, intg=intg, threshold=1e30)
            plan.set_history(True)
            plan.enqueue(st, b0)
            hists.append(plan.fetch_history(intg))
            offs.append(b.start)
            plan.close()
        merged = sdr.shard.merge_doppler_slabs(hists, offs, nsamp, threshold=3.0)     # ACQTH
        for m, w in zip(merged, want):
            for k in ("acquired", "freqi", "codei", "nintg", "acqfreq", "peakr", "cn0", "maxp", "maxp2"):
                assert m[k] == w[k], (world, k, m, w)


def test_acq_cluster_exchange_repeatable(sdr):
    """The sub-transforms of a row cross the cluster through distributed shared memory behind
    a CTA barrier, one cluster-scope release fence and a relaxed cluster barrier.  A read that
    overtook its producer changes power values (seen when the fence was left out): 60 repeated
    searches of the same window must give the identical power matrix, bit for bit."""
    cfg, vis = util.scene_c1(sdr, nsats=4)
    prns = vis + [2, 4, 5, 7]
    nsamp, avail = 16368, 65536
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    acq = sdr.Acquisition(util.FS1, 1, nsamp, 16, 1e-3, freq, codes, 1.023e6, intg=1, keep_power=True)
    b0 = avail - 2 * nsamp
    first = acq.run(st, b0)
    P0 = [acq.power(i).copy() for i in range(len(prns))]
    plain = sdr.Acquisition(util.FS1, 1, nsamp, 16, 1e-3, freq, codes, 1.023e6, intg=1)
    for rep in range(120):
        assert acq.run(st, b0) == first
        assert plain.run(st, b0) == first                       # register-only variant (no P)
        if rep % 6 == 0:
            for i in range(len(prns)):
                assert np.array_equal(acq.power(i), P0[i]), (rep, i)


@pytest.fixture
def pow2_env(request, monkeypatch):
    """SDRB200_ACQ_POW2 (read once, at plan creation) keeps the transform length a power of two."""
    if request.param:
        monkeypatch.setenv("SDRB200_ACQ_POW2", "1")
    else:
        monkeypatch.delenv("SDRB200_ACQ_POW2", raising=False)
    return request.param


@pytest.mark.parametrize("pow2_env", [0, 1], indirect=True)
def test_acq_weak_signal_full_size_coherent(sdr, oracle, pow2_env):
    """config 4 at its real size: 20 ms coherent blocks (n = 327 360 lags; M = 655 360 = 160
    sub-transforms of 4096, radix 16 x 10 - or, with SDRB200_ACQ_POW2, 2^20 = 256 x 4096), 100 Hz Doppler step,
    2 non-coherent steps, a 30 dB-Hz GPS PRN, an SBAS PRN (133) and an absent PRN, no data
    transitions in the window.  Oracle: the reference's pcorrelator on a 20-period replica
    with its own accumulation flag, bins limited to 5 to keep the CPU side short."""
    ncoh, nsamp, fs = 20, 16368, util.FS1
    n = ncoh * nsamp
    sats = [dict(prn=9, cn0=31.0, doppler=1312.0, code_phase=311.4, carr_phase=0.3, rate=0.0, nav_on=0),
            dict(prn=133, cn0=30.0, doppler=-2093.0, code_phase=875.2, carr_phase=0.7, rate=0.0, nav_on=0)]
    cfg = sdr.make_ifgen(fs, util.IF1, util.CF, 1, sats, sigma=24.0, seed=0xB2000004)
    intg = 2
    avail = (intg + 1) * n + 4096
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    host = st.read(0, avail)
    b0 = avail - (intg + 1) * n
    prns = [9, 133, 4]
    ci = 1.023e6 / fs
    for i, p in enumerate(prns):
        dop = np.round((sats[i]["doppler"] if i < 2 else 0.0) / 100) * 100
        freq = np.array([util.IF1 + dop + (k - 2) * 100.0 for k in range(5)])
        code = sdr.gencode(p)[0]
        acq = sdr.Acquisition(fs, 1, nsamp, 16, 1e-3, freq, code[None, :], 1.023e6, intg=intg, ncoh=ncoh,
                              keep_power=True)
        g = acq.run(st, b0)[0]
        rep, _ = oracle.rescode(code.astype(np.int16), 0.0, 0, ci, n)
        x = np.zeros(2 * n, np.complex64)
        x[:n] = rep
        X = oracle.cpxfft(x)
        ref = oracle.pcorrelator(host[b0:b0 + 2 * n], 1, 1 / fs, n, freq, 1.023e6, 2 * n, X)
        if g["nintg"] == 2:                                   # no early exit: second block accumulated
            ref = oracle.pcorrelator(host[b0 + n:b0 + 3 * n], 1, 1 / fs, n, freq, 1.023e6, 2 * n, X, P=ref)
        P = acq.power(0).ravel()
        assert np.abs(P - ref).max() <= 2e-4 * ref.max(), p
        flat = int(np.argmax(ref))
        assert (g["freqi"], g["codei"]) == (flat // n, flat % n), (p, g)
        if i < 2:
            # the true Doppler bin, and a code phase equal to the injected delay modulo one period
            assert g["freqi"] == 2, (p, g)
            assert g["maxp"] > 8 * np.median(ref), p


@pytest.mark.parametrize("pow2_env", [0, 1], indirect=True)
@pytest.mark.parametrize("ncoh,dtype", [(10, 1), (40, 1), (40, 2), (6, 1), (12, 1), (24, 1)])
def test_acq_two_pass_path_all_radices(sdr, oracle, ncoh, dtype, pow2_env):
    """The two-pass path at its other radices.  Power-of-two lengths (SDRB200_ACQ_POW2): 10 ms real
    (M = 2^19, R = 128, G = 8), 40 ms real (M = 2^21, R = 256 x 8192, G = 16) and 40 ms I/Q at 2.048 MS/s
    (M = 2^18, R = 64, G = 4, the I/Q wipe-off of k_fwd_split).  Default lengths: M = 5 * 2^16 = 80 x 4096
    (G = 5), 5 * 2^18 = 160 x 8192 (G = 10) and, for the I/Q case, 3 * 2^16 = 48 x 4096 (G = 3; 5 * 2^15 has no radix);
    ncoh = 6, 12, 24: M = 3 * 2^16, 2^17, 2^18 = 48, 96, 192 x 4096 (G = 3, 6, 12; 2^18 .. 2^20 with POW2).
    Against the reference's pcorrelator on an ncoh-period replica."""
    fs, f_if, nsamp, scene = (util.FS1, util.IF1, 16368, util.scene_c1) if dtype == 1 else (util.FS5, 0.0, 2048, util.scene_c5)
    n = ncoh * nsamp
    cfg, vis = scene(sdr, nsats=2, seed=0xB2000046 + ncoh)
    for i in range(cfg.nsat):
        cfg.sat[i].nav_on = 0
    avail = 2 * n + 4096
    st = sdr.Stream(dtype, avail)
    st.push_ifgen(cfg, avail)
    host = st.read(0, avail)
    b0 = avail - 2 * n
    dop = np.round(cfg.sat[0].doppler_hz / 100) * 100
    freq = np.array([f_if + dop + (i - 1) * 100.0 for i in range(3)])
    code = sdr.gencode(vis[0])[0]
    acq = sdr.Acquisition(fs, dtype, nsamp, int(nsamp / 1023), 1e-3, freq, code[None, :], 1.023e6, intg=1, ncoh=ncoh,
                          keep_power=True)
    g = acq.run(st, b0)[0]
    rep, _ = oracle.rescode(code.astype(np.int16), 0.0, 0, 1.023e6 / fs, n)
    x = np.zeros(2 * n, np.complex64)
    x[:n] = rep
    ref = oracle.pcorrelator(host[b0 * dtype:(b0 + 2 * n) * dtype], dtype, 1 / fs, n, freq, 1.023e6, 2 * n, oracle.cpxfft(x))
    P = acq.power(0).ravel()
    assert np.abs(P - ref).max() <= 2e-4 * ref.max()
    flat = int(np.argmax(ref))
    assert (g["freqi"], g["codei"]) == (flat // n, flat % n)
    assert g["freqi"] == 1


@pytest.mark.parametrize("pow2_env", [0, 1], indirect=True)
@pytest.mark.parametrize("fs,dtype", [(10.0e6, 1), (20.0e6, 2), (20.0e6, 1), (5.0e6, 2), (2.5e6, 2),
                                      (24.0e6, 2), (12.0e6, 1), (6.0e6, 2), (3.0e6, 2)])
def test_acq_cluster_kernel_five_times_power_of_two(sdr, oracle, fs, dtype, pow2_env):
    """Windows of 2 500 .. 20 000 samples (1 ms at 2.5 .. 20 MS/s): the cluster kernel runs them as
    5 x 1024, 5 x 2048, 5 x 4096 = 20 480 and 10 x 4096 = 40 960 points, or 3 x 2048, 3 x 4096, 6 x 4096, 12 x 4096 (radix-5 / radix-10 recombination across the
    cluster, the last round of slice columns partial) instead of 2^15 / 2^16; with SDRB200_ACQ_POW2
    as 8 x 4096 / 16 x 4096.  Oracle: the reference's pcorrelator + argmax on the same window."""
    nsamp = int(fs * 1e-3)
    n = nsamp
    f_if = 0.0 if dtype == 2 else fs / 4
    sats = [dict(prn=7, cn0=47.0, doppler=1040.0, code_phase=400.3, carr_phase=0.2, rate=0.0, nav_on=0),
            dict(prn=19, cn0=45.0, doppler=-2030.0, code_phase=77.7, carr_phase=1.1, rate=0.0, nav_on=0)]
    cfg = sdr.make_ifgen(fs, f_if, util.CF, dtype, sats, sigma=24.0, seed=0xB2000050 + dtype)
    avail = 2 * n + 4096
    st = sdr.Stream(dtype, avail)
    st.push_ifgen(cfg, avail)
    host = st.read(0, avail)
    b0 = avail - 2 * n
    freq = np.array([f_if + (i - 5) * 500.0 for i in range(11)])
    prns = [7, 19, 3]
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    acq = sdr.Acquisition(fs, dtype, nsamp, int(nsamp / 1023), 1e-3, freq, codes, 1.023e6, intg=1, keep_power=True)
    got = acq.run(st, b0)
    plain = sdr.Acquisition(fs, dtype, nsamp, int(nsamp / 1023), 1e-3, freq, codes, 1.023e6, intg=1)
    assert plain.run(st, b0) == got                           # register-only search: same tuples
    for i, p in enumerate(prns):
        rep, _ = oracle.rescode(codes[i].astype(np.int16), 0.0, 0, 1.023e6 / fs, n)
        x = np.zeros(2 * n, np.complex64)
        x[:n] = rep
        ref = oracle.pcorrelator(host[b0 * dtype:(b0 + 2 * n) * dtype], dtype, 1 / fs, n, freq, 1.023e6, 2 * n, oracle.cpxfft(x))
        P = acq.power(i).ravel()
        assert np.abs(P - ref).max() <= 2e-4 * ref.max(), p
        flat = int(np.argmax(ref))
        assert (got[i]["freqi"], got[i]["codei"]) == (flat // n, flat % n), (p, got[i])
        row, c0, w = ref.reshape(len(freq), n)[flat // n], flat % n, 2 * int(nsamp / 1023)
        keep = np.ones(n, bool)
        keep[np.arange(c0 - w, c0 + w + 1) % n] = False                       # src/sdracq.c:81-82
        peakr = row[c0] / max(row[keep].max(), row[0])
        assert abs(got[i]["peakr"] - peakr) <= 1e-3 * peakr, (p, got[i])
        assert got[i]["acquired"] == int(peakr > 3.0), (p, got[i])
    assert got[0]["freqi"] == 7 and got[1]["freqi"] == 1 and got[0]["acquired"] == 1      # bins next to +1040 / -2030 Hz
