"""GPU parity at the sizes BASELINE.json quotes, against the reference build.

C1  GPS L1CA acquisition, ALL 32 PRNs x 71 bins x 16368 code phases, one 1 ms period
    (sdracquisition/pcorrelator/checkacquisition, src/sdracq.c:14-95, src/sdrcmn.c:738-773):
    acquired set, freqi, codei, nintg equal; peakr rel 1e-3, cn0 0.01 dB (float FFT vs the shim's
    double DFT).
C2  GPS L1CA tracking, 12 channels x 30 s (sdrtracking + correlator + pll/dll + sdrnavigation bit
    sync, src/sdrtrk.c:15-150, src/sdrnav.c:15-58,198-282): currnsamp, flagsync, synci, every decided
    nav bit and the frame-sync flags equal over all ~30 000 periods; the measured drift of the float
    class at 30 s is PRINTED next to the stated bound (|dII| <= 508 = two flipped samples per tap;
    carrfreq rel 1e-9) - SURVEY 8c asked for the figure.  The reference runs one process per channel.
C3  GLONASS G1 FDMA acquisition, all 14 frequency channels (k = -7..6) x 101 bins (+-10 kHz / 200 Hz)
    x 20000 code phases, I/Q int8 at 20 MS/s, 10 non-coherent integrations.  This fork cannot
    dispatch G1 (src/sdrcode.c:523-539), so the oracle is the port (pinned bit-for-bit to the
    reference on L1CA, its G1 code pinned to the reference's gencode_G1G2).
"""
import multiprocessing as mp
import os
import tempfile

import numpy as np
import pytest

import oracle_bind as ob
import util

pytestmark = pytest.mark.gpu

TRK_PRNS = [1, 3, 6, 9, 11, 14, 17, 19, 22, 25, 28, 31]
AVAIL10 = 3 * 65536
NSAMP = 16368


def _scene12(sdr, seed):
    """the bench's C2 scene: 12 PRNs at 55-58 dB-Hz within 30 Hz of a search bin."""
    rng = np.random.default_rng(seed)
    sats = []
    for p in TRK_PRNS:
        b = int(rng.integers(-20, 21))
        sats.append(dict(prn=p, cn0=float(rng.uniform(55, 58)), doppler=b * 200.0 + float(rng.uniform(-30, 30)),
                         code_phase=float(rng.uniform(0, 1023 * 20)), carr_phase=float(rng.uniform(0, 1)),
                         rate=float(rng.uniform(-1, 1))))
    return sdr.make_ifgen(util.FS1, util.IF1, util.CF, 1, sats, sigma=24.0, seed=seed)


def test_c1_all_32_prns_against_reference(sdr, oracle):
    cfg = _scene12(sdr, 0xB2000000)
    avail = 65536
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    host = st.read(0, avail)
    prns = list(range(1, 33))
    ref, _ = oracle.acquire(ob.cfg_iffile(acq_intg=1), host, avail, prns)
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    acq = sdr.Acquisition(util.FS1, 1, NSAMP, 16, 1e-3, freq, codes, 1.023e6, intg=1)
    got = acq.run(st, avail - 2 * NSAMP)
    for r, g in zip(ref, got):
        assert (r.acquired, r.freqi, r.codei, r.nintg_used) == (g["acquired"], g["freqi"], g["codei"], g["nintg"]), (r.prn, g)
        assert abs(r.peakr - g["peakr"]) <= 1e-3 * r.peakr and abs(r.cn0 - g["cn0"]) <= 0.01
    assert sorted(p for p, g in zip(prns, got) if g["acquired"]) == TRK_PRNS


def _ref_track_worker(job):
    path, total, prn, nper = job
    host = np.load(path, mmap_mode="r")[:total]
    o = ob.best()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)                                   # the reference printf()s per channel
    try:
        acq, rec, ntr = o.acq_track(ob.cfg_iffile(acq_intg=10), np.asarray(host), AVAIL10, [prn], nper)
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    n = ntr[0]
    r = rec[0][:n]
    keep = {k: r[k].copy() for k in ("currnsamp", "flagsync", "synci", "swsync", "bit", "flagsyncf", "carrfreq",
                                     "codefreq", "remcode", "remcarr", "L0", "flagobs")}
    keep["II"], keep["QQ"] = r["II"][:, :13].copy(), r["QQ"][:, :13].copy()
    return prn, (acq[0].acquired, acq[0].freqi, acq[0].codei, acq[0].acqfreq), n, keep, o.name


def test_c2_12_channels_30_s_against_reference(sdr):
    if not (ob.have_ref() or ob.have_port()):
        pytest.skip("no oracle library built")
    T = 30.0
    nper = int(T * 1000) - 8
    total = int(T * util.FS1) + 2 * 65536
    cfg = _scene12(sdr, 0xB2000000)
    st = sdr.Stream(1, total)
    st.push_ifgen(cfg, total)
    # the reference sees the very bytes the GPU tracks: the stream is read back once and shared
    # with the worker processes through a memory-mapped file
    tmp = tempfile.NamedTemporaryFile(suffix=".npy", delete=False, dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    tmp.close()
    try:
        np.save(tmp.name, st.read(0, total))
        with mp.get_context("spawn").Pool(min(len(TRK_PRNS), os.cpu_count() or 1)) as pool:
            refs = pool.map(_ref_track_worker, [(tmp.name, total, p, nper) for p in TRK_PRNS])
    finally:
        os.unlink(tmp.name)
    # GPU: the reference's hand-over (10 integrations, early exit), then 30 s of tracking
    tcodes = np.stack([sdr.gencode(p)[0] for p in TRK_PRNS])
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    acq10 = sdr.Acquisition(util.FS1, 1, NSAMP, 16, 1e-3, freq, tcodes, 1.023e6, intg=10)
    b10 = AVAIL10 - 11 * NSAMP
    got = acq10.run(st, b10)
    chans = []
    for (prn, a, n, keep, name), g in zip(refs, got):
        assert a[:3] == (g["acquired"], g["freqi"], g["codei"]) and a[3] == g["acqfreq"], (prn, a, g)
        assert g["acquired"] == 1
        chans.append(sdr.make_trk_chan(util.FS1, util.IF1, util.CF, 1, prn, sdr.gencode(prn)[0], b10 + g["codei"],
                                       g["acqfreq"], 6, 3, 6, (5.0, 1.0), (30.0, 10.0), (200.0, 50.0)))
    trk = sdr.Tracking(chans)
    rec, nrec = trk.run(st, nper)
    dII = dcf = dcode = drem = dL0 = 0.0
    nbits = 0
    for k, (prn, a, n, r, name) in enumerate(refs):
        assert nrec[k] == n == nper, (prn, nrec[k], n)
        g = rec[k]
        for key in ("currnsamp", "flagsync", "synci", "swsync", "flagsyncf"):
            assert np.array_equal(g[key], r[key]), (prn, key, int(np.argmax(g[key] != r[key])))
        sel = r["swsync"] == 1
        assert np.array_equal(g["bit"][sel], r["bit"][sel]), prn           # every decided nav bit
        nbits += int(sel.sum())
        assert g["flagsync"][-1] == 1                                       # the channel reached bit sync
        dII = max(dII, float(np.abs(g["II"][:, :13] - r["II"]).max()), float(np.abs(g["QQ"][:, :13] - r["QQ"]).max()))
        dcf = max(dcf, float(np.abs(g["carrfreq"] - r["carrfreq"]).max()))
        dcode = max(dcode, float(np.abs(g["codefreq"] - r["codefreq"]).max()))
        drem = max(drem, float(np.abs(g["remcode"] - r["remcode"]).max()), float(np.abs(g["remcarr"] - r["remcarr"]).max()))
        assert np.array_equal(g["flagobs"], r["flagobs"])                  # setobsdata scheduling
        dL0 = max(dL0, float(np.abs(g["L0"] - r["L0"]).max()))
        assert np.allclose(g["carrfreq"], r["carrfreq"], rtol=1e-9, atol=0) and np.allclose(g["codefreq"], r["codefreq"], rtol=1e-9, atol=0)
    assert dII <= 508.0 and drem <= 1e-6 and dL0 <= 2e-2
    print(f"\nC2 drift over {T:g} s x {len(refs)} channels vs {refs[0][4]}: max |dII,dQQ| = {dII:g} (bound 508), "
          f"max |dcarrfreq| = {dcf:.3g} Hz, max |dcodefreq| = {dcode:.3g} Hz, max |dremcode,dremcarr| = {drem:.3g}, max |dL0| = {dL0:.3g} cycles (bound 2e-2), "
          f"{nbits} nav bits equal")


def _port_g1_worker(job):
    path, avail, ks = job
    host = np.load(path, mmap_mode="r")
    o = ob.port()
    cfgo = ob.OhCfg(f_sf=20.0e6, f_if=0.0, f_cf=1.602e9, dtype=2, corrn=4, corrd=1, corrp=1, acq_hband=10000.0, acq_step=200.0)
    cfgo.dllb[:] = [5.0, 2.0]; cfgo.pllb[:] = [30.0, 20.0]; cfgo.fllb[:] = [200.0, 50.0]
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        ref, _ = o.acquire(cfgo, np.asarray(host), avail, list(ks), ctypes=[20] * len(ks))
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    return [(k, r.acquired, r.freqi, r.codei, r.nintg_used, r.peakr, r.cn0, r.acqfreq) for k, r in zip(ks, ref)]


def test_c3_glonass_14_channels_101_bins(sdr):
    if not ob.have_port():
        pytest.skip("oracle port not built")
    fs, nsamp, intg, nf = 20.0e6, 20000, 10, 101
    rng = np.random.default_rng(0xB2000033)
    present = [-7, -4, -1, 0, 2, 5, 6]                       # frequency numbers in view
    sats = [dict(prn=1, cn0=float(rng.uniform(44, 48)), doppler=k * 562500.0 + float(rng.uniform(-9000, 9000)),
                 code_phase=float(rng.uniform(0, 511 * 10)), carr_phase=float(rng.uniform(0, 1)), rate=0.0,
                 nav_on=0) for k in present]
    cfg = sdr.make_ifgen(fs, 0.0, 1.602e9, 2, sats, sigma=2.5, seed=0xB2000033, crate=0.511e6, clen=511, bitlen=10)
    need = (intg + 1) * nsamp
    avail = 65536 * (need // 65536 + 1)
    st = sdr.Stream(2, avail)
    st.push_ifgen(cfg, avail)
    ks = list(range(-7, 7))                                  # all 14 FDMA channels
    tmp = tempfile.NamedTemporaryFile(suffix=".npy", delete=False)
    tmp.close()
    try:
        np.save(tmp.name, st.read(0, avail))
        nproc = min(len(ks), os.cpu_count() or 1)
        with mp.get_context("spawn").Pool(nproc) as pool:
            parts = pool.map(_port_g1_worker, [(tmp.name, avail, ks[i::nproc]) for i in range(nproc)])
    finally:
        os.unlink(tmp.name)
    ref = {t[0]: t[1:] for part in parts for t in part}
    code, crate = sdr.gencode(1, 20)
    b0 = avail - need
    acquired = []
    for k in ks:
        freq = np.array([(i - nf // 2) * 200.0 + 562500.0 * k for i in range(nf)])        # src/sdrinit.c:633-635
        acq = sdr.Acquisition(fs, 2, nsamp, int(nsamp / 511), 1e-3, freq, code[None, :], crate, intg=intg)
        g = acq.run(st, b0)[0]
        r = ref[k]
        assert r[:4] == (g["acquired"], g["freqi"], g["codei"], g["nintg"]), (k, r, g)
        assert abs(r[4] - g["peakr"]) <= 1e-3 * r[4] and abs(r[5] - g["cn0"]) <= 0.01 and r[6] == g["acqfreq"]
        if g["acquired"]:
            acquired.append(k)
        acq.close()
    # which of the weak channels clear the threshold is the reference's call (equal above); none is invented
    assert set(acquired) <= set(present) and len(acquired) >= 4
    # the same 14 searches as ONE plan: the G1 code once per channel, every channel over its own grid
    # (freq [14][101], cfg.grid_per_prn) - one launch sequence per integration instead of fourteen
    grids = np.array([[(i - nf // 2) * 200.0 + 562500.0 * k for i in range(nf)] for k in ks])
    acq = sdr.Acquisition(fs, 2, nsamp, int(nsamp / 511), 1e-3, grids, np.repeat(code[None, :], len(ks), 0), crate, intg=intg)
    for k, g in zip(ks, acq.run(st, b0)):
        r = ref[k]
        assert r[:4] == (g["acquired"], g["freqi"], g["codei"], g["nintg"]), (k, r, g)
        assert abs(r[4] - g["peakr"]) <= 1e-3 * r[4] and abs(r[5] - g["cn0"]) <= 0.01 and r[6] == g["acqfreq"]
    acq.close()
