"""Generate the golden vectors under tests/golden/ from the REFERENCE build of the
oracle (oracle/_ref/libreplay_ref.so = the reference's own C sources compiled
where they lie + oracle FFT shim).  Run in the container that has /root/reference:

    python tests/golden/make_golden.py

The inputs are stored next to the outputs, so the fixtures do not depend on any
libm of the machine that replays them."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_bind as ob  # noqa: E402
import sdrb200_pkg  # noqa: E402
import util  # noqa: E402

OBS_FIELDS = ["cnt", "buffloc", "flagsync", "flagloopfilter", "sumI0", "carrfreq", "codefreq", "remcarr",
              "L0", "D0", "S0", "remcout0", "Isum", "obs_codei", "obs_cnt", "flagobs"]
REC_FIELDS = ["cnt", "buffloc", "currnsamp", "flagsync", "flagloopfilter", "swsync", "swloop", "synci",
              "biti", "bit", "II", "QQ", "sumI0", "sumQ0", "codefreq", "codeErr", "codeNco", "carrfreq",
              "carrErr", "carrNco", "freqErr", "remcode", "remcarr",
              "L0", "D0", "S0", "remcout0", "Isum", "obs_codei", "obs_cnt", "flagobs"]


def main():
    import testkit
    sdr = testkit.attach(sdrb200_pkg.load())
    o = ob.ref()
    assert o.name == "reference"
    rng = np.random.default_rng(0xB200)

    # ---- primitives: correlator / mixcarr / rescode on random int8 IF -------------
    prim = {}
    code7 = o.gencode(7, 1)[0]
    d_real = rng.integers(-128, 128, 16500, dtype=np.int8)
    d_iq = rng.integers(-128, 128, 2 * 2100, dtype=np.int8)
    cases = []
    for i in range(6):
        real = i < 3
        n = (16368 if real else 2048) + int(rng.integers(-1, 2))
        c = dict(dtype=1 if real else 2, n=n, ti=1.0 / (util.FS1 if real else util.FS5),
                 freq=(util.IF1 if real else 0.0) + float(rng.uniform(-5000, 5000)),
                 phi0=float(rng.uniform(0, 2 * np.pi)), crate=1.023e6 + float(rng.uniform(-4, 4)),
                 coff=float(rng.uniform(0, 1023)), s=[3, 6, 9, 12, 15, 18] if real else [1, 2, 3, 4])
        d = d_real[:n] if real else d_iq[:2 * n]
        II, QQ, remc, remp = o.correlator(d, c["dtype"], c["ti"], n, c["freq"], c["phi0"], c["crate"], c["coff"], c["s"], code7)
        mI, mQ, mrem = o.mixcarr(d, c["dtype"], c["ti"], n, c["freq"], c["phi0"])
        rc, rrem = o.rescode(code7, c["coff"], c["s"][-1], c["ti"] * c["crate"], n)
        cases.append((c, II, QQ, remc, remp, mI, mQ, mrem, rc, rrem))
    prim["d_real"], prim["d_iq"], prim["code7"] = d_real, d_iq, code7
    prim["params"] = np.array([[c["dtype"], c["n"], c["ti"], c["freq"], c["phi0"], c["crate"], c["coff"]] for c, *_ in cases])
    for i, (c, II, QQ, remc, remp, mI, mQ, mrem, rc, rrem) in enumerate(cases):
        prim[f"s{i}"] = np.array(c["s"], np.int32)
        prim[f"II{i}"], prim[f"QQ{i}"] = II, QQ
        prim[f"rem{i}"] = np.array([remc, remp, mrem, rrem])
        prim[f"mI{i}"], prim[f"mQ{i}"], prim[f"rc{i}"] = mI, mQ, rc.astype(np.int8)
    np.savez_compressed(os.path.join(HERE, "primitives.npz"), **prim)

    # ---- codes: every L1CA/SBAS PRN and the G1 code, packed to bits ---------------
    codes = np.stack([o.gencode(p, 1)[0] for p in range(1, 211)])
    # G1: gencode() has no CTYPE_G1 case in this fork (src/sdrcode.c:523-539); the static
    # generator gencode_G1G2 (src/sdrcode.c:426-444) is reached through oracle/g1_wrap.c
    import ctypes as C
    g1 = np.zeros(511, np.int16)
    cr = C.c_double()
    assert o.lib.oh_gencode_g1_static(g1.ctypes.data_as(C.POINTER(C.c_short)), C.byref(cr)) == 511 and cr.value == 0.511e6
    np.savez_compressed(os.path.join(HERE, "codes.npz"), l1ca=np.packbits(codes > 0, axis=1), g1=g1.astype(np.int8))

    # ---- front-end byte unpack: all 256 byte values through the reference's own
    # rtlsdr_exp / stereo_exp / gn3s_exp_v3 (driver files compiled into oracle/_ref)
    raw = np.arange(256, dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "ingest.npz"),
                        **{f: o.unpack(f, raw) for f in ("rtlsdr", "stereo1", "stereo2", "gn3s2", "gn3s4")})

    # ---- FFT-domain primitives on small, non power-of-two sizes ---------------------
    f = {}
    for n in (12, 62, 1000):
        a = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
        b = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
        f[f"a{n}"], f[f"b{n}"] = a, b
        f[f"fft{n}"], f[f"ifft{n}"] = o.cpxfft(a), o.cpxfft(a, inverse=True)
        f[f"conv{n}"] = o.cpxconv(a, b, n // 2)
        f[f"pspec{n}"] = o.cpxpspec(a)
    x, nsamp, nfft = o.codefft(3, 1, util.FS5)
    f["codefft_prn3_fs2048"] = x
    np.savez_compressed(os.path.join(HERE, "fftprims.npz"), **f)

    # ---- acquisition, config-1 style (real IF, 16.368 MS/s, 71 bins, intg 2) -------
    cfg, vis = util.scene_c1(sdr, nsats=4, seed=0xB2000021)
    for i in range(cfg.nsat):
        cfg.sat[i].cn0_dbhz = 43.0 + 2.0 * i
    host = sdr.ifgen_host(cfg, 0, 65536)
    prns = vis + [2, 32]
    res, pw = o.acquire(ob.cfg_iffile(acq_intg=2), host, 65536, prns, want_power=True)
    np.savez_compressed(os.path.join(HERE, "acq_c1.npz"), if_data=host, prns=np.array(prns),
                        result=np.array([[r.acquired, r.freqi, r.codei, r.nintg_used, r.acqfreq, r.peakr, r.cn0,
                                          r.maxp, r.buffloc] for r in res]),
                        power_row=np.stack([pw[i, res[i].freqi] for i in range(len(prns))]).astype(np.float32))

    # ---- acquisition + tracking, config-5 style (I/Q, 2.048 MS/s) ------------------
    cfg, vis = util.scene_c5(sdr, nsats=3, seed=0xB2000025)
    nper = 150
    total = 65536 + (nper + 4) * 2048
    host = sdr.ifgen_host(cfg, 0, total)
    prns = vis[:2] + [11]
    acq, rec, ntr = o.acq_track(ob.cfg_rtlsdr(acq_intg=2), host, 65536, prns, nper)
    out = dict(if_data=host, prns=np.array(prns), nper=nper, ntracked=np.array(ntr),
               acq=np.array([[r.acquired, r.freqi, r.codei, r.nintg_used, r.acqfreq, r.peakr, r.cn0, r.maxp,
                              r.buffloc] for r in acq]))
    for fld in REC_FIELDS:
        out["rec_" + fld] = np.ascontiguousarray(rec[fld])
    np.savez_compressed(os.path.join(HERE, "trk_c5.npz"), **out)

    # ---- observables (setobsdata, src/sdrtrk.c:160-209): a run long enough to pass bit sync.
    # The IF (5.4 MB) is not stored: it is regenerated from the seeded generator and pinned
    # by its CRC; the reference's records after bit sync are stored.
    import zlib
    cfg, vis = util.scene_c5(sdr, nsats=2, seed=0xB2000026)
    nper = 4800
    total = 65536 + (nper + 4) * 2048
    host = sdr.ifgen_host(cfg, 0, total)
    acq, rec, ntr = o.acq_track(ob.cfg_rtlsdr(acq_intg=2), host, 65536, vis, nper)
    assert int(rec["flagobs"].max()) == 2
    k0 = int(min(np.argmax(rec["flagobs"][c] > 0) for c in range(len(vis)))) - 5
    out = dict(prns=np.array(vis), nper=nper, k0=k0, if_crc=np.uint32(zlib.crc32(host.tobytes())),
               ntracked=np.array(ntr),
               acq=np.array([[r.acquired, r.freqi, r.codei, r.nintg_used, r.acqfreq, r.peakr, r.cn0, r.maxp,
                              r.buffloc] for r in acq]))
    for fld in OBS_FIELDS:
        out["rec_" + fld] = np.ascontiguousarray(rec[fld][:, k0:])
    np.savez_compressed(os.path.join(HERE, "obs_c5.npz"), **out)
    for fn in sorted(os.listdir(HERE)):
        print(fn, os.path.getsize(os.path.join(HERE, fn)))


if __name__ == "__main__":
    main()
