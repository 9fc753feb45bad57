"""GPU parity against the committed golden vectors (tests/golden/, recorded from
the reference build): the product's batched engines on the stored IF."""
import os

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_golden_primitives(sdr):
    g = np.load(os.path.join(G, "primitives.npz"))
    for i, prm in enumerate(g["params"]):
        dtype, n = int(prm[0]), int(prm[1])
        ti, freq, phi0, crate, coff = prm[2:]
        d = g["d_real"][:n] if dtype == 1 else g["d_iq"][:2 * n]
        II, QQ, remc, remp = sdr.correlator(d, dtype, ti, n, freq, phi0, crate, coff, g[f"s{i}"], g["code7"])
        assert np.array_equal(II, g[f"II{i}"]) and np.array_equal(QQ, g[f"QQ{i}"])
        assert remc == g[f"rem{i}"][0] and remp == g[f"rem{i}"][1]


def test_golden_acquisition(sdr):
    g = np.load(os.path.join(G, "acq_c1.npz"))
    prns = [int(p) for p in g["prns"]]
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    acq = sdr.Acquisition(util.FS1, 1, 16368, 16, 1e-3, freq, codes, 1.023e6, intg=2, keep_power=True)
    b0 = 65536 - 3 * 16368
    got = acq.run_host(g["if_data"][b0:], 3 * 16368)
    for i, (row, r) in enumerate(zip(g["result"], got)):
        assert (int(row[0]), int(row[1]), int(row[2]), int(row[3])) == (r["acquired"], r["freqi"], r["codei"], r["nintg"])
        assert row[4] == r["acqfreq"] and abs(row[5] - r["peakr"]) <= 1e-3 * row[5] and abs(row[6] - r["cn0"]) <= 0.01
        if r["acquired"]:
            assert int(row[8]) == b0 + r["codei"]
        P = acq.power(i)[r["freqi"]]
        assert np.abs(P - g["power_row"][i]).max() <= 2e-4 * g["power_row"][i].max()


def test_golden_tracking(sdr):
    g = np.load(os.path.join(G, "trk_c5.npz"))
    prns = [int(p) for p in g["prns"]]
    nper = int(g["nper"])
    st = sdr.Stream(2, len(g["if_data"]) // 2)
    st.push(g["if_data"])
    chans, rows = [], []
    for i, (p, a) in enumerate(zip(prns, g["acq"])):
        if int(a[0]):
            chans.append(sdr.make_trk_chan(util.FS5, 0.0, util.CF, 2, p, sdr.gencode(p)[0], int(a[8]), a[4],
                                           4, 1, 1, (5.0, 2.0), (30.0, 20.0), (200.0, 50.0)))
            rows.append(i)
    rec, nrec = sdr.Tracking(chans).run(st, nper)
    for k, i in enumerate(rows):
        n = min(nrec[k], int(g["ntracked"][i]))
        assert n == nper
        for f in ("cnt", "buffloc", "currnsamp", "flagsync", "flagloopfilter", "swloop", "synci", "biti"):
            assert np.array_equal(rec[k][f][:n], g["rec_" + f][i][:n]), f
        assert np.abs(rec[k]["II"][:n, :9] - g["rec_II"][i][:n, :9]).max() <= 508.0
        assert np.abs(rec[k]["QQ"][:n, :9] - g["rec_QQ"][i][:n, :9]).max() <= 508.0
        assert np.abs(rec[k]["carrfreq"][:n] - g["rec_carrfreq"][i][:n]).max() <= 1e-6
        assert np.abs(rec[k]["codefreq"][:n] - g["rec_codefreq"][i][:n]).max() <= 1e-6


def test_golden_observables(sdr):
    """setobsdata on the device (src/sdrtrk.c:160-209) against the reference's records after
    bit sync; IF regenerated from the seeded generator on the device and pinned by its CRC."""
    import zlib
    g = np.load(os.path.join(G, "obs_c5.npz"))
    prns, nper, k0 = [int(p) for p in g["prns"]], int(g["nper"]), int(g["k0"])
    cfg, vis = util.scene_c5(sdr, nsats=2, seed=0xB2000026)
    total = 65536 + (nper + 4) * 2048
    st = sdr.Stream(2, total)
    st.push_ifgen(cfg, total)
    assert np.uint32(zlib.crc32(st.read(0, total).tobytes())) == g["if_crc"]
    chans = [sdr.make_trk_chan(util.FS5, 0.0, util.CF, 2, p, sdr.gencode(p)[0], int(a[8]), a[4],
                               4, 1, 1, (5.0, 2.0), (30.0, 20.0), (200.0, 50.0)) for p, a in zip(prns, g["acq"])]
    rec, nrec = sdr.Tracking(chans).run(st, nper)
    assert list(nrec) == [int(x) for x in g["ntracked"]]
    for k in range(len(prns)):
        r = rec[k][k0:]
        for f in ("cnt", "buffloc", "flagsync", "flagloopfilter", "flagobs", "obs_codei", "obs_cnt"):
            assert np.array_equal(r[f], g["rec_" + f][k]), f
        assert int(r["flagobs"].max()) == 2
        assert np.abs(r["D0"] - g["rec_D0"][k]).max() <= 5e-3            # follows carrfreq (rel 1e-9 of IF)
        assert np.abs(r["L0"] - g["rec_L0"][k]).max() <= 2e-2            # integrates D0
        assert np.abs(r["remcout0"] - g["rec_remcout0"][k]).max() <= 2e-5
        assert np.abs(r["S0"] - g["rec_S0"][k]).max() <= 1e-6
