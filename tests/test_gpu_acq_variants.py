"""GPU: the two warp-per-sub-transform search kernels (k_roww: remote loads, k_rowd: bulk-copy
exchange; csrc/acq.cu) against the default cluster kernel k_rowc on the same IF window.

Oracle here is the product's own default path (itself pinned to the reference build in
tests/test_gpu_acquisition.py / test_gpu_headline.py): acquired / freqi / codei / nintg equal,
peak ratio and C/N0 within the float tolerance of those tests, the power matrix within
2e-4 * max(P), and - the race check - 40 repeated searches bit-identical."""
import os

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


def _plan(sdr, variant, freq, codes, intg, keep_power):
    old = os.environ.get("SDRB200_ACQ_WARP")
    if variant:
        os.environ["SDRB200_ACQ_WARP"] = variant                # read once, at plan creation
    else:
        os.environ.pop("SDRB200_ACQ_WARP", None)
    try:
        return sdr.Acquisition(util.FS1, 1, 16368, 16, 1e-3, freq, codes, 1.023e6, intg=intg, keep_power=keep_power)
    finally:
        if old is None:
            os.environ.pop("SDRB200_ACQ_WARP", None)
        else:
            os.environ["SDRB200_ACQ_WARP"] = old


@pytest.mark.parametrize("variant", ["pull", "push"])
@pytest.mark.parametrize("intg", [1, 3])
def test_warp_kernels_equal_cluster_kernel(sdr, variant, intg):
    cfg, vis = util.scene_c1(sdr, nsats=5)
    prns = vis + [2, 4, 5, 7, 8]
    nsamp = 16368
    avail = 65536 * (((intg + 1) * nsamp) // 65536 + 1)
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    b0 = avail - (intg + 1) * nsamp
    base = _plan(sdr, None, freq, codes, intg, True)
    want = base.run(st, b0)
    alt = _plan(sdr, variant, freq, codes, intg, True)
    got = alt.run(st, b0)
    assert sum(w["acquired"] for w in want) >= 4
    for w, g, p in zip(want, got, prns):
        for k in ("acquired", "freqi", "codei", "nintg", "acqfreq"):
            assert w[k] == g[k], (variant, p, k, w, g)
        assert abs(w["peakr"] - g["peakr"]) <= 1e-3 * w["peakr"], (variant, p, w, g)
        assert abs(w["cn0"] - g["cn0"]) <= 0.01, (variant, p, w, g)
    for i in range(len(prns)):
        P0, P1 = base.power(i), alt.power(i)
        assert np.max(np.abs(P0 - P1)) <= 2e-4 * np.max(P0), (variant, i)
    # register-only search (no power matrix) gives the same tuples as the accumulating one
    plain = _plan(sdr, variant, freq, codes, intg, False)
    assert plain.run(st, b0) == got


@pytest.mark.parametrize("variant", ["pull", "push"])
def test_warp_kernels_repeatable(sdr, variant):
    """READY/FREE mbarrier hand-over between the warps of a cluster: a read that overtook its
    producer, or a tile handed back before its readers were done, changes power values."""
    cfg, vis = util.scene_c1(sdr, nsats=4)
    prns = vis + [2, 4, 5, 7]
    nsamp, avail = 16368, 65536
    st = sdr.Stream(1, avail)
    st.push_ifgen(cfg, avail)
    freq = np.array([util.IF1 + (i - 35) * 200.0 for i in range(71)])
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    acq = _plan(sdr, variant, freq, codes, 1, True)
    plain = _plan(sdr, variant, freq, codes, 1, False)
    b0 = avail - 2 * nsamp
    first = acq.run(st, b0)
    P0 = [acq.power(i).copy() for i in range(len(prns))]
    for rep in range(40):
        assert acq.run(st, b0) == first, rep
        assert plain.run(st, b0) == first, rep
        if rep % 8 == 0:
            for i in range(len(prns)):
                assert np.array_equal(acq.power(i), P0[i]), (rep, i)
