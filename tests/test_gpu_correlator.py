"""GPU parity: sdrb200_correlator vs the oracle's correlator() (src/sdrcmn.c:687-722).
Bit-exact: the tap sums are integers/32 and the remainders identical doubles."""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


def _cases(rng, fs, f_if, n0, smax_taps):
    for _ in range(6):
        n = n0 + int(rng.integers(-1, 2))
        yield dict(n=n, freq=f_if + rng.uniform(-5000, 5000), phi0=rng.uniform(0, 2 * np.pi),
                   crate=1.023e6 + rng.uniform(-4, 4), coff=rng.uniform(0, 0.07), s=smax_taps, ti=1.0 / fs)


def test_correlator_real_bitexact(sdr, oracle):
    rng = np.random.default_rng(11)
    cfg, prns = util.scene_c1(sdr, nsats=3)
    data = sdr.ifgen_host(cfg, 12345, 16400)
    code = util.l1ca_code(sdr, prns[0])
    for off in (0, 1, 7, 13):                       # exercise every load alignment
        for c in _cases(rng, util.FS1, util.IF1, 16368, [3, 6, 9, 12, 15, 18]):
            d = data[off:off + c["n"]]
            ref = oracle.correlator(d, 1, c["ti"], c["n"], c["freq"], c["phi0"], c["crate"], c["coff"], c["s"], code)
            got = sdr.correlator(d, 1, c["ti"], c["n"], c["freq"], c["phi0"], c["crate"], c["coff"], c["s"], code)
            assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1])
            assert ref[2] == got[2] and ref[3] == got[3]


def test_correlator_iq_bitexact(sdr, oracle):
    rng = np.random.default_rng(12)
    cfg, prns = util.scene_c5(sdr, nsats=3)
    data = sdr.ifgen_host(cfg, 999, 2100)
    code = util.l1ca_code(sdr, prns[1])
    for off in (0, 3):
        for c in _cases(rng, util.FS5, 0.0, 2048, [1, 2, 3, 4]):
            d = data[2 * off:2 * (off + c["n"])]
            c["coff"] = rng.uniform(0, 1023)
            ref = oracle.correlator(d, 2, c["ti"], c["n"], c["freq"], c["phi0"], c["crate"], c["coff"], c["s"], code)
            got = sdr.correlator(d, 2, c["ti"], c["n"], c["freq"], c["phi0"], c["crate"], c["coff"], c["s"], code)
            assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1])
            assert ref[2] == got[2] and ref[3] == got[3]


def test_correlator_edge_cases(sdr, oracle):
    code = util.l1ca_code(sdr, 7)
    rng = np.random.default_rng(13)
    d = rng.integers(-128, 128, 64, dtype=np.int8)        # full int8 range incl. -128
    for n in (1, 2, 17, 64):
        ref = oracle.correlator(d[:n], 1, 1 / util.FS1, n, util.IF1, 0.0, 1.023e6, 1022.99, [1], code)
        got = sdr.correlator(d[:n], 1, 1 / util.FS1, n, util.IF1, 0.0, 1.023e6, 1022.99, [1], code)
        assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1])
        assert ref[2] == got[2] and ref[3] == got[3]
    with pytest.raises(sdr.Error):
        sdr.correlator(d, 1, 1 / util.FS1, 0, util.IF1, 0.0, 1.023e6, 0.0, [1], code)


def test_stream_and_ifgen_device(sdr):
    cfg, _ = util.scene_c1(sdr, nsats=4)
    st = sdr.Stream(1, 1 << 20)
    st.push_ifgen(cfg, 300000)
    dev = st.read(1000, 200000)
    host = sdr.ifgen_host(cfg, 1000, 200000)
    diff = np.abs(dev.astype(int) - host.astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 1e-4     # libm vs CUDA rounding only
    st2 = sdr.Stream(1, 70000)                              # ring wrap + guard mirror
    st2.push(host[:60000]); st2.push(host[60000:130000])
    assert np.array_equal(st2.read(65000, 60000), host[65000:125000])
    with pytest.raises(sdr.Error):
        st2.read(0, 10)


def test_device_nco_walk_matches_host_walker(sdr):
    """the device NCO walk of the tracking kernel (device arithmetic) against the host walker (which
    tests/test_nco_exact.py fuzzes against the literal loops of src/sdrcmn.c:608-668):
    carriers, codes with wrap, tie-heavy short mantissas, negative and zero steps."""
    import testkit
    for mode in (0, 1, 2, 3, 4, -1):
        bad = testkit.nco_selftest(3000, 0xB200 + mode + 7, mode)
        assert bad == 0, (mode, bad)


def test_packed_front_end_formats_on_device(sdr, oracle):
    """STEREO / GN3S / RTL-SDR bytes expanded on the device while they are
    appended to the IF stream, against the reference's own unpack routines; ragged
    lengths, unaligned starts and a ring wrap included."""
    rng = np.random.default_rng(11)
    for name, fmt, dtype in (("stereo1", sdr.FMT_STEREO_FE1, 1), ("stereo2", sdr.FMT_STEREO_FE2, 2),
                             ("gn3s2", sdr.FMT_GN3S_2BIT, 1), ("gn3s4", sdr.FMT_GN3S_4BIT, 2)):
        cap = 70000
        st = sdr.Stream(dtype, cap)
        pushed = []
        for n in (1, 15, 16, 17, 4099, 50000, 33333):          # the last push wraps the ring
            raw = rng.integers(0, 256, n).astype(np.uint8)
            st.push_packed(raw, fmt)
            pushed.append(raw)
        allraw = np.concatenate(pushed)
        want = oracle.unpack(name, allraw)
        tot = len(allraw)
        got = st.read(tot - 60000, 60000)
        assert np.array_equal(got, want[(tot - 60000) * dtype:]), name
        st.close()
    st = sdr.Stream(2, 4096)
    raw = rng.integers(0, 256, 2 * 3000).astype(np.uint8)
    st.push_packed(raw, sdr.FMT_RTLSDR)
    assert np.array_equal(st.read(0, 3000), oracle.unpack("rtlsdr", raw))
    with pytest.raises(sdr.Error):
        st.push_packed(raw[:100], sdr.FMT_STEREO_FE1)          # dtype mismatch is refused
