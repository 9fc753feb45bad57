"""GPU parity: the single-call primitives (host buffers in/out) against the oracle.

Bit-exact (integer work): mixcarr I/Q streams and remainder, rescode output and
remainder, dot_2x sums, cpxcpx.  Tolerance (float32 FFTs of different structure;
the oracle's FFT is double inside): |d| <= 2e-5*max|X| for transforms (Bluestein
for the non-power-of-two lengths), 2e-4*max for powers, stated per assert."""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


def test_mixcarr_rescode_bitexact(sdr, oracle):
    rng = np.random.default_rng(21)
    code = util.l1ca_code(sdr, 9)
    for dtype, n, fs, f0 in ((1, 16368, util.FS1, util.IF1), (2, 2048, util.FS5, 0.0), (1, 40000, util.FS1, util.IF1)):
        d = rng.integers(-128, 128, n * dtype, dtype=np.int8)
        for _ in range(3):
            freq, phi0 = f0 + rng.uniform(-7000, 7000), rng.uniform(0, 2 * np.pi)
            r = oracle.mixcarr(d, dtype, 1 / fs, n, freq, phi0)
            g = sdr.mixcarr(d, dtype, 1 / fs, n, freq, phi0)
            assert np.array_equal(r[0], g[0]) and np.array_equal(r[1], g[1]) and r[2] == g[2]
            coff, smax, ci = rng.uniform(0, 1023), int(rng.integers(0, 19)), (1.023e6 + rng.uniform(-5, 5)) / fs
            r = oracle.rescode(code, coff, smax, ci, n)
            g = sdr.rescode(code, coff, smax, ci, n)
            assert np.array_equal(r[0], g[0]) and r[1] == g[1]


def test_dot_and_cpxcpx_bitexact(sdr):
    """dot_21/22/23 (src/sdrcmn.c:287-354) and cpxcpx (:185-195) as single calls.  Oracle: numpy int64 /
    float32 arithmetic written out here, NOT the reference build - the integer sums are exact in any order
    and the cpxcpx products are single float32 operations, so nothing can differ; the reference's own
    functions are exercised on the same paths by the drop-in tests (tests/test_gpu_dropin.py)."""
    import ctypes as C
    rng = np.random.default_rng(22)
    n = 16371
    a1 = rng.integers(-4064, 4065, n).astype(np.int16)
    a2 = rng.integers(-4064, 4065, n).astype(np.int16)
    b = [rng.choice(np.array([-1, 1], np.int16), n) for _ in range(3)]
    for nb in (1, 2, 3):
        arr = (C.POINTER(C.c_short) * nb)(*[x.ctypes.data_as(C.POINTER(C.c_short)) for x in b[:nb]])
        d1, d2 = np.zeros(3), np.zeros(3)
        sdr.check(sdr.lib.sdrb200_dot2x(a1.ctypes.data_as(C.POINTER(C.c_short)), a2.ctypes.data_as(C.POINTER(C.c_short)),
                                        arr, nb, n, d1.ctypes.data_as(C.POINTER(C.c_double)),
                                        d2.ctypes.data_as(C.POINTER(C.c_double))), "dot2x")
        for j in range(nb):
            assert d1[j] == float(np.dot(a1.astype(np.int64), b[j])) and d2[j] == float(np.dot(a2.astype(np.int64), b[j]))
    out = np.zeros(n, np.complex64)
    sdr.check(sdr.lib.sdrb200_cpxcpx(a1.ctypes.data_as(C.POINTER(C.c_short)), a2.ctypes.data_as(C.POINTER(C.c_short)),
                                     C.c_double(1 / 32 / 32736), n, out.ctypes.data_as(C.POINTER(C.c_float))), "cpxcpx")
    sc = np.float32(1 / 32 / 32736)
    assert np.array_equal(out.real, a1.astype(np.float32) * sc) and np.array_equal(out.imag, a2.astype(np.float32) * sc)


@pytest.mark.parametrize("n", [1, 12, 62, 256, 1000, 4096, 32736, 32768, 40000, 65536])
def test_cpxfft_any_length(sdr, oracle, n):
    rng = np.random.default_rng(n)
    a = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    for inv in (False, True):
        ref = np.fft.ifft(a.astype(np.complex128)) * n if inv else np.fft.fft(a.astype(np.complex128))
        got = sdr.cpxfft(a, inverse=inv)
        assert np.abs(got - ref).max() <= 2e-5 * np.abs(ref).max() + 1e-6
    if n <= 4096:
        assert np.abs(sdr.cpxfft(a) - oracle.cpxfft(a)).max() <= 2e-5 * np.abs(oracle.cpxfft(a)).max() + 1e-6


@pytest.mark.parametrize("m", [1000, 4096, 32736])
def test_cpxconv_cpxpspec(sdr, oracle, m):
    rng = np.random.default_rng(m)
    a = (rng.standard_normal(m) + 1j * rng.standard_normal(m)).astype(np.complex64)
    b = (rng.standard_normal(m) + 1j * rng.standard_normal(m)).astype(np.complex64)
    n = m // 2
    ref = oracle.cpxconv(a, b, n)
    got = sdr.cpxconv(a, b, n)
    assert np.abs(got - ref).max() <= 2e-4 * ref.max()
    acc = rng.uniform(0, 1, n)
    assert np.abs(sdr.cpxconv(a, b, n, conv=acc) - oracle.cpxconv(a, b, n, conv=acc)).max() <= 2e-4 * (ref.max() + 1)
    ref = oracle.cpxpspec(a)
    assert np.abs(sdr.cpxpspec(a) - ref).max() <= 2e-4 * ref.max()


def test_pcorrelator_checkacquisition(sdr, oracle):
    """the reference's own call: m = 2n = 4096 and the non power-of-two m = 32736"""
    for (fs, f_if, dtype, scene) in ((util.FS5, 0.0, 2, util.scene_c5), (util.FS1, util.IF1, 1, util.scene_c1)):
        cfg, vis = scene(sdr, nsats=3)
        n = int(fs * 1e-3)
        data = sdr.ifgen_host(cfg, 0, 2 * n)
        x, nsamp, nfft = oracle.codefft(vis[0], 1, fs)
        assert nsamp == n and nfft == 2 * n
        freq = np.array([f_if + (i - 10) * 200.0 + np.round(cfg.sat[0].doppler_hz / 200) * 200 for i in range(21)])
        ref = oracle.pcorrelator(data, dtype, 1 / fs, n, freq, 1.023e6, 2 * n, x)
        got = sdr.pcorrelator(data, dtype, 1 / fs, n, freq, 1.023e6, 2 * n, x)
        assert np.abs(got - ref).max() <= 2e-4 * ref.max()
        assert int(np.argmax(got)) == int(np.argmax(ref))
        r2 = oracle.pcorrelator(data, dtype, 1 / fs, n, freq, 1.023e6, 2 * n, x, P=ref)     # accumulation
        g2 = sdr.pcorrelator(data, dtype, 1 / fs, n, freq, 1.023e6, 2 * n, x, P=got)
        assert np.abs(g2 - r2).max() <= 2e-4 * r2.max()
        ca = sdr.checkacquisition(ref, n, len(freq), int(n / 1023), 1e-3)
        flat = int(np.argmax(ref))
        assert (ca["codei"], ca["freqi"]) == (flat % n, flat // n)
        row, ci, w = ref.reshape(len(freq), n)[flat // n], flat % n, 2 * int(n / 1023)
        keep = np.ones(n, bool)
        keep[np.arange(ci - w, ci + w + 1) % n] = False                # src/sdracq.c:81-82
        p2 = max(row[keep].max(), row[0])                              # maxvd seed, src/sdrcmn.c:464
        assert abs(ca["peakr"] - row[ci] / p2) <= 1e-12 * ca["peakr"]
        assert abs(ca["cn0"] - 10 * np.log10(row[ci] / row[keep].mean() / 1e-3)) <= 1e-9
    # the maxvd seed quirk: peak next to lag 0 (src/sdrcmn.c:464)
    P = np.random.default_rng(5).uniform(0, 1, (3, 2048))
    P[1, 1] = 50.0
    P[1, 0] = 20.0
    ca = sdr.checkacquisition(P.ravel(), 2048, 3, 2, 1e-3)
    assert (ca["codei"], ca["freqi"]) == (1, 1) and abs(ca["peakr"] - 2.5) < 1e-12


def test_gencode_on_device(sdr):
    """gencode (src/sdrcode.c:101-154, 426-444) on the GPU: all 210 C/A-family PRNs (GPS, SBAS
    120-158, QZSS) and the GLONASS ST code, chip for chip against the host tables, which
    tests/test_host_cpu.py and tests/golden/codes.npz pin to the reference's generator."""
    prns = list(range(1, 211))
    dev = sdr.gencode_device(prns, 1)
    assert dev.shape == (210, 1023)
    for i, p in enumerate(prns):
        assert np.array_equal(dev[i], sdr.gencode(p, 1)[0]), p
    assert np.array_equal(sdr.gencode_device([133, 120], 27), np.stack([sdr.gencode(133, 27)[0], sdr.gencode(120, 27)[0]]))
    g1 = sdr.gencode_device([1], 20)
    assert g1.shape == (1, 511) and np.array_equal(g1[0], sdr.gencode(1, 20)[0])
    with pytest.raises(sdr.Error):
        sdr.gencode_device([211], 1)
