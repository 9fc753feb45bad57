"""CPU: host-side logic of the product and the C-ABI surface (no compute calls)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_bind as ob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(sdrb200_[a-z0-9_]+)\s*\(", txt)))


def test_abi_exports_every_declared_symbol(sdr):
    names = _declared("sdrb200.h")
    assert len(names) > 40
    missing = [n for n in names if not hasattr(sdr.lib, n)]
    assert not missing, missing


def test_no_cpu_fallback(sdr):
    """without a device every compute entry point refuses loudly (ENODEV)"""
    if sdr.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(sdr.Error, match="no CUDA device"):
        sdr.Stream(1, 1024)
    code = np.ones(1023, np.int16)
    with pytest.raises(sdr.Error):
        sdr.correlator(np.zeros(64, np.int8), 1, 1e-6, 64, 0.0, 0.0, 1.023e6, 0.0, [1], code)
    with pytest.raises(sdr.Error):
        sdr.Acquisition(2.048e6, 2, 2048, 2, 1e-3, np.zeros(3), np.ones((1, 1023), np.int8), 1.023e6)


def test_plan_shapes_of_the_baseline_configs(sdr, monkeypatch):
    """The transform a plan builds holds the 2n samples the reference transforms (acq.nfft = 2*nsamp,
    src/sdrinit.c:625) and is the smallest of 2^k and 5*2^(k-3) with kernels; host logic, no device."""
    for v in ("SDRB200_ACQ_POW2", "SDRB200_LSUB", "SDRB200_ACQ_SCRATCH", "SDRB200_ACQ_R5", "SDRB200_ACQ_R8", "SDRB200_ACQ_WARP"):
        monkeypatch.delenv(v, raising=False)
    want = {
        (16368, 1, False): (32768, 4096, 8, True),       # C1: 2n = 32 736
        (2048, 1, False): (4096, 4096, 1, True),         # C5
        (20000, 1, True): (40960, 4096, 10, True),       # C3: 2n = 40 000, one grid per channel
        (10000, 1, False): (20480, 4096, 5, True),
        (5000, 1, False): (10240, 2048, 5, True),
        (2500, 1, False): (5120, 1024, 5, True),
        (16368, 20, False): (655360, 4096, 160, False),  # C4: 2n = 654 720
        (16368, 10, False): (327680, 4096, 80, False),
        (16368, 40, False): (1310720, 8192, 160, False),
        (2048, 40, False): (196608, 4096, 48, False),    # 5 * 2^15 has no radix: 3 * 2^16
        (26000, 1, False): (65536, 4096, 16, True),      # 2n = 52 000 > 49 152
        (24000, 1, False): (49152, 4096, 12, True),      # 3 * 2^k: 0.75 of the next power of two
        (12000, 1, False): (24576, 4096, 6, True),
        (6000, 1, False): (12288, 4096, 3, True),
        (3000, 1, False): (6144, 2048, 3, True),
        (16368, 6, False): (196608, 4096, 48, False),    # 2n = 196 416
        (16368, 12, False): (393216, 4096, 96, False),
        (16368, 24, False): (786432, 4096, 192, False),
        (16368, 48, False): (1572864, 8192, 192, False),
    }
    for (nsamp, ncoh, gpp), (M, L, R, fused) in want.items():
        got = sdr.plan_shape(nsamp, ncoh, gpp)
        assert (got["M"], got["L"], got["R"], got["fused"]) == (M, L, R, fused), ((nsamp, ncoh), got)
        assert got["M"] >= 2 * nsamp * ncoh and got["M"] == got["L"] * got["R"]
    monkeypatch.setenv("SDRB200_ACQ_POW2", "1")
    assert sdr.plan_shape(16368, 20)["M"] == 1 << 20 and sdr.plan_shape(20000, 1, True)["M"] == 65536
    with pytest.raises(sdr.Error):
        sdr.plan_shape(8)


def test_product_codes_match_oracle(sdr):
    """codegen.c (LFSR form) vs the oracle's +-1 register form, all 210 PRNs + G1"""
    o = ob.best()
    for prn in range(1, 211):
        ref, cr = o.gencode(prn, 1)
        got, cr2 = sdr.gencode(prn, 1)
        assert cr == cr2 and np.array_equal(ref, got.astype(np.int16)), prn
    if o.name == "port":
        ref, cr = o.gencode(0, 20)
        got, cr2 = sdr.gencode(0, 20)
        assert cr == cr2 and np.array_equal(ref, got.astype(np.int16))
    with pytest.raises(sdr.Error):
        sdr.gencode(211, 1)


def test_g1_code_against_port(sdr):
    if not ob.have_port():
        pytest.skip("port not built")
    ref, _ = ob.port().gencode(-3, 20)
    got, _ = sdr.gencode(-3, 20)
    assert np.array_equal(ref, got.astype(np.int16))


def test_ifgen_host_statistics(sdr):
    """the synthetic front end: noise level, clipping, determinism, dtype layout"""
    import util
    cfg, _ = util.scene_c1(sdr, nsats=3)
    a = sdr.ifgen_host(cfg, 5000, 40000)
    b = sdr.ifgen_host(cfg, 5000, 40000)
    assert np.array_equal(a, b) and a.dtype == np.int8
    assert abs(float(a.std()) - 24.0) < 1.0 and int(np.abs(a.astype(int)).max()) <= 127
    assert np.array_equal(sdr.ifgen_host(cfg, 5100, 100), a[100:200])       # pure function of the index
    cfg5, _ = util.scene_c5(sdr, nsats=2)
    iq = sdr.ifgen_host(cfg5, 0, 1000)
    assert len(iq) == 2000


def test_loop_coefficients(sdr):
    l = sdr.loop_coeffs(5.0, 30.0, 200.0)                                   # sdrinit.c:411-416
    assert l.dllw2 == (5.0 / 0.53) ** 2 and l.pllaw == 1.414 * (30.0 / 0.53) and l.fllw == 800.0


def test_trk_chan_fill(sdr):
    code, _ = sdr.gencode(7, 1)
    c = sdr.make_trk_chan(16.368e6, 4.092e6, 1575.42e6, 1, 7, code, 1234, 4.0925e6, 6, 3, 6,
                          (5.0, 1.0), (30.0, 10.0), (200.0, 50.0))
    assert (c.nsamp, c.ne, c.nl, c.loopms, c.navrate) == (16368, 3, 4, 10, 20)   # sdrinit.c:441-474
    assert list(c.corrp[:6]) == [3, 6, 9, 12, 15, 18]
