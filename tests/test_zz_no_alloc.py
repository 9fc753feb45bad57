"""GPU: sdrb200_checkacquisition_noalloc (static device scratch, no allocation in the call)
against sdrb200_checkacquisition, through the C ABI.

The entry point has no counterpart in the reference;
its oracle is the existing entry point, which tests/test_gpu_primitives.py pins
to the reference's checkacquisition (src/sdracq.c:71-95).  Both run the same two kernels on
the same doubles, so the comparison is EXACT (codei, freqi equal; cn0, peakr bit-identical).
Two calls in a row on different grids: the second finds the first call's contents in the
shared scratch and must not be disturbed by them.  Collected last (file name) on purpose.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _grid(rng, nsamp, nfreq, peak_at, peak):
    P = rng.random(nsamp * nfreq) * 10.0 + 1.0
    P[peak_at] = peak
    return P


def test_checkacquisition_noalloc_equals_allocating_entry_point(sdr):
    assert sdr.device_count() >= 1, "libsdrb200 has no CPU fallback"
    rng = np.random.default_rng(2025)
    # first call: the C1 grid (71 bins x 16368 code phases), peak near the row end so that the
    # exclusion window wraps; second call: a SMALLER grid, so stale cells of call one lie behind it
    cases = [(16368, 71, 16, 35 * 16368 + 16360, 5000.0), (2048, 3, 2, 1 * 2048 + 1, 900.0),
             (16368, 71, 16, 3 * 16368 + 7, 4000.0)]
    for nsamp, nfreq, nsc, at, pk in cases:
        P = _grid(rng, nsamp, nfreq, at, pk)
        want = sdr.checkacquisition(P, nsamp, nfreq, nsc, 1e-3)
        got = sdr.checkacquisition_noalloc(P, nsamp, nfreq, nsc, 1e-3)
        assert (got["codei"], got["freqi"]) == (want["codei"], want["freqi"]) == (at % nsamp, at // nsamp)
        assert got["cn0"] == want["cn0"] and got["peakr"] == want["peakr"]      # same kernels, same doubles
        again = sdr.checkacquisition_noalloc(P, nsamp, nfreq, nsc, 1e-3)        # back to back on the same scratch
        assert again == got


def test_checkacquisition_noalloc_rejects_oversize(sdr):
    lib = sdr.lib
    nsamp, nfreq = 4096, 513                                   # 2 101 248 cells > 2 097 152
    P = np.ones(nsamp * nfreq)
    cn0, peakr, codei, freqi = C.c_double(), C.c_double(), C.c_int(), C.c_int()
    rc = lib.sdrb200_checkacquisition_noalloc(P.ctypes.data_as(C.POINTER(C.c_double)), nsamp, nfreq, 4,
                                              C.c_double(1e-3), C.byref(cn0), C.byref(peakr),
                                              C.byref(codei), C.byref(freqi))
    assert rc == -5                                            # SDRB200_ERANGE
    lib.sdrb200_last_error.restype = C.c_char_p
    assert b"exceed the static scratch" in lib.sdrb200_last_error()
