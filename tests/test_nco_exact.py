"""CPU: csrc/nco_exact.h (the parallel-friendly exact replay of the reference's
sequential NCO sums, src/sdrcmn.c:608-621,647-668) against the plain loops, via
tests/csrc/nco_check.c compiled with gcc."""
import ctypes as C
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def nco(tmp_path_factory):
    out = tmp_path_factory.mktemp("nco") / "libnco_check.so"
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared",
                           "-I", os.path.join(ROOT, "erlangnetwork-gnsslib-sdr_b200", "csrc"),
                           os.path.join(ROOT, "tests", "csrc", "nco_check.c"), "-o", str(out), "-lm"])
    lib = C.CDLL(str(out))
    lib.fuzz.argtypes = [C.c_ulonglong, C.c_int, C.c_int]
    lib.check_carrier.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int]
    lib.check_code.argtypes = [C.c_double, C.c_int, C.c_double, C.c_int, C.c_int]
    return lib


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4, 5])
def test_nco_fuzz(nco, mode):
    assert nco.fuzz(1000 + mode, 1500, mode) == 0


def test_nco_edge_cases(nco):
    assert nco.check_carrier(0.0, 0.0, 1 / 2.048e6, 4096) == 0          # bin at exactly 0 Hz
    assert nco.check_carrier(-7000.0, 0.0, 1 / 2.048e6, 4096) == 0      # negative frequency from phase 0
    assert nco.check_carrier(4.092e6, 6.283185307179586, 1 / 16.368e6, 1) == 0
    assert nco.check_carrier(4.099e6, 0.0, 1 / 16.368e6, 654720) == 0   # 20 ms coherent window
    assert nco.check_code(1022.9999, 18, 1.023e6 / 16.368e6, 1023, 16368) == 0
    assert nco.check_code(0.0, 0, 1.023e6 / 16.368e6, 1023, 3 * 16368) == 0   # three code periods
    assert nco.check_code(0.0, 0, 1.023e6 / 16.368e6, 1023, 327360) == -1     # table overflow is reported
    assert nco.check_code(510.5, 2, 0.511e6 / 20e6, 511, 20000) == 0    # GLONASS G1 at 20 MS/s


def test_reduce_2pi_fuzz(nco):
    """mixcarr's `while (prem > DPI) prem -= DPI` (src/sdrcmn.c:666-668): closed-form per-binade
    jumps against the literal loop, random and structured arguments up to 2^21."""
    nco.fuzz_reduce.restype = C.c_int
    assert nco.fuzz_reduce(C.c_int(40000)) == 0
