"""Host-side Python mirror of the libsdrb200 C ABI (include/sdrb200.h).

This package is plumbing only: it loads the in-tree ``libsdrb200.so`` with
ctypes and mirrors the C structs.  All arithmetic of the hot path runs in the
CUDA library; there is no Python or CPU fallback - if the library is missing
the import fails loudly, and every compute entry point raises when no CUDA
device is present.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsdrb200.so")
if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with __graft_entry__.build() "
        "(make -C erlangnetwork-gnsslib-sdr_b200); there is no fallback path")
lib = C.CDLL(LIB_PATH)

OK, EINVAL, ENODEV, ENOMEM, ECUDA, ERANGE = 0, -1, -2, -3, -4, -5
CTYPE_L1CA, CTYPE_G1, CTYPE_L1SBAS = 1, 20, 27
FMT_RTLSDR, FMT_STEREO_FE1, FMT_STEREO_FE2, FMT_GN3S_2BIT, FMT_GN3S_4BIT = 1, 2, 3, 4, 5
MAXCORRN = 16
MAXTAPS = 1 + 2 * MAXCORRN


class Error(RuntimeError):
    pass


lib.sdrb200_last_error.restype = C.c_char_p
lib.sdrb200_version.restype = C.c_char_p
lib.sdrb200_launch_count.restype = C.c_uint64
lib.sdrb200_stream_count.restype = C.c_uint64
lib.sdrb200_stream_count.argtypes = [C.c_void_p]


def check(rc: int, what: str = "") -> None:
    if rc != OK:
        raise Error(f"{what} failed ({rc}): {lib.sdrb200_last_error().decode()}")


# ----------------------------------------------------------------- structs
class AcqCfg(C.Structure):
    _fields_ = [("f_sf", C.c_double), ("dtype", C.c_int), ("nsamp", C.c_int),
                ("nsampchip", C.c_int), ("ctime", C.c_double), ("nfreq", C.c_int),
                ("freq", C.POINTER(C.c_double)), ("nprn", C.c_int), ("clen", C.c_int),
                ("crate", C.c_double), ("codes", C.POINTER(C.c_int8)), ("intg", C.c_int),
                ("ncoh", C.c_int), ("threshold", C.c_double), ("grid_per_prn", C.c_int)]


class AcqResult(C.Structure):
    _fields_ = [("acquired", C.c_int), ("freqi", C.c_int), ("codei", C.c_int),
                ("nintg", C.c_int), ("acqfreq", C.c_double), ("peakr", C.c_double),
                ("cn0", C.c_double), ("maxp", C.c_double), ("maxp2", C.c_double),
                ("meanp", C.c_double)]


class Loop(C.Structure):
    _fields_ = [("dllw2", C.c_double), ("dllaw", C.c_double), ("pllw2", C.c_double),
                ("pllaw", C.c_double), ("fllw", C.c_double)]


class TrkChan(C.Structure):
    _fields_ = [("f_sf", C.c_double), ("f_if", C.c_double), ("f_cf", C.c_double),
                ("foffset", C.c_double), ("dtype", C.c_int), ("clen", C.c_int),
                ("crate", C.c_double), ("ctime", C.c_double), ("nsamp", C.c_int),
                ("prn", C.c_int), ("corrn", C.c_int), ("corrp", C.c_int * MAXCORRN),
                ("ne", C.c_int), ("nl", C.c_int), ("loopms", C.c_int), ("navrate", C.c_int),
                ("prm1", Loop), ("prm2", Loop), ("code", C.POINTER(C.c_int8)),
                ("buffloc", C.c_uint64), ("acqfreq", C.c_double)]


class TrkRec(C.Structure):
    _fields_ = [("cnt", C.c_uint64), ("buffloc", C.c_uint64), ("currnsamp", C.c_int32),
                ("flagsync", C.c_int32), ("flagloopfilter", C.c_int32), ("swsync", C.c_int32),
                ("swloop", C.c_int32), ("synci", C.c_int32), ("biti", C.c_int32),
                ("bit", C.c_int32), ("II", C.c_double * MAXTAPS), ("QQ", C.c_double * MAXTAPS),
                ("sumI0", C.c_double), ("sumQ0", C.c_double), ("codefreq", C.c_double),
                ("codeErr", C.c_double), ("codeNco", C.c_double), ("carrfreq", C.c_double),
                ("carrErr", C.c_double), ("carrNco", C.c_double), ("freqErr", C.c_double),
                ("remcode", C.c_double), ("remcarr", C.c_double),
                ("L0", C.c_double), ("D0", C.c_double), ("S0", C.c_double), ("remcout0", C.c_double),
                ("Isum", C.c_double), ("obs_codei", C.c_uint64), ("obs_cnt", C.c_uint64),
                ("flagobs", C.c_int32), ("flagsyncf", C.c_int32), ("polarity", C.c_int32), ("flagtow", C.c_int32),
                ("firstsf", C.c_uint64), ("firstsfcnt", C.c_uint64), ("obs_dcnt", C.c_int64)]


OBSINTERPN = 80


class TrkObs(C.Structure):
    _fields_ = [("L", C.c_double * OBSINTERPN), ("D", C.c_double * OBSINTERPN), ("S", C.c_double * OBSINTERPN),
                ("remcout", C.c_double * OBSINTERPN), ("codei", C.c_uint64 * OBSINTERPN),
                ("cntout", C.c_uint64 * OBSINTERPN), ("codeisum", C.c_uint64 * OBSINTERPN)]


def loop_coeffs(dllb: float, pllb: float, fllb: float) -> Loop:
    """inittrkprmstruct(), src/sdrinit.c:411-422."""
    return Loop(dllw2=(dllb / 0.53) * (dllb / 0.53), dllaw=1.414 * (dllb / 0.53),
                pllw2=(pllb / 0.53) * (pllb / 0.53), pllaw=1.414 * (pllb / 0.53),
                fllw=fllb / 0.25)


def plan_shape(nsamp: int, ncoh: int = 1, grid_per_prn: bool = False) -> dict:
    """Transform length and decomposition an acquisition plan will use for windows of nsamp*ncoh lags
    (sdrb200_acq_plan_shape, host logic only): M = R sub-transforms of L points, fused = cluster kernel."""
    M, L, R, fused = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    check(lib.sdrb200_acq_plan_shape(int(nsamp), int(ncoh), int(bool(grid_per_prn)), C.byref(M), C.byref(L), C.byref(R),
                                     C.byref(fused)), "acq_plan_shape")
    return {"M": M.value, "L": L.value, "R": R.value, "fused": bool(fused.value)}


from .host import *  # noqa: E402,F401,F403  (convenience wrappers)
from . import shard  # noqa: E402,F401
