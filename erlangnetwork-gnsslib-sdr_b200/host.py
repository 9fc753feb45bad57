"""numpy-facing wrappers over the C ABI; names follow the reference (sdr.h)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import (AcqCfg, AcqResult, Error, Loop, TrkChan, TrkObs, TrkRec,  # noqa: F401
               MAXCORRN, MAXTAPS, check, lib)

__all__ = ["gencode", "gencode_device", "correlator", "mixcarr", "rescode",
           "cpxfft", "cpxconv", "cpxpspec", "pcorrelator", "checkacquisition", "checkacquisition_noalloc",
           "Stream", "Acquisition", "Tracking", "DeviceBytes", "device_count", "launch_count", "make_trk_chan", "trk_log"]

_i8p = C.POINTER(C.c_int8)
_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_sp = C.POINTER(C.c_short)
_ip = C.POINTER(C.c_int)


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(t)


def device_count() -> int:
    return int(lib.sdrb200_device_count())


def launch_count() -> int:
    return int(lib.sdrb200_launch_count())


def gencode(prn: int, ctype: int = 1):
    """gencode(), src/sdrcode.c:523-539 -> (int8 code, chip rate)."""
    buf = np.zeros(1023, np.int8)
    crate = C.c_double()
    n = lib.sdrb200_gencode(int(prn), int(ctype), _p(buf, _i8p), C.byref(crate))
    if n <= 0:
        raise Error(f"gencode prn:{prn} ctype:{ctype}")
    return buf[:n].copy(), crate.value


def gencode_device(prns, ctype: int = 1) -> np.ndarray:
    """gencode on the GPU for a list of PRNs -> [n][len] int8 (sdrb200_gencode_device)."""
    n = len(prns)
    out = np.zeros((n, 1023), np.int8)
    arr = (C.c_int * n)(*[int(p) for p in prns])
    ln = lib.sdrb200_gencode_device(n, arr, int(ctype), _p(out, _i8p), 1023)
    check(ln if ln < 0 else 0, "gencode_device")
    return out[:, :ln]


# ------------------------------------------------------ single-call primitives
def correlator(data, dtype, ti, n, freq, phi0, crate, coff, s, code):
    """correlator(), src/sdrcmn.c:687-722 -> (II, QQ, remc, remp)."""
    data = np.ascontiguousarray(data, np.int8)
    s = np.ascontiguousarray(s, np.int32)
    code = np.ascontiguousarray(code, np.int16)
    nt = 1 + 2 * len(s)
    II, QQ = np.zeros(nt), np.zeros(nt)
    remc, remp = C.c_double(), C.c_double()
    check(lib.sdrb200_correlator(_p(data, _i8p), int(dtype), C.c_double(ti), int(n), C.c_double(freq),
                                 C.c_double(phi0), C.c_double(crate), C.c_double(coff), _p(s, _ip),
                                 len(s), _p(II, _dp), _p(QQ, _dp), C.byref(remc), C.byref(remp),
                                 _p(code, _sp), len(code)), "correlator")
    return II, QQ, remc.value, remp.value


def mixcarr(data, dtype, ti, n, freq, phi0):
    """mixcarr(), src/sdrcmn.c:633-669 -> (II, QQ, remainder)."""
    data = np.ascontiguousarray(data, np.int8)
    II, QQ = np.zeros(n, np.int16), np.zeros(n, np.int16)
    rem = C.c_double()
    check(lib.sdrb200_mixcarr(_p(data, _i8p), int(dtype), C.c_double(ti), int(n), C.c_double(freq),
                              C.c_double(phi0), _p(II, _sp), _p(QQ, _sp), C.byref(rem)), "mixcarr")
    return II, QQ, rem.value


def rescode(code, coff, smax, ci, n):
    """rescode(), src/sdrcmn.c:608-621 -> (rcode[n+2*smax], remainder)."""
    code = np.ascontiguousarray(code, np.int16)
    out = np.zeros(n + 2 * smax, np.int16)
    rem = C.c_double()
    check(lib.sdrb200_rescode(_p(code, _sp), len(code), C.c_double(coff), int(smax), C.c_double(ci),
                              int(n), _p(out, _sp), C.byref(rem)), "rescode")
    return out, rem.value


def cpxfft(cpx, inverse=False):
    """cpxfft()/cpxifft(), src/sdrcmn.c:134-175; cpx: complex64[n], returns new array."""
    a = np.ascontiguousarray(cpx, np.complex64).copy()
    check(lib.sdrb200_cpxfft(_p(a, _fp), len(a), int(bool(inverse))), "cpxfft")
    return a


def cpxconv(cpxa, cpxb, n, conv=None):
    """cpxconv(), src/sdrcmn.c:228-251; flagsum iff conv is given."""
    a = np.ascontiguousarray(cpxa, np.complex64).copy()
    b = np.ascontiguousarray(cpxb, np.complex64)
    flagsum = conv is not None
    out = np.ascontiguousarray(conv, np.float64).copy() if flagsum else np.zeros(n)
    check(lib.sdrb200_cpxconv(_p(a, _fp), _p(b, _fp), len(a), int(n), int(flagsum), _p(out, _dp)), "cpxconv")
    return out


def cpxpspec(cpx, pspec=None):
    """cpxpspec(), src/sdrcmn.c:261-276."""
    a = np.ascontiguousarray(cpx, np.complex64).copy()
    flagsum = pspec is not None
    out = np.ascontiguousarray(pspec, np.float64).copy() if flagsum else np.zeros(len(a))
    check(lib.sdrb200_cpxpspec(_p(a, _fp), len(a), int(flagsum), _p(out, _dp)), "cpxpspec")
    return out


def pcorrelator(data, dtype, ti, n, freq, crate, m, codex, P=None):
    """pcorrelator(), src/sdrcmn.c:738-773; accumulates into P[nfreq*n]."""
    data = np.ascontiguousarray(data, np.int8)
    freq = np.ascontiguousarray(freq, np.float64)
    codex = np.ascontiguousarray(codex, np.complex64)
    out = np.zeros(len(freq) * n) if P is None else np.ascontiguousarray(P, np.float64).copy()
    check(lib.sdrb200_pcorrelator(_p(data, _i8p), int(dtype), C.c_double(ti), int(n), _p(freq, _dp),
                                  len(freq), C.c_double(crate), int(m), _p(codex, _fp), _p(out, _dp)),
          "pcorrelator")
    return out


def checkacquisition(P, nsamp, nfreq, nsampchip, ctime):
    """checkacquisition(), src/sdracq.c:71-95 -> dict."""
    P = np.ascontiguousarray(P, np.float64)
    cn0, peakr = C.c_double(), C.c_double()
    codei, freqi = C.c_int(), C.c_int()
    check(lib.sdrb200_checkacquisition(_p(P, _dp), int(nsamp), int(nfreq), int(nsampchip), C.c_double(ctime),
                                       C.byref(cn0), C.byref(peakr), C.byref(codei), C.byref(freqi)),
          "checkacquisition")
    return dict(cn0=cn0.value, peakr=peakr.value, codei=codei.value, freqi=freqi.value)


def checkacquisition_noalloc(P, nsamp, nfreq, nsampchip, ctime):
    """sdrb200_checkacquisition_noalloc: the allocation-free variant (static device scratch,
    at most SDRB200_CHECKACQ_NOALLOC_MAXCELLS cells; no counterpart in the reference)."""
    P = np.ascontiguousarray(P, np.float64)
    cn0, peakr = C.c_double(), C.c_double()
    codei, freqi = C.c_int(), C.c_int()
    check(lib.sdrb200_checkacquisition_noalloc(_p(P, _dp), int(nsamp), int(nfreq), int(nsampchip),
                                               C.c_double(ctime), C.byref(cn0), C.byref(peakr),
                                               C.byref(codei), C.byref(freqi)),
          "checkacquisition_noalloc")
    return dict(cn0=cn0.value, peakr=peakr.value, codei=codei.value, freqi=freqi.value)


class DeviceBytes:
    """A span of device memory owned by the library, exposed through the CUDA array interface
    so that torch.as_tensor(span, device="cuda") wraps it without a copy (for NCCL)."""

    def __init__(self, ptr: int, nbytes: int):
        if not ptr:
            raise Error("device buffer not available")
        self.ptr, self.nbytes = int(ptr), int(nbytes)
        self.__cuda_array_interface__ = {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False),
                                         "version": 2, "strides": None}


# --------------------------------------------------------------- IF stream
class Stream:
    """Device-resident IF buffer with absolute sample indexing (rcvgetbuff
    semantics, src/sdrrcv.c:406-463)."""

    def __init__(self, dtype: int, capacity: int):
        self.h = C.c_void_p()
        self.dtype = dtype
        check(lib.sdrb200_stream_create(C.byref(self.h), int(dtype), C.c_uint64(capacity)), "stream_create")

    def push(self, host: np.ndarray) -> None:
        host = np.ascontiguousarray(host, np.int8)
        check(lib.sdrb200_stream_push_host(self.h, _p(host, _i8p), C.c_uint64(len(host) // self.dtype)), "push")

    def push_rtlsdr(self, host: np.ndarray) -> None:
        host = np.ascontiguousarray(host, np.uint8)
        check(lib.sdrb200_stream_push_host_rtlsdr(self.h, host.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                  C.c_uint64(len(host) // 2)), "push_rtlsdr")

    def push_packed(self, host: np.ndarray, fmt: int) -> None:
        """one packed byte per sample: STEREO / GN3S front-end formats (FMT_* of sdrb200.h)."""
        host = np.ascontiguousarray(host, np.uint8)
        check(lib.sdrb200_stream_push_host_packed(self.h, host.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                  C.c_uint64(len(host) // 2 if int(fmt) == 1 else len(host)), int(fmt)),
              "push_packed")                      # FMT_RTLSDR: two bytes (I, Q) per sample

    def push_device(self, d_ptr: int, n: int) -> None:
        """n samples that are already in device memory at address d_ptr (device-to-device append)."""
        check(lib.sdrb200_stream_push_device(self.h, C.c_void_p(d_ptr), C.c_uint64(n)), "push_device")

    @property
    def count(self) -> int:
        return int(lib.sdrb200_stream_count(self.h))

    def read(self, buffloc: int, n: int) -> np.ndarray:
        out = np.empty(n * self.dtype, np.int8)
        check(lib.sdrb200_stream_read(self.h, C.c_uint64(buffloc), C.c_uint64(n), _p(out, _i8p)), "stream_read")
        return out

    def close(self):
        if self.h:
            lib.sdrb200_stream_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# -------------------------------------------------------------- acquisition
class Acquisition:
    """Batched sdracquisition/pcorrelator/checkacquisition over PRN x Doppler
    (src/sdracq.c:14-95, src/sdrcmn.c:738-773)."""

    def __init__(self, f_sf, dtype, nsamp, nsampchip, ctime, freq, codes, crate,
                 intg=10, ncoh=1, threshold=3.0, keep_power=False):
        self.freq = np.ascontiguousarray(freq, np.float64)
        self.codes = np.ascontiguousarray(codes, np.int8)
        assert self.codes.ndim == 2
        self.nprn, clen = self.codes.shape
        # freq [nfreq]: one Doppler grid for all replicas; freq [nprn][nfreq]: one grid per replica
        # (GLONASS FDMA channels, each around its own carrier, src/sdrinit.c:633-635)
        per_prn = self.freq.ndim == 2
        assert not per_prn or self.freq.shape[0] == self.nprn
        self.nsamp, self.nfreq = nsamp, self.freq.shape[-1]
        self.n_eff = nsamp * max(int(ncoh), 1)               # lags per row (coherent block length)
        cfg = AcqCfg(f_sf=f_sf, dtype=dtype, nsamp=nsamp, nsampchip=nsampchip, ctime=ctime,
                     nfreq=self.nfreq, freq=_p(self.freq, _dp), nprn=self.nprn, clen=clen,
                     crate=crate, codes=_p(self.codes, _i8p), intg=intg, ncoh=ncoh, threshold=threshold,
                     grid_per_prn=int(per_prn))
        self.h = C.c_void_p()
        check(lib.sdrb200_acq_create(C.byref(self.h), C.byref(cfg)), "acq_create")
        if keep_power:
            self.keep_power(True)

    def keep_power(self, on: bool):
        """Store the power matrix in HBM too, so that power() can read it back."""
        check(lib.sdrb200_acq_keep_power(self.h, int(bool(on))), "acq_keep_power")

    def _results(self, arr):
        return [dict(acquired=r.acquired, freqi=r.freqi, codei=r.codei, nintg=r.nintg,
                     acqfreq=r.acqfreq, peakr=r.peakr, cn0=r.cn0, maxp=r.maxp, maxp2=r.maxp2,
                     meanp=r.meanp) for r in arr]

    def run(self, stream: Stream, buffloc0: int):
        res = (AcqResult * self.nprn)()
        check(lib.sdrb200_acq_run(self.h, stream.h, C.c_uint64(buffloc0), res), "acq_run")
        return self._results(res)

    def run_host(self, host_if: np.ndarray, nsamples: int):
        host_if = np.ascontiguousarray(host_if, np.int8)
        res = (AcqResult * self.nprn)()
        check(lib.sdrb200_acq_run_host(self.h, _p(host_if, _i8p), C.c_uint64(nsamples), res), "acq_run_host")
        return self._results(res)

    def enqueue(self, stream: Stream, buffloc0: int):
        check(lib.sdrb200_acq_enqueue(self.h, stream.h, C.c_uint64(buffloc0)), "acq_enqueue")

    def fetch(self):
        res = (AcqResult * self.nprn)()
        check(lib.sdrb200_acq_fetch(self.h, res), "acq_fetch")
        return self._results(res)

    # ---- multi-GPU plumbing: tuples stay in device memory until one collective gathers them
    RESULT_BYTES = C.sizeof(AcqResult)

    def sync(self):
        check(lib.sdrb200_acq_sync(self.h), "acq_sync")

    def results_device(self) -> "DeviceBytes":
        """the [nprn] tuples of the last enqueue, in place in device memory (no copy)."""
        lib.sdrb200_acq_results_device.restype = C.c_void_p
        return DeviceBytes(lib.sdrb200_acq_results_device(self.h), self.nprn * self.RESULT_BYTES)

    def copy_results_device(self, d_ptr: int):
        """enqueue a device-to-device copy of the tuples to caller memory on the plan's stream."""
        check(lib.sdrb200_acq_copy_results_device(self.h, C.c_void_p(d_ptr)), "acq_copy_results_device")

    def set_history(self, on: bool = True):
        """keep the tuples of every integration step (Doppler-slab searches)."""
        self.intg = getattr(self, "intg", None)
        check(lib.sdrb200_acq_set_history(self.h, int(bool(on))), "acq_set_history")

    def history_device(self, intg: int) -> "DeviceBytes":
        lib.sdrb200_acq_history_device.restype = C.c_void_p
        return DeviceBytes(lib.sdrb200_acq_history_device(self.h), intg * self.nprn * self.RESULT_BYTES)

    def fetch_history(self, intg: int):
        """[intg][nprn] dicts."""
        arr = (AcqResult * (intg * self.nprn))()
        check(lib.sdrb200_acq_fetch_history(self.h, arr), "acq_fetch_history")
        flat = self._results(arr)
        return [flat[i * self.nprn:(i + 1) * self.nprn] for i in range(intg)]

    @staticmethod
    def results_from_bytes(buf) -> list:
        """tuples from raw bytes (a gathered buffer): list of dicts."""
        raw = bytes(buf)
        n = len(raw) // C.sizeof(AcqResult)
        arr = (AcqResult * n).from_buffer_copy(raw)
        return [dict(acquired=r.acquired, freqi=r.freqi, codei=r.codei, nintg=r.nintg,
                     acqfreq=r.acqfreq, peakr=r.peakr, cn0=r.cn0, maxp=r.maxp, maxp2=r.maxp2,
                     meanp=r.meanp) for r in arr]

    def power(self, prn_index: int) -> np.ndarray:
        out = np.zeros(self.nfreq * self.n_eff)
        check(lib.sdrb200_acq_power(self.h, int(prn_index), _p(out, _dp)), "acq_power")
        return out.reshape(self.nfreq, self.n_eff)

    @property
    def cuda_stream(self) -> int:
        lib.sdrb200_acq_cuda_stream.restype = C.c_void_p
        return lib.sdrb200_acq_cuda_stream(self.h) or 0

    def close(self):
        if self.h:
            lib.sdrb200_acq_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ----------------------------------------------------------------- tracking
class Tracking:
    """Batched sdrtracking + cumsumcorr/pll/dll + bit sync for all channels
    (src/sdrtrk.c:15-150, src/sdrnav.c:15-41,198-282, loop of src/sdrmain.c:264-315)."""

    def __init__(self, chans):
        self.nch = len(chans)
        self._keep = chans
        arr = (TrkChan * self.nch)(*chans)
        self.h = C.c_void_p()
        check(lib.sdrb200_trk_create(C.byref(self.h), self.nch, arr), "trk_create")

    def run(self, stream: Stream, nperiods: int, want_records=True):
        nrec = (C.c_int * self.nch)()
        if want_records:
            rec = np.zeros(self.nch * nperiods, dtype=np.dtype(TrkRec))
            check(lib.sdrb200_trk_run(self.h, stream.h, int(nperiods),
                                      rec.ctypes.data_as(C.POINTER(TrkRec)), nrec), "trk_run")
            return rec.reshape(self.nch, nperiods), list(nrec)
        check(lib.sdrb200_trk_run(self.h, stream.h, int(nperiods), None, nrec), "trk_run")
        return None, list(nrec)

    def enqueue(self, stream: Stream, nperiods: int):
        check(lib.sdrb200_trk_enqueue(self.h, stream.h, int(nperiods)), "trk_enqueue")

    def sync(self):
        check(lib.sdrb200_trk_sync(self.h), "trk_sync")

    def reset(self):
        check(lib.sdrb200_trk_reset(self.h), "trk_reset")

    def fetch_obs(self, ch: int) -> dict:
        """the 80-deep interpolation stocks of one channel, newest first (setobsdata, src/sdrtrk.c:160-209)."""
        o = TrkObs()
        check(lib.sdrb200_trk_fetch_obs(self.h, int(ch), C.byref(o)), "trk_fetch_obs")
        return {k: np.array(getattr(o, k)) for k in ("L", "D", "S", "remcout", "codei", "cntout", "codeisum")}

    def set_cluster(self, cluster_size: int, ctrl_works: bool = False):
        """cluster size per channel (0: choose at the next launch, else 1, 2, 4, 8)."""
        check(lib.sdrb200_trk_set_cluster(self.h, int(cluster_size), int(bool(ctrl_works))), "trk_set_cluster")

    @property
    def cuda_stream(self) -> int:
        lib.sdrb200_trk_cuda_stream.restype = C.c_void_p
        return lib.sdrb200_trk_cuda_stream(self.h) or 0

    def close(self):
        if self.h:
            lib.sdrb200_trk_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def trk_log(recs, corrn: int, corrd: int, ctime=1e-3, firstsftow=0.0, flagdec=0) -> str:
    """the reference's tracking log (writelog_header + writelog, src/sdrout.c:386-436) of one channel's
    records: header and one CSV line per code period."""
    buf = C.create_string_buffer(4096)
    n = lib.sdrb200_trk_log_header(int(corrn), int(corrd), buf, C.c_size_t(4096))
    if n < 0:
        raise Error("trk_log_header failed")
    out = [buf.raw[:n].decode()]
    recs = np.ascontiguousarray(recs)
    for i in range(len(recs)):
        n = lib.sdrb200_trk_log_line(recs[i:i + 1].ctypes.data_as(C.POINTER(TrkRec)), int(corrn), C.c_double(firstsftow),
                                     C.c_double(ctime), int(flagdec), buf, C.c_size_t(4096))
        if n < 0:
            raise Error("trk_log_line failed")
        out.append(buf.raw[:n].decode())
    return "".join(out)


def make_trk_chan(f_sf, f_if, f_cf, dtype, prn, code, buffloc, acqfreq, corrn, corrd, corrp,
                  dllb, pllb, fllb, ctype=1, foffset=0.0):
    """Fill a TrkChan the way initsdrch/inittrkstruct/initnavstruct fill sdrch_t
    (src/sdrinit.c:583-657, 432-480, 489-568) for L1CA / SBAS / G1."""
    from . import loop_coeffs
    code = np.ascontiguousarray(code, np.int8)
    clen = len(code)
    crate = 0.511e6 if ctype == 20 else 1.023e6
    ctime = clen / crate
    c = TrkChan()
    c.f_sf, c.f_if, c.f_cf, c.foffset, c.dtype = f_sf, f_if, f_cf, foffset, dtype
    c.clen, c.crate, c.ctime = clen, crate, ctime
    c.nsamp = int(f_sf * ctime)
    c.prn, c.corrn = prn, corrn
    for i in range(corrn):
        c.corrp[i] = corrd * (i + 1)
        if c.corrp[i] == corrp:
            c.ne, c.nl = 2 * (i + 1) - 1, 2 * (i + 1)
    loop = {1: 10, 20: 10, 27: 2}[ctype]                      # LOOP_*, src/sdr.h:152-154
    c.loopms = loop * int(ctime * 1000)
    c.navrate = {1: 20, 20: 10, 27: 2}[ctype]                 # NAVRATE_*, src/sdr.h:159-171
    c.prm1 = loop_coeffs(dllb[0], pllb[0], fllb[0])
    c.prm2 = loop_coeffs(dllb[1], pllb[1], fllb[1])
    c._code_keep = code
    c.code = _p(code, _i8p)
    c.buffloc, c.acqfreq = int(buffloc), float(acqfreq)
    return c
