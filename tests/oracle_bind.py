"""ctypes bindings of the parity harness (oracle/harness.h).  TEST INFRASTRUCTURE:
only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libreplay_ref.so")
DROPIN_SO = os.path.join(ROOT, "oracle", "_ref", "libreplay_b200.so")
PORT_SO = os.path.join(ROOT, "oracle", "liboracle_port.so")
OH_MAXTAPS = 33


class OhCfg(C.Structure):
    _fields_ = [("f_sf", C.c_double), ("f_if", C.c_double), ("f_cf", C.c_double),
                ("dtype", C.c_int), ("corrn", C.c_int), ("corrd", C.c_int), ("corrp", C.c_int),
                ("dllb", C.c_double * 2), ("pllb", C.c_double * 2), ("fllb", C.c_double * 2),
                ("acq_intg", C.c_int), ("acq_hband", C.c_double), ("acq_step", C.c_double)]


class OhAcq(C.Structure):
    _fields_ = [("prn", C.c_int), ("ctype", C.c_int), ("acquired", C.c_int), ("freqi", C.c_int),
                ("codei", C.c_int), ("nintg_used", C.c_int), ("acqfreq", C.c_double),
                ("peakr", C.c_double), ("cn0", C.c_double), ("maxp", C.c_double),
                ("buffloc", C.c_uint64)]


class OhTrkRec(C.Structure):
    _fields_ = [("cnt", C.c_uint64), ("buffloc", C.c_uint64), ("currnsamp", C.c_int),
                ("flagsync", C.c_int), ("flagsyncf", C.c_int), ("flagtow", C.c_int),
                ("flagdec", C.c_int), ("flagloopfilter", C.c_int), ("swsync", C.c_int),
                ("swloop", C.c_int), ("synci", C.c_int), ("biti", C.c_int), ("bit", C.c_int),
                ("II", C.c_double * OH_MAXTAPS), ("QQ", C.c_double * OH_MAXTAPS),
                ("sumI0", C.c_double), ("sumQ0", C.c_double), ("codefreq", C.c_double),
                ("codeErr", C.c_double), ("codeNco", C.c_double), ("carrfreq", C.c_double),
                ("carrErr", C.c_double), ("carrNco", C.c_double), ("freqErr", C.c_double),
                ("remcode", C.c_double), ("remcarr", C.c_double), ("L0", C.c_double),
                ("D0", C.c_double), ("S0", C.c_double), ("remcout0", C.c_double), ("Isum", C.c_double),
                ("obs_codei", C.c_uint64), ("obs_cnt", C.c_uint64), ("flagobs", C.c_int), ("polarity", C.c_int)]


class OhStocks(C.Structure):
    _fields_ = [("L", C.c_double * 80), ("D", C.c_double * 80), ("S", C.c_double * 80), ("remcout", C.c_double * 80),
                ("codei", C.c_uint64 * 80), ("cntout", C.c_uint64 * 80), ("codeisum", C.c_uint64 * 80)]


# the INI defaults of frontend/iffile.ini:29-48 and frontend/rtlsdr_L1.ini:29-48
def cfg_iffile(**kw) -> OhCfg:
    c = OhCfg(f_sf=16.368e6, f_if=4.092e6, f_cf=1575.42e6, dtype=1, corrn=6, corrd=3, corrp=6)
    c.dllb[:] = [5.0, 1.0]; c.pllb[:] = [30.0, 10.0]; c.fllb[:] = [200.0, 50.0]
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def cfg_rtlsdr(**kw) -> OhCfg:
    c = OhCfg(f_sf=2.048e6, f_if=0.0, f_cf=1575.42e6, dtype=2, corrn=4, corrd=1, corrp=1)
    c.dllb[:] = [5.0, 2.0]; c.pllb[:] = [30.0, 20.0]; c.fllb[:] = [200.0, 50.0]
    for k, v in kw.items():
        setattr(c, k, v)
    return c


_i8p = C.POINTER(C.c_int8)
_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_sp = C.POINTER(C.c_short)
_ip = C.POINTER(C.c_int)


class Oracle:
    def __init__(self, path: str):
        self.lib = C.CDLL(path)
        self.lib.oh_impl.restype = C.c_char_p
        self.lib.oh_mixcarr.restype = C.c_double
        self.lib.oh_rescode.restype = C.c_double
        self.name = self.lib.oh_impl().decode()

    def acquire(self, cfg, if_data, avail, prns, ctypes=None, want_power=False):
        if_data = np.ascontiguousarray(if_data, np.int8)
        n = len(prns)
        prn_a = (C.c_int * n)(*prns)
        ct_a = (C.c_int * n)(*(ctypes or [1] * n))
        out = (OhAcq * n)()
        nsamples = len(if_data) // cfg.dtype
        power = None
        pp = None
        if want_power:
            nsamp = int(cfg.f_sf * 1e-3)
            nfreq = 71 if cfg.acq_step <= 0 else 2 * int(cfg.acq_hband / cfg.acq_step) + 1
            power = np.zeros((n, nfreq, nsamp))
            pp = power.ctypes.data_as(_dp)
        rc = self.lib.oh_acquire(C.byref(cfg), if_data.ctypes.data_as(_i8p), C.c_uint64(nsamples),
                                 C.c_uint64(avail), n, prn_a, ct_a, out, pp)
        assert rc == 0, f"oh_acquire rc={rc}"
        return list(out), power

    def set_lookahead(self, on: bool):
        """drop-in build only: sdrb200_sdrh_lookahead of libsdrb200_sdrh.so (1: sdrtracking served from
        the batched engine's records, 0: one correlator call per code period)"""
        lib = C.CDLL(os.path.join(ROOT, "erlangnetwork-gnsslib-sdr_b200", "libsdrb200_sdrh.so"))
        C.c_int.in_dll(lib, "sdrb200_sdrh_lookahead").value = 1 if on else 0

    def lookahead_served(self) -> int:
        lib = C.CDLL(os.path.join(ROOT, "erlangnetwork-gnsslib-sdr_b200", "libsdrb200_sdrh.so"))
        lib.sdrb200_sdrh_lookahead_served.restype = C.c_ulonglong
        return int(lib.sdrb200_sdrh_lookahead_served())

    # ---- observables (reference builds of replay.c only)
    def set_logfile(self, path, chno=1):
        """the reference's own writelog output of channel number chno during the next acq_track"""
        rc = self.lib.oh_set_logfile(path.encode() if path else None, int(chno))
        assert rc == 0, "oh_set_logfile"

    def last_stocks(self, i):
        """stocks of channel index i after the last acq_track: dict of 80-element arrays, newest first"""
        o = OhStocks()
        assert self.lib.oh_last_stocks(int(i), C.byref(o)) == 0
        return {k: np.array(getattr(o, k)) for k in ("L", "D", "S", "remcout", "codei", "cntout", "codeisum")}

    def interp_obs(self, stocks, sampbase, samprefd):
        """syncthread's interpolation (src/sdrsync.c:96-121) with the reference's own interp1"""
        o = OhStocks()
        for k in ("L", "D", "S", "remcout", "codei", "cntout", "codeisum"):
            arr = getattr(o, k)
            for j, v in enumerate(stocks[k]):
                arr[j] = int(v) if k in ("codei", "cntout", "codeisum") else float(v)
        L, D = C.c_double(), C.c_double()
        assert self.lib.oh_interp_obs(C.byref(o), C.c_uint64(int(sampbase)), C.c_double(samprefd), C.byref(L), C.byref(D)) == 0
        return L.value, D.value

    def fft_seconds(self, reset=True):
        """seconds spent inside the FFT shim's fftwf_execute_dft in this process (reference builds
        only; None for the port, which calls its FFT directly)."""
        f = getattr(self.lib, "oracle_fft_seconds", None)
        if f is None:
            return None
        f.restype = C.c_double
        return float(f(1 if reset else 0))

    def acq_track(self, cfg, if_data, avail, prns, nperiods, ctypes=None, threaded=False):
        """threaded=True: one pthread per channel (oh_acq_track_mt, replay builds only)"""
        if_data = np.ascontiguousarray(if_data, np.int8)
        n = len(prns)
        prn_a = (C.c_int * n)(*prns)
        ct_a = (C.c_int * n)(*(ctypes or [1] * n))
        acq = (OhAcq * n)()
        rec = np.zeros(n * nperiods, dtype=np.dtype(OhTrkRec))
        ntr = (C.c_int * n)()
        nsamples = len(if_data) // cfg.dtype
        fn = self.lib.oh_acq_track_mt if threaded else self.lib.oh_acq_track
        rc = fn(C.byref(cfg), if_data.ctypes.data_as(_i8p), C.c_uint64(nsamples),
                C.c_uint64(avail), n, prn_a, ct_a, int(nperiods), acq,
                rec.ctypes.data_as(C.POINTER(OhTrkRec)), ntr)
        assert rc == 0, f"oh_acq_track rc={rc}"
        return list(acq), rec.reshape(n, nperiods), list(ntr)

    def correlator(self, data, dtype, ti, n, freq, phi0, crate, coff, s, code):
        data = np.ascontiguousarray(data, np.int8)
        s = np.ascontiguousarray(s, np.int32)
        code = np.ascontiguousarray(code, np.int16)
        nt = 1 + 2 * len(s)
        II, QQ = np.zeros(nt), np.zeros(nt)
        remc, remp = C.c_double(), C.c_double()
        self.lib.oh_correlator(data.ctypes.data_as(_i8p), int(dtype), C.c_double(ti), int(n),
                               C.c_double(freq), C.c_double(phi0), C.c_double(crate), C.c_double(coff),
                               s.ctypes.data_as(_ip), len(s), II.ctypes.data_as(_dp), QQ.ctypes.data_as(_dp),
                               C.byref(remc), C.byref(remp), code.ctypes.data_as(_sp), len(code))
        return II, QQ, remc.value, remp.value

    def mixcarr(self, data, dtype, ti, n, freq, phi0):
        data = np.ascontiguousarray(data, np.int8)
        II, QQ = np.zeros(n + 64, np.int16), np.zeros(n + 64, np.int16)
        rem = self.lib.oh_mixcarr(data.ctypes.data_as(_i8p), int(dtype), C.c_double(ti), int(n),
                                  C.c_double(freq), C.c_double(phi0), II.ctypes.data_as(_sp),
                                  QQ.ctypes.data_as(_sp))
        return II[:n], QQ[:n], rem

    def rescode(self, code, coff, smax, ci, n):
        code = np.ascontiguousarray(code, np.int16)
        out = np.zeros(n + 2 * smax, np.int16)
        rem = self.lib.oh_rescode(code.ctypes.data_as(_sp), len(code), C.c_double(coff), int(smax),
                                  C.c_double(ci), int(n), out.ctypes.data_as(_sp))
        return out, rem

    def unpack(self, fmt, raw):
        """front-end byte unpack: 'rtlsdr', 'stereo1', 'stereo2', 'gn3s2', 'gn3s4' -> int8 samples.
        The reference build calls the reference's own rtlsdr_exp / stereo_exp / gn3s_exp_v3."""
        raw = np.ascontiguousarray(raw, np.uint8)
        n = len(raw)
        two = fmt in ("stereo2", "gn3s4")
        out = np.zeros(n * (2 if two else 1), np.int8)
        u8p, i8p = raw.ctypes.data_as(C.POINTER(C.c_uint8)), out.ctypes.data_as(_i8p)
        ref = self.name == "reference"
        if fmt == "rtlsdr":
            (self.lib.rtlsdr_exp if ref else self.lib.port_rtlsdr_exp)(u8p, n, i8p)
        elif fmt in ("stereo1", "stereo2"):
            (self.lib.stereo_exp if ref else self.lib.port_stereo_exp)(u8p, n, 2 if two else 1, i8p)
        else:
            f = getattr(self.lib, "_Z11gn3s_exp_v3PhiiPc") if ref else self.lib.port_gn3s_exp_v3
            f(u8p, n, 136 if two else 132, i8p)
        return out

    def gencode(self, prn, ctype=1):
        buf = np.zeros(10230, np.int16)
        cr = C.c_double()
        n = self.lib.oh_gencode(int(prn), int(ctype), buf.ctypes.data_as(_sp), C.byref(cr))
        if n <= 0:
            return None, 0.0
        return buf[:n].copy(), cr.value

    def framesync(self, ctype, prn, fbits):
        """(found, polarity) of the frame synchronisation on one nav.fbits register: the reference's
        findpreamble() after its own initnavstruct()/predecodefec() (oracle/_ref), or the port's
        GLONASS string sync (ctype 20, 200 units)."""
        fb = np.ascontiguousarray(fbits, np.int32)
        pol = C.c_int(0)
        ip = C.POINTER(C.c_int)
        if hasattr(self.lib, "oh_ref_framesync"):
            r = self.lib.oh_ref_framesync(int(ctype), int(prn), fb.ctypes.data_as(ip), int(len(fb)), C.byref(pol))
        else:
            assert ctype == 20 and len(fb) == 200
            r = self.lib.oh_port_framesync_g1(fb.ctypes.data_as(ip), C.byref(pol))
        return int(r), pol.value

    def codefft(self, prn, ctype, f_sf):
        nsamp, nfft = C.c_int(), C.c_int()
        self.lib.oh_codefft(int(prn), int(ctype), C.c_double(f_sf), None, C.byref(nsamp), C.byref(nfft))
        x = np.zeros(nfft.value, np.complex64)
        self.lib.oh_codefft(int(prn), int(ctype), C.c_double(f_sf), x.ctypes.data_as(_fp),
                            C.byref(nsamp), C.byref(nfft))
        return x, nsamp.value, nfft.value

    def cpxfft(self, cpx, inverse=False):
        a = np.ascontiguousarray(cpx, np.complex64).copy()
        self.lib.oh_cpxfft(a.ctypes.data_as(_fp), len(a), int(bool(inverse)))
        return a

    def cpxconv(self, cpxa, cpxb, n, conv=None):
        a = np.ascontiguousarray(cpxa, np.complex64).copy()
        b = np.ascontiguousarray(cpxb, np.complex64)
        flagsum = conv is not None
        out = np.ascontiguousarray(conv, np.float64).copy() if flagsum else np.zeros(n)
        self.lib.oh_cpxconv(a.ctypes.data_as(_fp), b.ctypes.data_as(_fp), len(a), int(n), int(flagsum),
                            out.ctypes.data_as(_dp))
        return out

    def cpxpspec(self, cpx, pspec=None):
        a = np.ascontiguousarray(cpx, np.complex64).copy()
        flagsum = pspec is not None
        out = np.ascontiguousarray(pspec, np.float64).copy() if flagsum else np.zeros(len(a))
        self.lib.oh_cpxpspec(a.ctypes.data_as(_fp), len(a), int(flagsum), out.ctypes.data_as(_dp))
        return out

    def pcorrelator(self, data, dtype, ti, n, freq, crate, m, codex, P=None):
        data = np.ascontiguousarray(data, np.int8)
        freq = np.ascontiguousarray(freq, np.float64)
        codex = np.ascontiguousarray(codex, np.complex64)
        out = np.zeros(len(freq) * n) if P is None else np.ascontiguousarray(P, np.float64).copy()
        self.lib.oh_pcorrelator(data.ctypes.data_as(_i8p), int(dtype), C.c_double(ti), int(n),
                                freq.ctypes.data_as(_dp), len(freq), C.c_double(crate), int(m),
                                codex.ctypes.data_as(_fp), out.ctypes.data_as(_dp))
        return out


def have_ref() -> bool:
    return os.path.exists(REF_SO)


REF_SHIPPED_SO = os.path.join(ROOT, "oracle", "_ref", "libreplay_ref_shipped.so")


def have_ref_shipped() -> bool:
    return os.path.exists(REF_SHIPPED_SO)


def ref_shipped() -> Oracle:
    """the reference compiled with its own bin/Makefile flags (-g, no -O, -DFFTMTX)"""
    return Oracle(REF_SHIPPED_SO)


def have_port() -> bool:
    return os.path.exists(PORT_SO)


def ref() -> Oracle:
    return Oracle(REF_SO)


def port() -> Oracle:
    return Oracle(PORT_SO)


def have_dropin() -> bool:
    return os.path.exists(DROPIN_SO)


def dropin() -> Oracle:
    """the reference's cold objects + replay.c with every hot sdr.h symbol bound to
    libsdrb200_sdrh.so (needs a GPU at call time)"""
    return Oracle(DROPIN_SO)


def best() -> Oracle:
    """the reference build when present (this container), else the CPU port."""
    return ref() if have_ref() else port()
