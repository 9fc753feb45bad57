"""Test and bench scaffolding (NOT product): the synthetic IF generator and two measurement helpers.

CPU side   oracle/libifgen.so        (oracle/ifgen_host.c + tests/testkit/ifgen.h; depends on nothing of the
                                      product, so the reference arm of bench.py maps oracle/ libraries only)
GPU side   tests/testkit/libsdrb200_testkit.so (tests/testkit/testkit.cu): the same generator as a CUDA
                                      kernel, appended to a product IF stream through the public
                                      sdrb200_stream_push_device(); ALU issue peaks; device NCO-walker self-test
``attach(sdr)`` hangs the helpers on the loaded product package under the names the tests use
(sdr.make_ifgen, sdr.ifgen_host, Stream.push_ifgen)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
IFGEN_LIB = os.path.join(ROOT, "oracle", "libifgen.so")
TESTKIT_LIB = os.path.join(ROOT, "tests", "testkit", "libsdrb200_testkit.so")
IFGEN_MAXSAT = 16


class IfgenSat(C.Structure):
    _fields_ = [("prn", C.c_int), ("nav_on", C.c_int), ("cn0_dbhz", C.c_double),
                ("doppler_hz", C.c_double), ("doppler_rate", C.c_double),
                ("code_phase", C.c_double), ("carr_phase", C.c_double),
                ("nav_seed", C.c_uint64), ("carr_offset_hz", C.c_double)]


class IfgenCfg(C.Structure):
    _fields_ = [("f_sf", C.c_double), ("f_if", C.c_double), ("f_cf", C.c_double),
                ("crate", C.c_double), ("sigma", C.c_double), ("dtype", C.c_int),
                ("clen", C.c_int), ("bitlen", C.c_int), ("nsat", C.c_int),
                ("seed", C.c_uint64), ("sat", IfgenSat * IFGEN_MAXSAT),
                ("amp", C.c_double * IFGEN_MAXSAT)]


def make_ifgen(f_sf, f_if, f_cf, dtype, sats, sigma=24.0, seed=0xB2000000,
               crate=1.023e6, clen=1023, bitlen=20) -> IfgenCfg:
    """sats: list of dicts(prn, cn0, doppler, rate, code_phase, carr_phase, nav_on, nav_seed, carr_offset)."""
    c = IfgenCfg()
    c.f_sf, c.f_if, c.f_cf, c.crate, c.sigma = f_sf, f_if, f_cf, crate, sigma
    c.dtype, c.clen, c.bitlen, c.nsat, c.seed = dtype, clen, bitlen, len(sats), seed
    for i, s in enumerate(sats):
        c.sat[i] = IfgenSat(prn=s["prn"], nav_on=s.get("nav_on", 1), cn0_dbhz=s["cn0"],
                            doppler_hz=s["doppler"], doppler_rate=s.get("rate", 0.0),
                            code_phase=s["code_phase"], carr_phase=s.get("carr_phase", 0.0),
                            nav_seed=s.get("nav_seed", 1000 + s["prn"]),
                            carr_offset_hz=s.get("carr_offset", 0.0))
    return c


_cpu = None
_gpu = None


def _cpu_lib():
    global _cpu
    if _cpu is None:
        if not os.path.exists(IFGEN_LIB):
            raise ImportError(f"{IFGEN_LIB} not found: make -C oracle port")
        _cpu = C.CDLL(IFGEN_LIB)
    return _cpu


def _gpu_lib():
    global _gpu
    if _gpu is None:
        if not os.path.exists(TESTKIT_LIB):
            raise ImportError(f"{TESTKIT_LIB} not found: make -C tests/testkit")
        _gpu = C.CDLL(TESTKIT_LIB)
    return _gpu


def ifgen_host(cfg: IfgenCfg, k0: int, n: int) -> np.ndarray:
    """samples [k0, k0 + n) on the CPU (oracle/libifgen.so)."""
    out = np.empty(n * cfg.dtype, np.int8)
    rc = _cpu_lib().oracle_ifgen_host(C.byref(cfg), C.c_uint64(k0), C.c_uint64(n),
                                      out.ctypes.data_as(C.POINTER(C.c_int8)))
    if rc:
        raise RuntimeError(f"oracle_ifgen_host failed ({rc})")
    return out


def push_ifgen(stream, cfg: IfgenCfg, n: int) -> None:
    """append the next n generator samples to a product IF stream (generated on the GPU)."""
    k0 = getattr(stream, "_ifgen_next", 0)
    if cfg.dtype != stream.dtype:
        raise ValueError("ifgen dtype != stream dtype")
    rc = _gpu_lib().sdrb200tk_stream_push_ifgen(stream.h, C.byref(cfg), C.c_uint64(k0), C.c_uint64(n))
    if rc:
        raise RuntimeError(f"sdrb200tk_stream_push_ifgen failed ({rc})")
    stream._ifgen_next = k0 + n


def set_device(dev: int) -> None:
    """the testkit has its own CUDA runtime instance: select the device for its kernels."""
    rc = _gpu_lib().sdrb200tk_set_device(int(dev))
    if rc:
        raise RuntimeError(f"sdrb200tk_set_device failed ({rc})")


def alu_peak(kind: int) -> float:
    """issue peak in operations per second: kind 0 fp32 FMA, 1 int32 multiply-add."""
    v = C.c_double()
    rc = _gpu_lib().sdrb200tk_alu_peak(int(kind), C.byref(v))
    if rc:
        raise RuntimeError(f"sdrb200tk_alu_peak failed ({rc})")
    return v.value


def nco_selftest(n: int, seed: int, mode: int) -> int:
    bad = C.c_int(-1)
    rc = _gpu_lib().sdrb200tk_nco_selftest(int(n), C.c_ulonglong(seed), int(mode), C.byref(bad))
    if rc:
        raise RuntimeError(f"sdrb200tk_nco_selftest failed ({rc})")
    return bad.value


def attach(sdr):
    """the names the tests and the bench use, on the loaded product package."""
    sdr.make_ifgen = make_ifgen
    sdr.ifgen_host = ifgen_host
    sdr.IfgenCfg, sdr.IfgenSat = IfgenCfg, IfgenSat
    sdr.Stream.push_ifgen = push_ifgen
    return sdr
