#!/usr/bin/env python
"""bench.py - throughput of the two GNSS-SDRLIB correlation hot paths on B200.

Headline (top-level keys): batched acquisition, BASELINE.json configs[0] shape -
GPS L1CA, 32 PRNs x 71 Doppler bins x 16368 code phases, real int8 IF at
16.368 MS/s, one 1 ms integration (sdracquisition/pcorrelator/checkacquisition,
src/sdracq.c:14-95, src/sdrcmn.c:738-773).  metric = cells/s, cell = one
(PRN, Doppler bin, code phase) power value.
Second half of the metric ("tracking" object): BASELINE.json configs[1] -
12 L1CA channels, E/P/L(+taps) correlators + DLL/PLL/FLL + bit sync over
`--trk-seconds` (default 30) of synthetic IF, channel.s/s (src/sdrtrk.c:15-150).

One step of the headline = one acquisition pass over all PRNs of one IF window.
Multi-GPU (torchrun, one rank per GPU): every rank searches its OWN IF stream
(weak scaling, no data-path collective); the per-rank peak tuples are gathered
with one NCCL all_gather, which is part of the e2e region.

  python bench.py [--gpus N] [--steps K] [--warmup W]
  python bench.py --impl reference ...     # the reference's CPU code (oracle/_ref)
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

FS, F_IF, F_CF, NSAMP, NFREQ, NPRN = 16.368e6, 4.092e6, 1575.42e6, 16368, 71, 32
CELLS_PER_STEP = NPRN * NFREQ * NSAMP
AVAIL = 65536                                   # samples "in the ring" at acquisition time
AVAIL10 = 3 * 65536                             # ... for the 10-integration hand-over acquisition
TRK_PRNS = [1, 3, 6, 9, 11, 14, 17, 19, 22, 25, 28, 31]
WORKLOAD_ACQ = ("GPS L1CA acquisition, 32 PRN x 71 Doppler bins (+-7 kHz/200 Hz) x 16368 code phases, "
                "real int8 IF 16.368 MS/s, 1 ms, intg=1 (BASELINE configs[0])")
WORKLOAD_TRK = ("GPS L1CA tracking, 12 channels, 13 taps (CORRN 6/CORRD 3/CORRP 6), DLL/PLL/FLL + bit sync, "
                "real int8 IF 16.368 MS/s (BASELINE configs[1])")


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def scene(sdr, seed):
    """12 visible PRNs at 55-58 dB-Hz, Doppler within 30 Hz of a search bin (so that the 1 ms acquisition picks the right bin and every channel locks) with Doppler, Doppler rate and 50 bps data."""
    import numpy as np
    rng = np.random.default_rng(seed)
    sats = []
    for p in TRK_PRNS:
        b = int(rng.integers(-20, 21))
        sats.append(dict(prn=p, cn0=float(rng.uniform(55, 58)), doppler=b * 200.0 + float(rng.uniform(-30, 30)),
                         code_phase=float(rng.uniform(0, 1023 * 20)), carr_phase=float(rng.uniform(0, 1)),
                         rate=float(rng.uniform(-1, 1))))
    return sdr.make_ifgen(FS, F_IF, F_CF, 1, sats, sigma=24.0, seed=seed)


# ------------------------------------------------------------------ CPU legs
def _cpu_acq_worker(job):
    """one process: the reference's sdracquisition for a list of PRNs."""
    seed, prns = job
    import oracle_bind as ob
    host = cached_if(seed, AVAIL)                         # made before the clock starts (pool initializer)
    o = ob.best()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)                                   # the reference printf()s per channel
    try:
        o.fft_seconds(reset=True)
        t0 = time.perf_counter()
        res, _ = o.acquire(ob.cfg_iffile(acq_intg=1), host, AVAIL, list(prns))
        dt = time.perf_counter() - t0
        fft_s = o.fft_seconds(reset=True)
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    with open("/proc/self/maps") as f:
        mapped = "libsdrb200" in f.read()              # must stay False: the CPU legs run on oracle/ libraries only
    return dt, [(r.prn, r.acquired, r.freqi, r.codei) for r in res], o.name, fft_s, mapped


def _cpu_trk_worker(job):
    seed, prn, nper = job
    import oracle_bind as ob
    total = AVAIL10 + (nper + 4) * NSAMP
    host = cached_if(seed, total)
    o = ob.best()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        cfg = ob.cfg_iffile(acq_intg=10)                     # ACQINTG_L1CA, src/sdr.h:140
        t0 = time.perf_counter()
        o.acquire(cfg, host, AVAIL10, [prn])
        t_acq = time.perf_counter() - t0
        t0 = time.perf_counter()
        acq, rec, ntr = o.acq_track(cfg, host, AVAIL10, [prn], nper)
        dt = time.perf_counter() - t0 - t_acq
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    n = ntr[0]
    r = rec[0][:n]
    keep = {"prn": prn, "acquired": acq[0].acquired, "freqi": acq[0].freqi, "codei": acq[0].codei, "n": n,
            "currnsamp": r["currnsamp"].copy(), "II": r["II"][:, :13].copy(), "QQ": r["QQ"][:, :13].copy(),
            "flagsync": r["flagsync"].copy(), "synci": r["synci"].copy(), "bit": r["bit"].copy(),
            "swsync": r["swsync"].copy(), "carrfreq": r["carrfreq"].copy(), "codefreq": r["codefreq"].copy()}
    return dt, n, o.name, keep


def _cpu_shipped_worker(job):
    """the reference compiled as its own bin/Makefile does (-g, no -O, -DFFTMTX) and run with its
    own threading model: acquisition by one thread (the channel threads' FFTs are serialised by
    hfftmtx anyway), tracking with one pthread per channel (oh_acq_track_mt)."""
    seed, nacq, nper = job
    import oracle_bind as ob
    if not ob.have_ref_shipped():
        return None
    o = ob.ref_shipped()
    total = AVAIL10 + (nper + 4) * NSAMP
    host = cached_if(seed, total)
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        t0 = time.perf_counter()
        o.acquire(ob.cfg_iffile(acq_intg=1), host[:AVAIL], AVAIL, list(range(1, nacq + 1)))
        t_acq = time.perf_counter() - t0
        cfg = ob.cfg_iffile(acq_intg=10)
        t0 = time.perf_counter()
        o.acq_track(cfg, host, AVAIL10, TRK_PRNS, 1, threaded=True)
        t_1 = time.perf_counter() - t0
        t0 = time.perf_counter()
        _, _, ntr = o.acq_track(cfg, host, AVAIL10, TRK_PRNS, nper, threaded=True)
        t_trk = time.perf_counter() - t0 - t_1
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    return {"acq_cells_per_s": nacq * NFREQ * NSAMP / t_acq, "acq_sample": f"{nacq} PRNs x 71 bins x 16368, one thread",
            "trk_channel_s_per_s": sum(ntr) * 1e-3 / t_trk,
            "trk_sample": f"{len(TRK_PRNS)} channels x {nper} code periods, one pthread per channel",
            "flags": "-g -DFFTMTX, no -O (bin/Makefile:17,29); FFT shim at -O2 in place of the FFTW binary"}


_IF_CACHE = {}


def cached_if(seed, n):
    """host IF of the bench scene, samples [0, n): generated once per process by the CPU generator
    (oracle/libifgen.so - the CPU legs never map the product library) and kept."""
    import testkit
    have = _IF_CACHE.get(seed)
    if have is None or len(have) < n:
        have = testkit.ifgen_host(scene(testkit, seed), 0, n)
        _IF_CACHE[seed] = have
    return have[:n]


def _pool_init(seed, n):
    cached_if(seed, n)


def cpu_pool(cores, seed=None, n=0):
    """worker processes; with a seed every worker generates its copy of the IF while it starts,
    i.e. before any clock of the CPU legs runs."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    if seed is None:
        return ctx.Pool(cores)
    return ctx.Pool(cores, initializer=_pool_init, initargs=(seed, n))


def cpu_acq_step(pool, cores, seed, prns):
    """one acquisition pass over `prns` spread over `cores` processes; wall seconds."""
    chunks = [prns[i::cores] for i in range(cores) if prns[i::cores]]
    t0 = time.perf_counter()
    out = pool.map(_cpu_acq_worker, [(seed, c) for c in chunks])
    return time.perf_counter() - t0, out


def oracle_kind():
    import oracle_bind as ob
    return "reference" if ob.have_ref() else "port"


def oracle_label():
    return ("reference C sources (gcc -O2) + oracle FFT shim, not FFTW" if oracle_kind() == "reference"
            else "oracle/port CPU restatement (gcc -O2) + oracle FFT")


# --------------------------------------------------------------- clocks
class ClockSampler:
    def __init__(self, dev):
        self.rows, self.proc, self.dev = [], None, dev
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(dev), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        self.names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        self.mark0 = 0

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def mark(self):
        self.mark0 = len(self.rows)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = self.rows[self.mark0:] or self.rows
        sm, mx, reasons = [], 0, set()
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0])); mx = max(mx, float(f[1]))
            except Exception:
                continue
            for name, v in zip(self.names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ GPU arm
class Evt:
    def __init__(self, sdr):
        self.sdr, self.h = sdr, C.c_void_p()
        sdr.check(sdr.lib.sdrb200_evt_create(C.byref(self.h)), "evt_create")

    def record(self, stream):
        self.sdr.check(self.sdr.lib.sdrb200_evt_record(self.h, C.c_void_p(stream)), "evt_record")

    def ms_since(self, start):
        ms = C.c_float()
        self.sdr.check(self.sdr.lib.sdrb200_evt_elapsed_ms(start.h, self.h, C.byref(ms)), "evt_elapsed")
        return float(ms.value)


def run_b200(args):
    import numpy as np
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    hbm_peak, peak_src = peaks()

    # CPU baseline first (rank 0, N=1 only), before CUDA is initialised in this process
    cpu_acq = cpu_trk = None
    ref_acq = ref_trk = host_par = None                  # what the GPU path is checked against (outside all clocks)
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        pool = cpu_pool(cores, 0xB2000000, AVAIL10 + 2004 * NSAMP)
        try:
            prns = list(range(1, NPRN + 1))
            cpu_acq_step(pool, cores, 0xB2000000, prns[:cores])         # warm (page in, FFT plans)
            dt, out = cpu_acq_step(pool, cores, 0xB2000000, prns)
            ref_acq = {t[0]: t[1:] for w in out for t in w[1]}          # prn -> (acquired, freqi, codei)
            fft_s = [w[3] for w in out if w[3] is not None]
            cpu_acq = {"value": CELLS_PER_STEP / dt, "unit": "cells/s", "cores": min(cores, NPRN),
                       "kind": oracle_kind(), "impl": oracle_label(), "fft": "shim",
                       "fftw3f_on_this_host": __import__("ctypes.util").util.find_library("fftw3f"),
                       "fft_share_of_cpu_time": (sum(fft_s) / sum(w[0] for w in out)) if fft_s else None,
                       "sample": f"1 full step (32 PRN x 71 bins x 16368), {dt:.2f} s wall on {min(cores, NPRN)} processes"}
            nper = 2000
            t0 = time.perf_counter()
            res = pool.map(_cpu_trk_worker, [(0xB2000000, p, nper) for p in TRK_PRNS])
            wall = time.perf_counter() - t0
            tot = sum(r[1] for r in res)
            per_core = sum(r[0] for r in res)
            ref_trk = [r[3] for r in res]
            host_par = pool.apply(cached_if, (0xB2000000, AVAIL10 + (nper + 4) * NSAMP))   # the bytes the CPU legs ran on
            cpu_trk = {"value": tot * 1e-3 / max(r[0] for r in res), "unit": "channel.s/s",
                       "cores": min(cores, len(TRK_PRNS)), "kind": oracle_kind(), "impl": oracle_label(),
                       "single_core_value": tot * 1e-3 / per_core,
                       "sample": f"12 channels x {nper} code periods (2 s of IF), one process per channel, "
                                 f"tracking loop only ({wall:.1f} s wall incl. acquisition)"}
            shipped = pool.apply(_cpu_shipped_worker, ((0xB2000000, 4, 600),))
            if shipped:
                cpu_acq["as_shipped"] = {"value": shipped["acq_cells_per_s"], "unit": "cells/s", "cores": 1,
                                         "sample": shipped["acq_sample"], "flags": shipped["flags"]}
                cpu_trk["as_shipped"] = {"value": shipped["trk_channel_s_per_s"], "unit": "channel.s/s",
                                         "cores": min(cores, len(TRK_PRNS)), "sample": shipped["trk_sample"],
                                         "flags": shipped["flags"]}
        finally:
            pool.close()

    import sdrb200_pkg
    import testkit
    sdr = testkit.attach(sdrb200_pkg.load())
    lib = sdr.lib
    if sdr.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device (libsdrb200 has no CPU fallback)")
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sdr.check(lib.sdrb200_set_device(local), "set_device")
    testkit.set_device(local)

    def barrier():
        if dist is not None:
            dist.barrier()
        sdr.check(lib.sdrb200_device_sync(), "sync")

    # ALU issue peaks of this GPU (micro-benchmark in the library), for the second roofline
    fp32_peak = testkit.alu_peak(0) / 1e12
    int32_peak = testkit.alu_peak(1) / 1e12

    # ------------------------------------------------------------ data
    seed = 0xB2000000 + rank
    cfg = scene(sdr, seed)
    trk_T = args.trk_seconds
    total = int(trk_T * FS) + 2 * AVAIL
    st = sdr.Stream(1, total)
    st.push_ifgen(cfg, total)
    prns = list(range(1, NPRN + 1))
    codes = np.stack([sdr.gencode(p)[0] for p in prns])
    freq = np.array([F_IF + (i - (NFREQ - 1) // 2) * 200.0 for i in range(NFREQ)])
    acq = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq, codes, 1.023e6, intg=1)
    b0 = AVAIL - 2 * NSAMP                                   # src/sdracq.c:24-26
    s_acq = acq.cuda_stream
    K, W = args.steps, args.warmup

    # one nvidia-smi poller per box (rank 0 watches its own GPU): eight of them polling every 100 ms
    # perturb the host-side legs of the other ranks
    clocks = ClockSampler(local) if rank == 0 else None
    launches0 = sdr.launch_count()
    # ------------------------------------------------ acquisition, HBM-resident
    for _ in range(W):
        acq.enqueue(st, b0)
    e0, e1 = Evt(sdr), Evt(sdr)
    ms4, c4 = (C.c_double * 4)(), (C.c_longlong * 4)()
    barrier()
    if clocks: clocks.mark()
    launches_t0 = sdr.launch_count()
    step_ms = []
    for _ in range(K):
        sdr.check(lib.sdrb200_l2_flush(C.c_void_p(s_acq)), "l2_flush")
        e0.record(s_acq)
        acq.enqueue(st, b0)
        e1.record(s_acq)
        step_ms.append(e1.ms_since(e0))
    barrier()
    launches_timed = sdr.launch_count() - launches_t0
    res = acq.fetch()
    t_acq = sum(step_ms) * 1e-3
    # per-kernel durations: the same K steps once more with a CUDA-event pair around every
    # launch on the plan's stream (the pairs cost a few % of a step, so they stay out of `value`)
    sdr.check(lib.sdrb200_acq_set_timing(acq.h, 1), "set_timing")
    for _ in range(K):
        sdr.check(lib.sdrb200_l2_flush(C.c_void_p(s_acq)), "l2_flush")
        acq.enqueue(st, b0)
        sdr.check(lib.sdrb200_acq_kernel_ms(acq.h, ms4, c4), "kernel_ms")
    barrier()
    sdr.check(lib.sdrb200_acq_set_timing(acq.h, 0), "set_timing")
    KNAMES = ["k_fwd_split_small+k_fft_rows", "k_rowc<search>", "k_comb", "k_rowc<peak>"]
    kern = {n: {"ms_total": ms4[i], "launches": int(c4[i]),
                "ms_per_launch": (ms4[i] / c4[i]) if c4[i] else None}
            for i, n in enumerate(KNAMES) if c4[i]}

    # ------------------------------------------------ acquisition, end to end (host buffers)
    import torch
    need = 2 * NSAMP
    host_pin = torch.empty(need, dtype=torch.int8, pin_memory=True)
    host_np = host_pin.numpy()
    host_np[:] = st.read(b0, need)
    # result tuples land in pinned host memory; at N > 1 they go straight back to the GPU
    # for one NCCL all-gather of the per-shard tuples (north_star: the only collective)
    rbytes = NPRN * C.sizeof(sdr.AcqResult)
    res_pin = torch.empty(rbytes, dtype=torch.uint8, pin_memory=True)
    res_ptr = C.cast(res_pin.data_ptr(), C.POINTER(sdr.AcqResult))
    host_ptr = C.cast(host_pin.data_ptr(), C.POINTER(C.c_int8))
    if dist is not None:
        # the per-step tuples also stay in a device ring [K][nprn]; ONE all-gather per K steps takes
        # them from device memory (round 1 bounced every step's tuples host -> device and gathered
        # them per 0.33 ms step, which cost 25 % at N = 8)
        res_ring = torch.empty((max(K, 3), rbytes), dtype=torch.uint8, device="cuda")
        res_all = torch.empty((world, max(K, 3), rbytes), dtype=torch.uint8, device="cuda")

    def e2e_steps(nsteps):
        for i in range(nsteps):
            sdr.check(lib.sdrb200_acq_run_host(acq.h, host_ptr, C.c_uint64(need), res_ptr), "acq_run_host")
            if dist is not None:
                acq.copy_results_device(res_ring[i].data_ptr())
        if dist is not None:
            acq.sync()
            dist.all_gather_into_tensor(res_all.view(-1), res_ring.view(-1))
            torch.cuda.synchronize()

    e2e_steps(3)
    barrier()
    t0 = time.perf_counter()
    e2e_steps(K)
    barrier()
    t_e2e_single = time.perf_counter() - t0
    # the same work with TWO plans in flight (sdrb200_acq_enqueue_host / sdrb200_acq_fetch): every step still
    # copies its own input from pinned host memory and reads its own tuples back, but the copy and the launches
    # of step i+1 are queued behind step i instead of waiting for its results - the batched form of the
    # reference's one acquisition thread per channel
    acq_b = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq, codes, 1.023e6, intg=1)
    host_pin_b = torch.empty(need, dtype=torch.int8, pin_memory=True)
    host_pin_b.numpy()[:] = host_np
    host_ptr_b = C.cast(host_pin_b.data_ptr(), C.POINTER(C.c_int8))
    res_b = (sdr.AcqResult * NPRN)()
    plans2 = [(acq, host_ptr, res_ptr), (acq_b, host_ptr_b, res_b)]

    def e2e_steps_2(nsteps):
        for i in range(min(2, nsteps)):
            pl, hp, _ = plans2[i & 1]
            sdr.check(lib.sdrb200_acq_enqueue_host(pl.h, hp, C.c_uint64(need)), "acq_enqueue_host")
        for i in range(nsteps):
            pl, hp, rp = plans2[i & 1]
            sdr.check(lib.sdrb200_acq_fetch(pl.h, rp), "acq_fetch")
            if dist is not None:
                pl.copy_results_device(res_ring[i].data_ptr())
            if i + 2 < nsteps:
                sdr.check(lib.sdrb200_acq_enqueue_host(pl.h, hp, C.c_uint64(need)), "acq_enqueue_host")
        if dist is not None:
            acq.sync(); acq_b.sync()
            dist.all_gather_into_tensor(res_all.view(-1), res_ring.view(-1))
            torch.cuda.synchronize()

    e2e_steps_2(4)
    barrier()
    t0 = time.perf_counter()
    e2e_steps_2(K)
    barrier()
    t_e2e = time.perf_counter() - t0
    rb = acq._results(res_b)
    assert [(r["acquired"], r["freqi"], r["codei"]) for r in rb] == \
           [(r["acquired"], r["freqi"], r["codei"]) for r in res], "two-plan host-buffer path disagrees with resident path"
    acq_b.close()
    r2 = acq._results((sdr.AcqResult * NPRN).from_buffer_copy(bytes(res_pin.numpy())))
    if dist is not None:
        mine = res_all[rank, K - 1].cpu().numpy()
        assert bytes(mine) == bytes(res_pin.numpy()), "gathered tuples differ from the local shard"
    assert [(r["acquired"], r["freqi"], r["codei"]) for r in r2] == \
           [(r["acquired"], r["freqi"], r["codei"]) for r in res], "host-buffer path disagrees with resident path"

    # ------------------------------------------------ tracking
    # hand-over as the reference does it: acquisition with the default 10 non-coherent
    # integrations and early exit (ACQINTG_L1CA, src/sdr.h:140; src/sdracq.c:29-43) on the
    # window that ends at AVAIL10, then tracking starts at the acquired code phase
    tcodes = np.stack([sdr.gencode(p)[0] for p in TRK_PRNS])
    acq10 = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq, tcodes, 1.023e6, intg=10)
    b10 = AVAIL10 - 11 * NSAMP                               # src/sdracq.c:24-26
    got = {p: r for p, r in zip(TRK_PRNS, acq10.run(st, b10))}
    chans = [sdr.make_trk_chan(FS, F_IF, F_CF, 1, p, sdr.gencode(p)[0],
                               b10 + got[p]["codei"],                      # src/sdracq.c:53
                               got[p]["acqfreq"], 6, 3, 6, (5.0, 1.0), (30.0, 10.0), (200.0, 50.0))
             for p in TRK_PRNS if got[p]["acquired"]]
    nch = len(chans)
    trk_obj = None
    if nch and not args.no_trk:
        trk = sdr.Tracking(chans)
        nper = int(trk_T * 1000) - 8
        s_trk = trk.cuda_stream
        trk_steps, trk_warm = max(1, min(args.trk_steps, K)), 1
        for _ in range(trk_warm):
            trk.reset(); trk.enqueue(st, nper); trk.sync()
        barrier()
        tms = []
        for _ in range(trk_steps):
            trk.reset()
            e0.record(s_trk); trk.enqueue(st, nper); e1.record(s_trk)
            tms.append(e1.ms_since(e0))
        barrier()
        trk.reset()
        _, nrec = trk.run(st, nper, want_records=False)
        done = sum(nrec)
        phases = None
        if args.trk_profile:
            # cycle counters of channel 0, outside every timed region (sdrb200_trk_profile)
            out = (C.c_longlong * 32)()
            sdr.check(lib.sdrb200_trk_profile(trk.h, 1, None), "trk_profile")
            trk.reset(); trk.enqueue(st, nper); trk.sync()
            sdr.check(lib.sdrb200_trk_profile(trk.h, 0, out), "trk_profile")
            per = max(1, int(out[6]))
            names = {0: "code_lane_table_and_walk", 1: "code_lane_waits_and_decisions", 2: "carrier_lane_table_and_walk",
                     3: "carrier_lane_2pi_reduction", 4: "carrier_lane_waits", 11: "carrier_lane_push",
                     5: "epilogue_warp_waiting_for_results", 12: "epilogue_nav", 13: "epilogue_loops", 14: "epilogue_obs_rec",
                     7: "team0_dot_products", 8: "team0_report", 9: "team0_waiting_for_hand_out"}
            phases = {v: round(out[k] / per, 1) for k, v in names.items()}
            tp = max(1, int(out[31]))
            sub = {26: "loads_wait_carrier_table", 16: "carrier_words", 27: "wait_code_table", 21: "code_bits", 22: "bar_code_bits",
                   25: "tap_dot_products", 24: "bar_partials"}
            phases["team0_per_own_period"] = {v: round(out[k] / tp, 1) for k, v in sub.items()}
            phases["team0_per_own_period"].update(period_total=round(out[7] / tp, 1), report=round(out[8] / tp, 1),
                                                  waiting=round(out[9] / tp, 1), periods=tp)
            phases.update(periods=per, quiet_periods=int(out[10]),
                          unit="cycles per period of channel 0 (team0_* per period of the channel, the team sees 1 in W)")
        t_trk = sum(tms) / len(tms) * 1e-3
        # channel-count sweep: the same 12 hand-over states replicated over more channels of the same
        # IF stream (every channel is still a full closed loop of its own), 4 s each
        sweep = []
        if not args.no_sweep:
            per_sw = 4000
            for nchs in (12, 64, 148, 512, 2048):
                eng = sdr.Tracking([chans[i % nch] for i in range(nchs)])
                s_sw = eng.cuda_stream
                eng.enqueue(st, 100); eng.sync(); eng.reset()
                barrier()
                e0.record(s_sw); eng.enqueue(st, per_sw); e1.record(s_sw)
                eng.sync()
                ms_sw = e1.ms_since(e0)
                _, nrec_sw = eng.run(st, 1, want_records=False)
                eng.close()
                b_sw = nchs * per_sw * (NSAMP + 704)
                sweep.append({"channels": nchs, "channel_s_per_s": nchs * per_sw * 1e-3 / (ms_sw * 1e-3),
                              "us_per_code_period_per_channel": ms_sw * 1e3 / per_sw,
                              "hbm_frac": b_sw / (ms_sw * 1e-3) / 1e9 / hbm_peak})
        # end to end: IF from pinned host memory -> device ring -> tracking -> per-channel record counts
        nbytes = int(nper + 8) * NSAMP + AVAIL
        host_if = torch.empty(nbytes, dtype=torch.int8, pin_memory=True)
        host_if.numpy()[:] = st.read(0, nbytes)
        st2 = sdr.Stream(1, nbytes)
        barrier()
        t0 = time.perf_counter()
        st2.push(host_if.numpy())
        trk.reset()
        _, nrec2 = trk.run(st2, nper, want_records=False)
        barrier()
        t_trk_e2e = time.perf_counter() - t0
        st2.close()
        a_bytes = done * (NSAMP + 704)                          # per launch: one launch per step
        trk_tmax = t_trk
        if dist is not None:
            tt = torch.tensor([t_trk, t_trk_e2e], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            trk_tmax, t_trk_e2e = float(tt[0]), float(tt[1])
            tn = torch.tensor([float(done), float(sum(nrec2))], dtype=torch.float64, device="cuda")
            dist.all_reduce(tn)
            done_all, done2_all = float(tn[0]), float(tn[1])
        else:
            done_all, done2_all = float(done), float(sum(nrec2))
        # DRAM bytes of one k_trk_cl launch: the ncu capture is a 2 s launch (profiles/r2d_ncu_full_trk_summary.csv:
        # the 12 channels read ONE IF stream, which comes from DRAM once), scaled to this launch's IF seconds
        trk_traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                per_s = json.load(f).get("k_trk_cl_per_if_second")
            if per_s:
                trk_traffic = int(per_s * args.trk_seconds)
        except Exception:
            pass
        trk_obj = {
            "metric": "tracking channel.s/s (channels x seconds of IF per wall second)",
            "value": done_all * 1e-3 / trk_tmax, "unit": "channel.s/s", "config": {"workload": WORKLOAD_TRK,
            "channels_per_gpu": nch, "if_seconds": trk_T, "code_periods_per_channel": nper},
            "steps": trk_steps, "warmup": trk_warm, "ms_per_step": trk_tmax * 1e3,
            "us_per_code_period": trk_tmax * 1e6 / nper, "gpu_launches": trk_steps, "dtype": "int32 sums, f64 loop filters",
            "e2e": {"value": done2_all * 1e-3 / t_trk_e2e, "unit": "channel.s/s",
                    "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 4 * (nch + 1)},
            "roofline": {"bound": "hbm", "achieved": a_bytes / t_trk / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": a_bytes / t_trk / 1e9 / hbm_peak, "traffic": trk_traffic, "peak_source": peak_src,
                         "kernel": "k_trk_cl", "algorithmic_bytes_per_launch": a_bytes,
                         "alu": {"bound": "int32 multiply-add issue", "unit": "Tops/s", "peak": int32_peak,
                                 "achieved": done * NSAMP * 28 * 2 / t_trk / 1e12,
                                 "frac": done * NSAMP * 28 * 2 / t_trk / 1e12 / int32_peak,
                                 "model": "SURVEY 8(d): currnsamp x (2 + 2 x 13 taps) MACs per channel-period"},
                         "note": "latency-bound time recurrence per channel (DESIGN.md); HBM fraction reported as asked"},
            "cpu_baseline": cpu_trk,
        }
        if phases:
            trk_obj["phases"] = phases
        if sweep:
            trk_obj["channel_sweep"] = {"if_seconds": 4.0, "note": "per GPU; first 4 s of every channel (bit sync falls inside), "
                                        "cluster size chosen per engine from co-residency", "points": sweep}
    # ------------------------------------------------ parity of the two headline configs, outside every clock
    # The CPU legs above ran the reference on host-generated IF; the same bytes go through the
    # CUDA path here and every decision is compared (the timed runs use the device-generated
    # copy of the same scene, which may differ from it in isolated samples by one LSB).
    parity = {"parity_checked": False, "why": "CPU legs skipped (--no-cpu or N > 1)"}
    if ref_acq is not None and host_par is not None:
        stp = sdr.Stream(1, len(host_par))
        stp.push(host_par)
        gp = acq.run(stp, b0)
        bad = [(p, ref_acq[p], (g["acquired"], g["freqi"], g["codei"])) for p, g in zip(prns, gp)
               if tuple(ref_acq[p]) != (g["acquired"], g["freqi"], g["codei"])]
        if bad:
            raise SystemExit(f"bench.py: acquisition parity FAILED against the reference (prn, ref, gpu): {bad[:4]}")
        parity = {"parity_checked": True, "oracle": oracle_kind(),
                  "acquisition": f"32 PRN: acquired / freqi / codei equal the reference's on the same IF bytes "
                                 f"({sum(1 for p in prns if ref_acq[p][0])} acquired)"}
        if ref_trk and not args.no_trk:
            g10 = {p: r for p, r in zip(TRK_PRNS, acq10.run(stp, b10))}
            nperp = min(r["n"] for r in ref_trk)
            chp, refs = [], []
            for r in ref_trk:
                g = g10[r["prn"]]
                if (r["acquired"], r["freqi"], r["codei"]) != (g["acquired"], g["freqi"], g["codei"]):
                    raise SystemExit(f"bench.py: hand-over parity FAILED for PRN {r['prn']}: ref "
                                     f"{(r['acquired'], r['freqi'], r['codei'])} gpu {(g['acquired'], g['freqi'], g['codei'])}")
                if g["acquired"]:
                    chp.append(sdr.make_trk_chan(FS, F_IF, F_CF, 1, r["prn"], sdr.gencode(r["prn"])[0], b10 + g["codei"],
                                                 g["acqfreq"], 6, 3, 6, (5.0, 1.0), (30.0, 10.0), (200.0, 50.0)))
                    refs.append(r)
            trp = sdr.Tracking(chp)
            recp, nrp = trp.run(stp, nperp)
            dII = dcf = 0.0
            for k, r in enumerate(refs):
                n_ = min(nrp[k], r["n"], nperp)
                if n_ < nperp:
                    raise SystemExit(f"bench.py: tracking parity FAILED: PRN {r['prn']} ran {n_} of {nperp} periods")
                g = recp[k][:n_]
                for key in ("currnsamp", "flagsync", "synci", "swsync"):
                    if not np.array_equal(g[key], r[key][:n_]):
                        raise SystemExit(f"bench.py: tracking parity FAILED: PRN {r['prn']} field {key}")
                if not np.array_equal(g["bit"][g["swsync"] == 1], r["bit"][:n_][r["swsync"][:n_] == 1]):
                    raise SystemExit(f"bench.py: tracking parity FAILED: PRN {r['prn']} nav bits differ")
                dII = max(dII, float(np.abs(g["II"][:, :13] - r["II"][:n_]).max()), float(np.abs(g["QQ"][:, :13] - r["QQ"][:n_]).max()))
                dcf = max(dcf, float(np.abs(g["carrfreq"] - r["carrfreq"][:n_]).max()))
            if dII > 508.0:
                raise SystemExit(f"bench.py: tracking parity FAILED: max |dII| {dII} > 508 (two flipped samples per tap)")
            parity["tracking"] = (f"{len(refs)} channels x {nperp} code periods: currnsamp, flagsync, synci, swsync, nav bits equal; "
                                  f"max |dII,dQQ| = {dII:g} (bound 508), max |dcarrfreq| = {dcf:.3g} Hz")
            trp.close()
        stp.close()

    # ------------------------------------------------ config-5 style: many independent IF streams
    # BASELINE configs[4]: RTL-SDR-like 8-bit I/Q at 2.048 MS/s, full acquisition + tracking per
    # stream, streams sharded across the GPUs (8 per GPU of a 64-stream job).  Every plan and
    # engine owns a CUDA stream; all of them are enqueued before the first wait, so the
    # channels of all streams share the SMs.  Reported, not the headline.
    streams_obj = None
    if args.c5_streams > 0:
        FS5, NS5 = 2.048e6, 2048
        T5 = args.c5_seconds
        nper5 = int(T5 * 1000) - 8
        tot5 = 65536 + (nper5 + 8) * NS5
        prn_pool = [2, 5, 7, 13, 19, 25, 29, 30, 4, 9, 16, 23]
        freq5 = np.array([(i - 35) * 200.0 for i in range(NFREQ)])
        codes5 = np.stack([sdr.gencode(p)[0] for p in range(1, NPRN + 1)])
        s5, a5 = [], []
        for k in range(args.c5_streams):
            rng = np.random.default_rng(0xB2000500 + 64 * rank + k)
            vis = sorted(rng.choice(prn_pool, size=8, replace=False).tolist())
            sats = [dict(prn=int(p_), cn0=float(rng.uniform(52, 56)),
                         doppler=float(int(rng.integers(-20, 21)) * 200.0 + rng.uniform(-30, 30)),
                         code_phase=float(rng.uniform(0, 1023 * 20)), carr_phase=float(rng.uniform(0, 1)),
                         rate=float(rng.uniform(-1, 1))) for p_ in vis]
            stk = sdr.Stream(2, tot5)
            stk.push_ifgen(sdr.make_ifgen(FS5, 0.0, F_CF, 2, sats, sigma=20.0, seed=0xB2000500 + 64 * rank + k), tot5)
            s5.append(stk)
            a5.append(sdr.Acquisition(FS5, 2, NS5, 2, 1e-3, freq5, codes5, 1.023e6, intg=10))
        b5 = 65536 - 11 * NS5
        for _ in range(2):                                   # warm
            for ak, sk in zip(a5, s5): ak.enqueue(sk, b5)
            got5 = [ak.fetch() for ak in a5]
        barrier()
        t0 = time.perf_counter()
        for ak, sk in zip(a5, s5): ak.enqueue(sk, b5)
        got5 = [ak.fetch() for ak in a5]
        sdr.check(lib.sdrb200_device_sync(), "sync")
        t_acq5 = time.perf_counter() - t0
        steps5 = sum(max(r["nintg"] for r in g) for g in got5)          # integration steps actually run
        engines = []
        for g, sk in zip(got5, s5):
            ch5 = [sdr.make_trk_chan(FS5, 0.0, F_CF, 2, p_, sdr.gencode(p_)[0], b5 + r["codei"], r["acqfreq"],
                                     4, 1, 1, (5.0, 2.0), (30.0, 20.0), (200.0, 50.0))
                   for p_, r in zip(range(1, NPRN + 1), g) if r["acquired"]]
            engines.append(sdr.Tracking(ch5) if ch5 else None)
        nch5 = sum(e.nch for e in engines if e)
        # all engines run at once: the cluster size follows the TOTAL number of channels on the GPU
        # (the small CTA instantiation fits two per SM)
        cl5 = 1
        while cl5 < 8 and nch5 * cl5 * 2 <= 2 * max(sdr.lib.sdrb200_sm_count(), 1):
            cl5 *= 2
        if os.environ.get("BENCH_C5_CLUSTER"):                       # A/B runs
            cl5 = int(os.environ["BENCH_C5_CLUSTER"])
        for e in engines:
            if e: e.set_cluster(cl5)
        for e, sk in zip(engines, s5):
            if e: e.enqueue(sk, 200)
        for e in engines:
            if e: e.sync(); e.reset()
        barrier()
        t0 = time.perf_counter()
        for e, sk in zip(engines, s5):
            if e: e.enqueue(sk, nper5)
        for e in engines:
            if e: e.sync()
        t_trk5 = time.perf_counter() - t0
        done5 = 0
        for e, sk in zip(engines, s5):
            if e:
                e.reset()
        tv = [t_acq5, t_trk5]
        if dist is not None:
            tt = torch.tensor(tv, dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            tv = [float(tt[0]), float(tt[1])]
            tn = torch.tensor([float(nch5), float(steps5)], dtype=torch.float64, device="cuda")
            dist.all_reduce(tn)
            nch5_all, steps5_all = float(tn[0]), float(tn[1])
        else:
            nch5_all, steps5_all = float(nch5), float(steps5)
        # the whole job of BASELINE configs[4] under ONE clock: acquisition of every stream, hand-over on
        # the host (tuples out, channel tables in, tracking engines created), tracking of every acquired PRN
        for e in engines:
            if e: e.close()
        barrier()
        t0 = time.perf_counter()
        for ak, sk in zip(a5, s5): ak.enqueue(sk, b5)
        gotf = [ak.fetch() for ak in a5]
        engf = []
        for g, sk in zip(gotf, s5):
            chf = [sdr.make_trk_chan(FS5, 0.0, F_CF, 2, p_, codes5[p_ - 1], b5 + r["codei"], r["acqfreq"],
                                     4, 1, 1, (5.0, 2.0), (30.0, 20.0), (200.0, 50.0))
                   for p_, r in zip(range(1, NPRN + 1), g) if r["acquired"]]
            ef = sdr.Tracking(chf) if chf else None
            if ef: ef.set_cluster(cl5); ef.enqueue(sk, nper5)
            engf.append(ef)
        donef = 0
        for ef in engf:
            if ef: ef.sync()
        barrier()
        t_full5 = time.perf_counter() - t0
        for ef, sk in zip(engf, s5):
            if ef:
                _, nrf = ef.run(sk, 1, want_records=False)      # one more period: every channel is still alive
                donef += ef.nch
        if dist is not None:
            tt = torch.tensor([t_full5], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t_full5 = float(tt[0])
            tn2 = torch.tensor([float(donef)], dtype=torch.float64, device="cuda")
            dist.all_reduce(tn2)
            donef_all = float(tn2[0])
        else:
            donef_all = float(donef)
        engines = engf
        streams_obj = {
            "full_job": {"what": "acquisition + hand-over + tracking of every stream under one clock (plans and IF resident; "
                                 "tracking engines are created inside the clock)",
                         "ms": t_full5 * 1e3, "streams": world * args.c5_streams, "if_seconds_per_stream": T5,
                         "stream_seconds_per_s": world * args.c5_streams * T5 / t_full5,
                         "channel_s_per_s": donef_all * nper5 * 1e-3 / t_full5, "channels": donef_all},
            "workload": f"{args.c5_streams} independent I/Q int8 streams per GPU at 2.048 MS/s (BASELINE configs[4] style): "
                        f"32 PRN x 71 bins x 2048 acquisition (10 integrations, early exit) then {T5:g} s tracking "
                        "of every acquired PRN, 9 taps",
            "streams_per_gpu": args.c5_streams, "channels_tracked": nch5_all,
            "acquisition": {"cells_per_s": NPRN * NFREQ * NS5 * steps5_all / tv[0], "ms": tv[0] * 1e3,
                            "integration_steps": steps5_all},
            "tracking": {"value": nch5_all * nper5 * 1e-3 / tv[1], "unit": "channel.s/s", "ms": tv[1] * 1e3, "cluster_size": cl5,
                         "us_per_code_period_per_channel": tv[1] * 1e6 / nper5},
            "timing": "host clock around enqueue-all .. sync-all (several CUDA streams), device idle before",
        }
        for e in engines:
            if e: e.close()
        for ak in a5: ak.close()
        for sk in s5: sk.close()
    # ------------------------------------------------ BASELINE configs[2]: GLONASS G1 FDMA acquisition
    # 14 frequency channels (k = -7..6, 562.5 kHz apart) x 101 bins (+-10 kHz / 200 Hz) x 20000 code
    # phases, I/Q int8 at 20 MS/s (stereo_L1G1 style), 10 non-coherent integrations (2 n = 40 000 <= M = 40 960 =
    # 10 x 4096: the fused cluster path, radix-10 recombination).  One plan per channel (its own Doppler grid, src/sdrinit.c:633-635),
    # all enqueued before the first wait.  Parity: tests/test_gpu_headline.py::test_c3_*.
    c3_obj = None
    if args.c3:
        fs3, ns3, nf3, intg3 = 20.0e6, 20000, 101, 10
        rng3 = np.random.default_rng(0xB2000300 + rank)
        ks3 = list(range(-7, 7))
        sats3 = [dict(prn=1, cn0=float(rng3.uniform(46, 50)), doppler=k_ * 562500.0 + float(rng3.uniform(-9000, 9000)),
                      code_phase=float(rng3.uniform(0, 511 * 10)), carr_phase=float(rng3.uniform(0, 1)), rate=0.0, nav_on=0)
                 for k_ in (-7, -4, -1, 0, 2, 5, 6)]
        need3 = (intg3 + 1) * ns3
        avail3 = 65536 * (need3 // 65536 + 1)
        st3 = sdr.Stream(2, avail3)
        st3.push_ifgen(sdr.make_ifgen(fs3, 0.0, 1.602e9, 2, sats3, sigma=2.5, seed=0xB2000300 + rank, crate=0.511e6, clen=511, bitlen=10), avail3)
        code3, crate3 = sdr.gencode(1, 20)
        grids3 = np.array([[(i - nf3 // 2) * 200.0 + 562500.0 * k_ for i in range(nf3)] for k_ in ks3])   # src/sdrinit.c:633-635
        b3 = avail3 - need3

        def _time3(plans):
            for _ in range(2):
                for pl in plans: pl.enqueue(st3, b3)
                for pl in plans: pl.sync()
            ts = []
            for _ in range(3):
                sdr.check(lib.sdrb200_l2_flush(C.c_void_p(plans[0].cuda_stream)), "l2_flush")
                barrier()
                t0 = time.perf_counter()
                for pl in plans: pl.enqueue(st3, b3)
                for pl in plans: pl.sync()
                ts.append(time.perf_counter() - t0)
            tm = min(ts)
            if dist is not None:
                tt = torch.tensor([tm], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                tm = float(tt[0])
            return tm
        # ONE plan: the G1 code once per channel, every channel over its own Doppler grid (cfg.grid_per_prn)
        one3 = sdr.Acquisition(fs3, 2, ns3, int(ns3 / 511), 1e-3, grids3, np.repeat(code3[None, :], len(ks3), 0), crate3,
                               intg=intg3, threshold=1e30)                                              # every step runs
        t3m = _time3([one3])
        g3 = one3.fetch()
        # per-kernel share of one search (events around every launch; adds their cost)
        sdr.check(lib.sdrb200_acq_set_timing(one3.h, 1), "set_timing")
        one3.enqueue(st3, b3); one3.sync()
        ms4_3 = (C.c_double * 4)(); c4_3 = (C.c_longlong * 4)()
        sdr.check(lib.sdrb200_acq_kernel_ms(one3.h, ms4_3, c4_3), "kernel_ms")
        sdr.check(lib.sdrb200_acq_set_timing(one3.h, 0), "set_timing")
        kern3 = {"forward (k_fwd_split_small + k_fft_rows)": ms4_3[0], "k_rowc<search>": ms4_3[1], "k_rowc<peak>": ms4_3[3]}
        # round-1/2 arrangement for comparison: one plan per channel on 14 CUDA streams
        plans3 = [sdr.Acquisition(fs3, 2, ns3, int(ns3 / 511), 1e-3, grids3[i], code3[None, :], crate3, intg=intg3, threshold=1e30)
                  for i in range(len(ks3))]
        t3p = _time3(plans3)
        g3p = [pl.fetch()[0] for pl in plans3]
        for a_, b_ in zip(g3, g3p):
            if (a_["freqi"], a_["codei"], a_["nintg"]) != (b_["freqi"], b_["codei"], b_["nintg"]):
                raise SystemExit("bench.py: batched GLONASS plan disagrees with the per-channel plans")
        cells3 = world * len(ks3) * nf3 * ns3 * intg3
        c3_obj = {"workload": f"GLONASS G1 FDMA: {len(ks3)} channels x {nf3} bins x {ns3} code phases x {intg3} integrations, "
                              "I/Q int8 at 20 MS/s (BASELINE configs[2]); one IF stream per GPU",
                  "cells_per_s": cells3 / t3m, "ms_per_search": t3m * 1e3,
                  "plan": "one plan, 14 replicas x their own 101-bin grid (grid_per_prn): 4 launches per integration",
                  "kernel_ms_per_search": {k_: round(v_, 3) for k_, v_ in kern3.items()},
                  "one_plan_per_channel": {"cells_per_s": cells3 / t3p, "ms_per_search": t3p * 1e3,
                                           "note": "14 plans on 14 CUDA streams (the arrangement of the earlier runs); tuples equal"},
                  "peak_ratio_per_channel": {str(k_): round(g_["peakr"], 2) for k_, g_ in zip(ks3, g3)},
                  "channels_with_signal": [-7, -4, -1, 0, 2, 5, 6],
                  "hbm": {"model": "SURVEY 8(d): IF + code spectrum + power matrix read/write per step",
                          "frac": world * len(ks3) * (intg3 * (2 * ns3 * 2 + 40960 * 8) + nf3 * ns3 * 4 * (2 * intg3 - 1)) / t3m / 1e9 / (hbm_peak * world)},
                  "timing": "host clock around enqueue .. sync, L2 flushed before"}
        one3.close()
        for pl in plans3: pl.close()
        st3.close()
    # ------------------------------------------------ config-4 style: long coherent blocks
    # BASELINE configs[3]: weak-signal search, 100 Hz Doppler step (141 bins), 20 ms coherent
    # blocks (n = 327 360 lags, M = 655 360 = 160 x 4096: the two-pass path k_fwd_split/k_fft_rows/k_row/
    # k_comb_fft through an L2-sized scratch), non-coherent accumulation in the power matrix.
    # A PRN slab of 16 (the 8-GPU share of 51 PRNs is 7) and 3 non-coherent steps; reported,
    # not the headline.
    c4_obj = None
    if args.c4_prns > 0:
        ncoh4, nf4, intg4 = 20, 141, 3
        n4 = ncoh4 * NSAMP
        need4 = (intg4 + 1) * n4 + 4096
        if int(st.count) >= need4:
            freq4 = np.array([F_IF + (i - nf4 // 2) * 100.0 for i in range(nf4)])
            codes4 = np.stack([sdr.gencode(p_)[0] for p_ in range(1, args.c4_prns + 1)])
            acq4 = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq4, codes4, 1.023e6, intg=intg4, ncoh=ncoh4,
                                   threshold=1e30)                     # no early exit: every step runs
            s4 = acq4.cuda_stream
            for _ in range(2):
                acq4.enqueue(st, 0); r4 = acq4.fetch()
            barrier()
            t4 = []
            for _ in range(3):
                sdr.check(lib.sdrb200_l2_flush(C.c_void_p(s4)), "l2_flush")
                e0.record(s4); acq4.enqueue(st, 0); e1.record(s4)
                r4 = acq4.fetch()
                t4.append(e1.ms_since(e0))
            ms4 = sum(t4) / len(t4)
            if dist is not None:                             # every rank searched its own stream
                tt4 = torch.tensor([ms4], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt4, op=dist.ReduceOp.MAX)
                ms4 = float(tt4[0])
            cells4 = world * args.c4_prns * nf4 * n4 * intg4
            m4 = 5 << 17                                     # transform length the plan chooses for 2 n = 654 720
            a_bytes4 = world * (intg4 * (2 * n4 + args.c4_prns * m4 * 8) + args.c4_prns * nf4 * n4 * 4 * (1 + 2 * (intg4 - 1)))
            c4_obj = {"workload": f"{args.c4_prns} PRN x {nf4} bins (100 Hz) x {n4} lags, {ncoh4} ms coherent blocks "
                                  f"(M = 655360), {intg4} non-coherent steps (BASELINE configs[3] style)",
                      "cells_per_s": cells4 / (ms4 * 1e-3), "ms_per_run": ms4, "ms_per_step": ms4 / intg4,
                      "algorithmic_bytes_per_run": a_bytes4,
                      "hbm": {"achieved_GBps": a_bytes4 / (ms4 * 1e-3) / 1e9, "peak_GBps": hbm_peak * world,
                              "frac": a_bytes4 / (ms4 * 1e-3) / 1e9 / (hbm_peak * world),
                              "model": "SURVEY 8(d): IF + code-spectrum bank + power matrix read/write per step"},
                      "l2": "flushed before every timed run", "peak_bins": [r["freqi"] for r in r4[:4]]}
            acq4.close()
    # ------------------------------------------------ the partition north_star names: ONE search over N GPUs
    # (a) strong scaling: BASELINE configs[3] at its full size - 51 PRN (GPS 1-32 + SBAS 120-138) x 141 bins
    #     (100 Hz) x 327 360 lags (20 ms coherent, M = 655 360) x 10 non-coherent steps with the reference's
    #     early exit - split by PRN slab; every rank searches the SAME IF stream; the tuples are gathered
    #     ONCE per search from device memory (sdrb200_acq_results_device) with one NCCL all-gather.
    # (b) Doppler slabs for PRNs < GPUs: 4 PRN x 71 bins, 4 integrations, bins split over the ranks,
    #     per-step histories gathered once, early exit replayed (shard.merge_doppler_slabs).
    strong_obj = dslab_obj = None
    if args.strong_prns > 0:
        ncoh4, nf4, intg4 = 20, 141, args.strong_steps
        n4 = ncoh4 * NSAMP
        tot4 = (intg4 + 1) * n4 + 65536
        prns4 = (list(range(1, 33)) + list(range(120, 139)))[:args.strong_prns]
        rng4 = np.random.default_rng(0xB2000400)
        sats4 = [dict(prn=p_, cn0=float(c_), doppler=float(int(rng4.integers(-60, 61)) * 100.0 + rng4.uniform(-20, 20)),
                      code_phase=float(rng4.uniform(0, 1023 * 20)), carr_phase=float(rng4.uniform(0, 1)), rate=0.0, nav_on=0)
                 for p_, c_ in ((3, 42.0), (11, 38.0), (17, 45.0), (25, 33.0), (31, 36.0), (133, 40.0), (122, 34.0)) if p_ in prns4]
        st4 = sdr.Stream(1, tot4)
        st4.push_ifgen(sdr.make_ifgen(FS, F_IF, F_CF, 1, sats4, sigma=24.0, seed=0xB2000400), tot4)     # same stream on every rank
        freq4 = np.array([F_IF + (i - nf4 // 2) * 100.0 for i in range(nf4)])
        mine4 = sdr.shard.prn_slab(prns4, rank, world)
        width4 = -(-len(prns4) // world)
        rb = C.sizeof(sdr.AcqResult)
        plan4 = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq4,
                                np.stack([sdr.gencode(p_)[0] for p_ in mine4]), 1.023e6, intg=intg4, ncoh=ncoh4) if mine4 else None
        loc4 = torch.zeros(width4 * rb, dtype=torch.uint8, device="cuda")
        all4 = torch.zeros(world * width4 * rb, dtype=torch.uint8, device="cuda")
        s4s = plan4.cuda_stream if plan4 else 0

        def strong_search():
            if plan4:
                plan4.enqueue(st4, 0)
                plan4.copy_results_device(loc4.data_ptr())
                plan4.sync()
            if dist is not None:
                dist.all_gather_into_tensor(all4, loc4)
                torch.cuda.synchronize()

        strong_search()                                      # warm
        t_s, t_k = [], []
        for _ in range(2):
            if plan4: sdr.check(lib.sdrb200_l2_flush(C.c_void_p(s4s)), "l2_flush")
            barrier()
            t0 = time.perf_counter()
            if plan4: e0.record(s4s)
            strong_search()
            if plan4: e1.record(s4s)
            barrier()
            t_s.append(time.perf_counter() - t0)
            t_k.append(e1.ms_since(e0) if plan4 else 0.0)
        ts4 = min(t_s)
        kern4 = None
        if plan4:                                            # per-kernel share of one search on this rank (event pairs)
            sdr.check(lib.sdrb200_acq_set_timing(plan4.h, 1), "set_timing")
            plan4.enqueue(st4, 0); plan4.sync()
            ms4_4 = (C.c_double * 4)(); c4_4 = (C.c_longlong * 4)()
            sdr.check(lib.sdrb200_acq_kernel_ms(plan4.h, ms4_4, c4_4), "kernel_ms")
            sdr.check(lib.sdrb200_acq_set_timing(plan4.h, 0), "set_timing")
            tot4_ms = max(sum(ms4_4[i] for i in range(4)), 1e-9)   # the plan keeps 256 event pairs: shares of those, not totals
            kern4 = {"forward (k_fwd_split + k_fft_rows)": round(ms4_4[0] / tot4_ms, 3), "k_row": round(ms4_4[1] / tot4_ms, 3),
                     "k_comb_fft": round(ms4_4[2] / tot4_ms, 3), "k_peak_hbm": round(ms4_4[3] / tot4_ms, 3),
                     "unit": "share of the timed launches (first 256 of the search)"}
        if dist is not None:
            tt = torch.tensor([ts4], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ts4 = float(tt[0])
            gathered = all4.cpu().numpy().reshape(world, width4 * rb)
            merged4 = []
            for r_ in range(world):
                n_r = len(sdr.shard.prn_slab(prns4, r_, world))
                merged4 += sdr.Acquisition.results_from_bytes(gathered[r_, :n_r * rb])
        else:
            merged4 = sdr.Acquisition.results_from_bytes(loc4.cpu().numpy()[:len(mine4) * rb])
        if plan4: plan4.close()
        eq4 = None
        if rank == 0 and world > 1:
            # the unsharded search on rank 0, outside the clock: the merged tuples must equal it exactly
            whole4 = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq4,
                                     np.stack([sdr.gencode(p_)[0] for p_ in prns4]), 1.023e6, intg=intg4, ncoh=ncoh4)
            want4 = whole4.run(st4, 0)
            whole4.close()
            if want4 != merged4:
                raise SystemExit("bench.py: PRN-slab search differs from the unsharded search")
            eq4 = True
        steps_run = max(m_["nintg"] for m_ in merged4)
        strong_obj = {"workload": f"ONE search, {len(prns4)} PRN (GPS 1-32 + SBAS 120-138) x {nf4} bins (100 Hz) x {n4} lags, "
                                  f"{ncoh4} ms coherent (M = 655360 = 160 x 4096), up to {intg4} non-coherent steps with early exit (BASELINE configs[3])",
                      "scaling": "strong", "partition": "PRN slabs, one NCCL all-gather of the tuples per search from device memory",
                      "kernel_share_this_rank": kern4,
                      "n_gpus": world, "prns_per_gpu": [len(sdr.shard.prn_slab(prns4, r_, world)) for r_ in range(world)],
                      "ms_per_search": ts4 * 1e3, "ms_search_kernels_this_rank": min(t_k),
                      "cells_per_s": len(prns4) * nf4 * n4 * steps_run / ts4, "integration_steps_run": steps_run,
                      "acquired_prns": [p_ for p_, m_ in zip(prns4, merged4) if m_["acquired"]],
                      "note_acquired": "checkacquisition's +-2-chip exclusion (src/sdracq.c:78-83) does not cover the code-period "
                                       "replicas of the peak inside a 20 ms block, so the peak ratio stays near 1 at this size; "
                                       "the decisions that are compared are (freqi, codei, maxp)",
                      "found_bins": {str(p_): [m_["freqi"], m_["codei"] % NSAMP] for p_, m_ in zip(prns4, merged4)
                                     if p_ in (3, 11, 17, 25, 31, 133, 122)},
                      "true_bins": {str(s_["prn"]): int(round(s_["doppler"] / 100.0)) + nf4 // 2 for s_ in sats4},
                      "merged_equals_unsharded": eq4,
                      "hbm_model_bytes": len(prns4) * nf4 * n4 * 4 * (2 * steps_run - 1),
                      "hbm_frac": len(prns4) * nf4 * n4 * 4 * (2 * steps_run - 1) / ts4 / 1e9 / (hbm_peak * world)}
        st4.close()
        if world > 1:
            prnsd = [1, 3, 11, 2]                             # fewer PRNs than GPUs: the Doppler grid is what shards
            intgd = 4
            bd = sdr.shard.doppler_slab(NFREQ, rank, world)
            codesd = np.stack([sdr.gencode(p_)[0] for p_ in prnsd])
            bd0 = AVAIL10 - (intgd + 1) * NSAMP
            rbh = intgd * len(prnsd) * rb
            loch = torch.zeros(rbh, dtype=torch.uint8, device="cuda")
            allh = torch.zeros(world * rbh, dtype=torch.uint8, device="cuda")
            pland = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq[bd.start:bd.stop], codesd, 1.023e6,
                                    intg=intgd, threshold=1e30) if len(bd) else None
            if pland: pland.set_history(True)
            # every rank holds the same scene in `st` only at N = 1; here each rank has its own stream, so the
            # slab search runs on a common stream made for it
            std = sdr.Stream(1, AVAIL10 + 65536)
            std.push_ifgen(scene(sdr, 0xB2000000), AVAIL10 + 65536)

            def dslab_search():
                if pland:
                    pland.enqueue(std, bd0)
                    pland.sync()
                    loch.copy_(torch.as_tensor(pland.history_device(intgd), device="cuda"))
                dist.all_gather_into_tensor(allh, loch)
                torch.cuda.synchronize()

            dslab_search()
            barrier()
            t0 = time.perf_counter()
            for _ in range(K):
                dslab_search()
            barrier()
            td = (time.perf_counter() - t0) / K
            tt = torch.tensor([td], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            td = float(tt[0])
            gh = allh.cpu().numpy().reshape(world, rbh)
            hists, offs = [], []
            for r_ in range(world):
                b_ = sdr.shard.doppler_slab(NFREQ, r_, world)
                if len(b_) == 0:
                    continue
                flat = sdr.Acquisition.results_from_bytes(gh[r_])
                hists.append([flat[i * len(prnsd):(i + 1) * len(prnsd)] for i in range(intgd)])
                offs.append(b_.start)
            mergedd = sdr.shard.merge_doppler_slabs(hists, offs, NSAMP, 3.0)
            eqd = None
            if rank == 0:
                wholed = sdr.Acquisition(FS, 1, NSAMP, int(NSAMP / 1023), 1e-3, freq, codesd, 1.023e6, intg=intgd)
                wantd = wholed.run(std, bd0)
                wholed.close()
                keys = ("acquired", "freqi", "codei", "nintg", "acqfreq", "peakr", "cn0", "maxp", "maxp2")
                if [[m_[k_] for k_ in keys] for m_ in mergedd] != [[w_[k_] for k_ in keys] for w_ in wantd]:
                    raise SystemExit("bench.py: Doppler-slab search differs from the unsharded search")
                eqd = True
            if pland: pland.close()
            std.close()
            dslab_obj = {"workload": f"ONE search, {len(prnsd)} PRN x {NFREQ} bins x {NSAMP} lags, {intgd} integrations, bins split over the GPUs",
                         "scaling": "strong", "partition": "Doppler slabs, per-step tuples gathered once per search (device memory), early exit replayed",
                         "n_gpus": world, "bins_per_gpu": [len(sdr.shard.doppler_slab(NFREQ, r_, world)) for r_ in range(world)],
                         "ms_per_search": td * 1e3, "cells_per_s": len(prnsd) * NFREQ * NSAMP * intgd / td,
                         "merged_equals_unsharded": eqd}
    # ------------------------------------------------ the sdr.h drop-in against the reference, same replay
    # oracle/_ref/libreplay_b200.so = the reference's cold objects + its channel-thread loop with every hot
    # sdr.h symbol bound to libsdrb200_sdrh.so; libreplay_ref.so = all reference.  Same IF bytes, the
    # reference's threading model (one pthread per channel), 12 channels (BASELINE configs[1] scene).
    dropin_obj = None
    if rank == 0 and world == 1 and host_par is not None and not args.no_dropin:
        import oracle_bind as ob
        if ob.have_ref() and ob.have_dropin():
            nperd = 2000
            cfgd = ob.cfg_iffile(acq_intg=10)

            def _quiet(fn):
                devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1); os.dup2(devnull, 1)
                try:
                    return fn()
                finally:
                    os.dup2(saved, 1); os.close(devnull); os.close(saved)

            def _run(o):
                # tracking time = difference of two replays that share the same 12 acquisitions
                # (best of 3 each: the acquisitions dominate the short one)
                nshort = 200
                ts, tl, ntl = [], [], 0
                for _ in range(3):
                    t0 = time.perf_counter()
                    _, _, nts = _quiet(lambda: o.acq_track(cfgd, host_par, AVAIL10, TRK_PRNS, nshort, threaded=True))
                    ts.append(time.perf_counter() - t0)
                    t0 = time.perf_counter()
                    _, _, ntr = _quiet(lambda: o.acq_track(cfgd, host_par, AVAIL10, TRK_PRNS, nperd, threaded=True))
                    tl.append(time.perf_counter() - t0)
                    ntl = sum(ntr) - sum(nts)
                return min(ts), max(min(tl) - min(ts), 1e-6), ntl
            od, orf = ob.dropin(), ob.ref()
            od.set_lookahead(True)
            _run(od)                                               # warm: plans, engines, mirror
            served0 = od.lookahead_served()
            ta_d, tt_d, n_d = _run(od)
            served = od.lookahead_served() - served0
            ta_r, tt_r, n_r = _run(orf)
            dropin_obj = {"what": "reference channel-thread loop (one pthread per channel) with the hot sdr.h symbols from "
                                  "libsdrb200_sdrh.so, against the all-reference build; same IF bytes, same host",
                          "channels": len(TRK_PRNS), "code_periods_per_channel": nperd,
                          "tracking_channel_s_per_s": {"dropin": n_d * 1e-3 / tt_d, "reference": n_r * 1e-3 / tt_r,
                                                       "ratio": (n_d / tt_d) / (n_r / tt_r)},
                          "acquisition_plus_200_periods_ms": {"dropin": ta_d * 1e3, "reference": ta_r * 1e3},
                          "timing": "tracking = (replay of 2000 periods) - (replay of 200 periods), best of 3 each, host clock",
                          "periods_served_from_engine_records": served, "host_cores": os.cpu_count()}
    clk = clocks.stop() if clocks else None

    # ------------------------------------------------ reduce over ranks, print
    tmax, te2e, te2e1 = t_acq, t_e2e, t_e2e_single
    if dist is not None:
        tt = torch.tensor([t_acq, t_e2e, t_e2e_single], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        tmax, te2e, te2e1 = float(tt[0]), float(tt[1]), float(tt[2])
    dom = max(kern, key=lambda n: kern[n]["ms_total"])
    per_launch_units = CELLS_PER_STEP * K / max(kern[dom]["launches"], 1)
    a_bytes = 0.226 * per_launch_units                      # 8.41e6 B / 37 188 096 cells
    a_flops = 170.0 * per_launch_units
    dur = kern[dom]["ms_per_launch"] * 1e-3
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(dom)
    except Exception:
        pass
    line = {
        "metric": "acquisition PRN x Doppler x code-phase cells/s", "value": world * CELLS_PER_STEP * K / tmax,
        "unit": "cells/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": tmax * 1e3 / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 FFT (int8 IF, int LUT wipe-off)",
        "data": "synthetic", "impl": "b200",
        "config": {"workload": WORKLOAD_ACQ, "if_streams": world, "cells_per_step_per_gpu": CELLS_PER_STEP,
                   "l2": "flushed between timed iterations (384 MB memset)", "multi_gpu": "one IF stream per rank"},
        "e2e": {"value": world * CELLS_PER_STEP * K / te2e, "unit": "cells/s", "h2d_bytes_per_step": need,
                "d2h_bytes_per_step": NPRN * C.sizeof(sdr.AcqResult),
                "api": "sdrb200_acq_enqueue_host + sdrb200_acq_fetch on two plans (pinned host IF in and result tuples out "
                       "every step; step i+1 is queued while step i runs)",
                "single_call": {"value": world * CELLS_PER_STEP * K / te2e1, "unit": "cells/s",
                                "api": "sdrb200_acq_run_host: one blocking call per step"}},
        "gpu_launches": int(launches_timed),
        "kernels": kern,
        "roofline": {"bound": "hbm", "achieved": a_bytes / dur / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": a_bytes / dur / 1e9 / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                     "kernel": dom, "algorithmic_bytes_per_launch": a_bytes,
                     "alu": {"bound": "fp32 issue", "unit": "TFLOP/s", "peak": fp32_peak,
                             "achieved": a_flops / dur / 1e12, "frac": a_flops / dur / 1e12 / fp32_peak,
                             "model": "SURVEY 8(d): 170 flop per cell"},
                     "note": "compulsory-traffic model: the path is FFT/shared-memory bound, not HBM bound (DESIGN.md)"},
        "cpu_baseline": cpu_acq,
        "parity": parity, "parity_checked": parity["parity_checked"],
        "clocks": clk,
        "acquired_prns": [p for p, r in zip(prns, res) if r["acquired"]],
        "tracking": trk_obj,
        "streams_c5": streams_obj,
        "coherent_c4": c4_obj,
        "glonass_c3": c3_obj,
        "strong_scaling_c4": strong_obj,
        "doppler_slabs": dslab_obj,
        "dropin": dropin_obj,
    }
    if rank == 0:
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


# ------------------------------------------------------------ reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    K, W = args.steps, args.warmup
    pool = cpu_pool(cores, 0xB2000000, AVAIL)             # every worker holds the IF before a clock starts
    prns = list(range(1, NPRN + 1))
    try:
        dt, _ = cpu_acq_step(pool, cores, 0xB2000000, prns[:min(cores, NPRN)])     # calibrate
        per_prn = dt                                         # seconds for one PRN per process
        rounds = -(-NPRN // cores)
        est = per_prn * rounds * (K + W)
        nsub = NPRN
        if est > 150.0:                                      # bounded sample: fewer PRNs per step
            nsub = max(min(cores, NPRN), int(NPRN * 150.0 / est) // max(min(cores, NPRN), 1) * min(cores, NPRN))
            nsub = max(1, min(NPRN, nsub))
        sub = prns[:nsub]
        for _ in range(W):
            cpu_acq_step(pool, cores, 0xB2000000, sub)
        t, fft_t, cpu_t, mapped = 0.0, 0.0, 0.0, False
        for _ in range(K):
            dt, out = cpu_acq_step(pool, cores, 0xB2000000, sub)
            t += dt
            fft_t += sum(w[3] or 0.0 for w in out); cpu_t += sum(w[0] for w in out)
            mapped = mapped or any(w[4] for w in out)
    finally:
        pool.close()
    cells = nsub * NFREQ * NSAMP
    v = cells * K / t
    used = min(cores, nsub)
    sample = f"{nsub} of 32 PRNs x 71 bins x 16368 per step, {used} processes"
    line = {
        "metric": "acquisition PRN x Doppler x code-phase cells/s", "value": v, "unit": "cells/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": t * 1e3 / K, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 FFT (int8 IF, int LUT wipe-off)", "data": "synthetic",
        "impl": "reference",
        "config": {"workload": WORKLOAD_ACQ, "sample": sample},
        "cpu_baseline": {"value": v, "unit": "cells/s", "cores": used, "kind": oracle_kind(),
                         "impl": oracle_label(), "sample": sample, "fft": "shim",
                         "fftw3f_on_this_host": __import__("ctypes.util").util.find_library("fftw3f"),
                         "fft_share_of_cpu_time": (fft_t / cpu_t) if cpu_t > 0 and fft_t > 0 else None,
                         "if_generated": "before the clock (pool initializer, oracle/libifgen.so)",
                         "product_library_mapped": bool(mapped)},
        "e2e": {"value": v, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


_REAL_STDOUT = None


def guard_stdout():
    """stdout carries exactly ONE JSON line: everything libraries print on fd 1 (NCCL's
    version banner, for one) is sent to stderr; emit() writes to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--trk-seconds", type=float, default=30.0)
    ap.add_argument("--trk-steps", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-trk", action="store_true", help="skip the tracking half")
    ap.add_argument("--c5-streams", type=int, default=8, help="config-5 style independent streams per GPU (0: skip)")
    ap.add_argument("--c5-seconds", type=float, default=10.0)
    ap.add_argument("--c4-prns", type=int, default=0, help="config-4 style 20 ms coherent search over this many PRNs (0: skip)")
    ap.add_argument("--no-c3", dest="c3", action="store_false", help="skip the GLONASS G1 FDMA acquisition object")
    ap.add_argument("--strong-prns", type=int, default=51, help="strong-scaling search (BASELINE configs[3]) over this many PRNs (0: skip)")
    ap.add_argument("--strong-steps", type=int, default=10, help="non-coherent steps of the strong-scaling search")
    ap.add_argument("--no-dropin", action="store_true", help="skip the drop-in vs reference replay object")
    ap.add_argument("--no-sweep", action="store_true", help="skip the tracking channel-count sweep")
    ap.add_argument("--trk-profile", action="store_true", help="add the per-phase cycle profile of channel 0")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3 if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
