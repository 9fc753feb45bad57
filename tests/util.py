"""shared helpers of the parity tests: synthetic scenes."""
import numpy as np

FS1, IF1, CF = 16.368e6, 4.092e6, 1575.42e6
FS5 = 2.048e6


def scene_c1(sdr, nsats=8, seed=0xB2000001):
    """config-1 style: real IF at 16.368 MS/s, visible PRNs at 47-50 dB-Hz."""
    rng = np.random.default_rng(seed)
    prns = [1, 3, 6, 9, 14, 17, 22, 31][:nsats]
    sats = []
    for p in prns:
        bin_ = rng.integers(-24, 25)
        dop = bin_ * 200.0 + rng.uniform(-70, 70)       # >= 30 Hz from a half-bin edge
        sats.append(dict(prn=p, cn0=rng.uniform(47, 50), doppler=float(dop),
                         code_phase=float(rng.uniform(0, 1023 * 20)),
                         carr_phase=float(rng.uniform(0, 1)), rate=float(rng.uniform(-1, 1))))
    return sdr.make_ifgen(FS1, IF1, CF, 1, sats, sigma=24.0, seed=seed), prns


def scene_c5(sdr, nsats=6, seed=0xB2000005):
    """config-5 style: RTL-SDR-like I/Q at 2.048 MS/s, zero IF."""
    rng = np.random.default_rng(seed)
    prns = [2, 5, 7, 13, 19, 25, 29, 30][:nsats]
    sats = []
    for p in prns:
        bin_ = rng.integers(-24, 25)
        dop = bin_ * 200.0 + rng.uniform(-70, 70)
        sats.append(dict(prn=p, cn0=rng.uniform(46, 50), doppler=float(dop),
                         code_phase=float(rng.uniform(0, 1023 * 20)),
                         carr_phase=float(rng.uniform(0, 1)), rate=float(rng.uniform(-1, 1))))
    return sdr.make_ifgen(FS5, 0.0, CF, 2, sats, sigma=20.0, seed=seed), prns


def l1ca_code(sdr, prn):
    c, _ = sdr.gencode(prn, 1)
    return c.astype(np.int16)
