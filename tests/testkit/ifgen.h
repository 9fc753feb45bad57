/* ifgen.h - deterministic synthetic IF generator (host + device).
 *
 * The reference ships no IF recordings (test/testdata_download_link.txt only
 * lists URLs), so every parity test and benchmark runs on IF produced here.
 * One sample is a pure function of (configuration, sample index): the same
 * code is compiled as plain C for the CPU tools and as a CUDA device function
 * for the on-GPU generator kernel, so a 30 s / 491 MS stream can be made in
 * HBM in milliseconds.  CPU and GPU outputs may differ in isolated samples by
 * one LSB (libm vs CUDA cos/log rounding); parity tests therefore always feed
 * the SAME byte stream to both the oracle and the CUDA path.
 *
 * Signal model per satellite s (t = k / f_sf):
 *   chips(t) = code_phase + crate*(1 + fd/f_cf)*t + 0.5*crate*fdr/f_cf*t^2
 *   cyc(t)   = carr_phase + (f_if + fd)*t + 0.5*fdr*t^2
 *   real : x += A * D[bit] * c[chip] * cos(2*pi*cyc)
 *   IQ   : x += A * D[bit] * c[chip] * exp(-i*2*pi*cyc)   (the convention the
 *          reference's complex carrier wipe-off expects, src/sdrcmn.c:653-657)
 * plus white Gaussian noise of std `sigma` LSB per component, rounded to
 * nearest and clipped to [-127,127] (signed 8 bit, the format every reference
 * front end expands to, src/sdr.h:126-127).
 */
#ifndef SDRB200_IFGEN_H
#define SDRB200_IFGEN_H

#include <stdint.h>
#include <math.h>

#ifdef __CUDACC__
#define IFGEN_HD __host__ __device__ __forceinline__
#else
#define IFGEN_HD static inline
#endif

#define IFGEN_MAXSAT   16
#define IFGEN_MAXCLEN  1023

typedef struct {
    int      prn;              /* code number (index into the code table)   */
    int      nav_on;           /* 0: all +1, 1: 50 bps random data, 2: 50 bps GPS LNAV subframes
                                * (preamble, TOW count, valid (32,26) parity in every word),
                                * 3: GLONASS strings (random 10 ms units + the 30-unit time mark),
                                * 4: SBAS L1 messages (preamble, type 12, CRC-24Q, rate-1/2 K = 7 code) */
    double   cn0_dbhz;         /* carrier-to-noise density                  */
    double   doppler_hz;       /* carrier Doppler at t=0                    */
    double   doppler_rate;     /* Hz/s                                      */
    double   code_phase;       /* chips at sample 0, in [0, clen*bitlen)    */
    double   carr_phase;       /* cycles at sample 0                        */
    uint64_t nav_seed;         /* data bit b = hash(nav_seed, b) & 1        */
    double   carr_offset_hz;   /* added to the carrier only (FDMA channel offset): no code Doppler */
} ifgen_sat_t;

typedef struct {
    double   f_sf, f_if, f_cf; /* sampling, intermediate, RF carrier (Hz)   */
    double   crate;            /* chip rate (chip/s)                        */
    double   sigma;            /* noise std per component (LSB)             */
    int      dtype;            /* 1 real, 2 interleaved I/Q                 */
    int      clen;             /* chips per code period                     */
    int      bitlen;           /* code periods per data bit (20 for L1CA)   */
    int      nsat;
    uint64_t seed;
    ifgen_sat_t sat[IFGEN_MAXSAT];
    double   amp[IFGEN_MAXSAT];/* filled by ifgen_prepare()                 */
} ifgen_cfg_t;

IFGEN_HD uint64_t ifgen_mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

/* amplitude in LSB for a given C/N0 (see file header) */
static inline void ifgen_prepare(ifgen_cfg_t *c)
{
    int s;
    for (s = 0; s < c->nsat; s++) {
        double cn0 = pow(10.0, c->sat[s].cn0_dbhz / 10.0);
        c->amp[s] = (c->dtype == 2) ? c->sigma * sqrt(2.0 * cn0 / c->f_sf)
                                    : 2.0 * c->sigma * sqrt(cn0 / c->f_sf);
    }
}

IFGEN_HD int ifgen_navbit(uint64_t nav_seed, int64_t bit)
{
    return (ifgen_mix64(nav_seed ^ (uint64_t)bit * 0xD6E8FEB86659FD93ULL) & 1) ? 1 : -1;
}

/* nav_on == 2: bit b of an endless train of GPS LNAV subframes (300 bits: ten 30-bit words,
 * IS-GPS-200 20.3.5).  Word 1 starts with the preamble 10001011, word 2 carries a non-zero
 * 17-bit TOW count and a subframe ID cycling 1..3, the other data bits are pseudo-random.
 * The last two data bits of EVERY word are solved so that its parity bits D29 = D30 = 0 (what
 * the real message does in words 2 and 10), which makes every word independent of its
 * predecessor: a bit is a pure function of (seed, b).  Logical 1 is sent as -1, the sign
 * convention of the reference's checkbit()/paritycheck_l1ca() (src/sdrnav.c:265-270,
 * src/sdrnav_gps.c:141-166).                                                            */
IFGEN_HD int ifgen_parity32(uint32_t v) { v ^= v >> 16; v ^= v >> 8; v ^= v >> 4; return (0x6996u >> (v & 15u)) & 1u; }
IFGEN_HD int ifgen_lnavbit(uint64_t nav_seed, int64_t bit)
{
    const int64_t sf = bit / 300;
    const int r = (int)(bit - sf * 300), wi = r / 30, k = r - wi * 30;      /* word, bit in word */
    uint32_t w = (uint32_t)(ifgen_mix64(nav_seed ^ (uint64_t)(sf * 10 + wi) * 0xD6E8FEB86659FD93ULL) >> 20) & 0xFFFFFCu;
    uint32_t word;
    int d24, d23;
    if (wi == 0) w = (w & 0x00FFFFu) | (0x8Bu << 16);                       /* preamble 10001011 */
    if (wi == 1) {
        const uint32_t tow = (uint32_t)((sf + 1 + (int64_t)(nav_seed & 0xFFFF)) % 100799) + 1;   /* 1..100799 */
        w = (tow << 7) | ((uint32_t)(sf % 3 + 1) << 2);                     /* TOW | alert, AS = 0 | ID | t t */
    }
    d24 = ifgen_parity32(w & 0xAEC7CCu);                                    /* forces D29 = 0 */
    d23 = ifgen_parity32(w & 0x2DEA24u) ^ d24;                              /* forces D30 = 0 */
    w |= ((uint32_t)d23 << 1) | (uint32_t)d24;
    word = (w << 6) | ((uint32_t)ifgen_parity32(w & 0xEC7CD2u) << 5) | ((uint32_t)ifgen_parity32(w & 0x763E69u) << 4)
         | ((uint32_t)ifgen_parity32(w & 0xBB1F34u) << 3) | ((uint32_t)ifgen_parity32(w & 0x5D8F9Au) << 2);
    return ((word >> (29 - k)) & 1u) ? -1 : 1;
}

/* nav_on == 3: unit u (one 10 ms half-symbol, bitlen = 10 code periods) of an endless train of
 * GLONASS strings: 200 units = 2 s, the first 170 pseudo-random (the 85 data bits in meander
 * form are not decoded on this path), the last 30 the time mark the reference's findpreamble()
 * looks for at the end of its 200-unit register (pre_g1, src/sdrinit.c:494-496,
 * src/sdrnav.c:396-400).  nav_seed bit 16 set: the whole train is sent inverted.             */
IFGEN_HD int ifgen_g1bit(uint64_t nav_seed, int64_t unit)
{
    const int64_t st = (unit >= 0) ? unit / 200 : -((-unit + 199) / 200);
    const int r = (int)(unit - st * 200);
    const int inv = ((nav_seed >> 16) & 1) ? -1 : 1;
    if (r >= 170) {
        const unsigned int tm = 0x3E375096u;              /* 111110001101110101000010010110, MSB first */
        return (((tm >> (29 - (r - 170))) & 1u) ? -1 : 1) * inv;
    }
    return ifgen_navbit(nav_seed, unit) * inv;
}

/* nav_on == 4: symbol s (500 symbols/s, bitlen = 2 code periods) of an endless train of SBAS L1 messages as the
 * reference's receiver expects them (src/sdrnav.c:300-321, 347-356, 383-389; src/sdrnav_sbs.c:88-141): 250-bit
 * messages = 8-bit preamble (0x53, 0x9A, 0xC6 in turn), message type 12 (the only one whose decoding yields the
 * week, without which the reference drops its frame sync again), 212 pseudo-random bits, CRC-24Q over the 226 bits
 * right-aligned in 29 bytes; rate-1/2 K = 7 convolutional code in the order of the ka9q decoder the reference
 * links (encstate = encstate << 1 | bit; symbols parity(encstate & 0x6d), parity(encstate & 0x4f),
 * lib/ka9q-fec/fec.h:13-14, viterbi27_port.c:48-56); data bit / symbol 0 is sent as +1, 1 as -1
 * (enc[i] = fbits == 1 ? 0 : 255, src/sdrnav.c:304-305).  nav_seed bit 16 set: the train is sent inverted.  */
IFGEN_HD unsigned int ifgen_crc24q_bits(unsigned int crc, int bit)
{
    crc ^= (unsigned int)bit << 23;
    crc = (crc & 0x800000u) ? ((crc << 1) ^ 0x1864CFBu) : (crc << 1);
    return crc & 0xFFFFFFu;
}
IFGEN_HD int ifgen_sbas_databit(uint64_t seed, int64_t m, int j)          /* j < 226 */
{
    if (j < 8) { const unsigned int pre[3] = {0x53u, 0x9Au, 0xC6u}; return (int)((pre[(int)(m % 3)] >> (7 - j)) & 1u); }
    if (j < 14) return (12 >> (13 - j)) & 1;
    return (int)(ifgen_mix64(seed ^ (uint64_t)(m * 256 + j) * 0xD6E8FEB86659FD93ULL) & 1u);
}
IFGEN_HD int ifgen_sbas_msgbit(uint64_t seed, int64_t m, int j, unsigned int *crc_cache, int64_t *crc_m)
{
    if (j < 226) return ifgen_sbas_databit(seed, m, j);
    if (*crc_m != m) {
        unsigned int crc = 0;
        int i;
        for (i = 0; i < 6; i++) crc = ifgen_crc24q_bits(crc, 0);          /* right alignment: 6 leading zero bits */
        for (i = 0; i < 226; i++) crc = ifgen_crc24q_bits(crc, ifgen_sbas_databit(seed, m, i));
        *crc_cache = crc; *crc_m = m;
    }
    return (int)((*crc_cache >> (23 - (j - 226))) & 1u);
}
IFGEN_HD int ifgen_sbassym(uint64_t nav_seed, int64_t sym)
{
    const int inv = ((nav_seed >> 16) & 1) ? -1 : 1;
    unsigned int reg = 0, crc = 0, v;
    int64_t crc_m = -1, b, i;
    int which;
    if (sym < 0) return inv;
    b = sym >> 1; which = (int)(sym & 1);
    for (i = b - 6; i <= b; i++) {
        int bit = 0;
        if (i >= 0) { const int64_t m = i / 250; bit = ifgen_sbas_msgbit(nav_seed, m, (int)(i - m * 250), &crc, &crc_m); }
        reg = (reg << 1) | (unsigned int)bit;
    }
    v = reg & (which ? 0x4Fu : 0x6Du);
    return ifgen_parity32(v) ? -inv : inv;
}

/* one sample; codes = [nsat][IFGEN_MAXCLEN] of +-1; out[0]=I (or real), out[1]=Q */
IFGEN_HD void ifgen_sample(const ifgen_cfg_t *c, const signed char *codes,
                           uint64_t k, signed char *out)
{
    const double t = (double)k / c->f_sf;
    double xi = 0.0, xq = 0.0;
    int s;
    for (s = 0; s < c->nsat; s++) {
        const ifgen_sat_t *sv = &c->sat[s];
        double fd = sv->doppler_hz, fdr = sv->doppler_rate;
        double chips = sv->code_phase + c->crate * (1.0 + fd / c->f_cf) * t
                     + 0.5 * c->crate * fdr / c->f_cf * t * t;
        double per = floor(chips / (double)c->clen);            /* code period # */
        int    chip = (int)(chips - per * (double)c->clen);
        double cyc, fr, a;
        int    d = 1;
        if (chip >= c->clen) chip = c->clen - 1;
        if (chip < 0) chip = 0;
        if (sv->nav_on == 4) d = ifgen_sbassym(sv->nav_seed, (int64_t)floor(per / (double)c->bitlen));
        else if (sv->nav_on == 3) d = ifgen_g1bit(sv->nav_seed, (int64_t)floor(per / (double)c->bitlen));
        else if (sv->nav_on == 2) d = ifgen_lnavbit(sv->nav_seed, (int64_t)floor(per / (double)c->bitlen));
        else if (sv->nav_on) d = ifgen_navbit(sv->nav_seed, (int64_t)floor(per / (double)c->bitlen));
        cyc = sv->carr_phase + (c->f_if + fd + sv->carr_offset_hz) * t + 0.5 * fdr * t * t;
        fr  = cyc - floor(cyc);
        a   = c->amp[s] * (double)(d * codes[s * IFGEN_MAXCLEN + chip]);
        xi += a * cos(6.283185307179586 * fr);
        if (c->dtype == 2) xq -= a * sin(6.283185307179586 * fr);
    }
    {
        uint64_t h1 = ifgen_mix64(c->seed ^ (2 * k) * 0x9E3779B97F4A7C15ULL);
        uint64_t h2 = ifgen_mix64(c->seed ^ (2 * k + 1) * 0x9E3779B97F4A7C15ULL);
        double u1 = ((double)(h1 >> 11) + 1.0) * (1.0 / 9007199254740992.0);
        double u2 = (double)(h2 >> 11) * (1.0 / 9007199254740992.0);
        double r  = c->sigma * sqrt(-2.0 * log(u1));
        double v;
        xi += r * cos(6.283185307179586 * u2);
        xq += r * sin(6.283185307179586 * u2);
        v = rint(xi); if (v > 127.0) v = 127.0; if (v < -127.0) v = -127.0;
        out[0] = (signed char)v;
        if (c->dtype == 2) {
            v = rint(xq); if (v > 127.0) v = 127.0; if (v < -127.0) v = -127.0;
            out[1] = (signed char)v;
        }
    }
}

#endif /* SDRB200_IFGEN_H */
