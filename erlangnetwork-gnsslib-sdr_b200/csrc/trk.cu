/* trk.cu - tracking correlator: single-call primitive and batched engine.
 *
 * sdrb200_correlator  <-> correlator()  src/sdrcmn.c:687-722
 * sdrb200_trk_*       <-> sdrtracking() src/sdrtrk.c:15-54 + cumsumcorr/pll/dll
 *                         src/sdrtrk.c:64-150 + bit sync src/sdrnav.c:15-41,198-282
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <mutex>
#include "sdrb200_internal.h"
#include "trk_device.cuh"
#include "trk_dot.cuh"
#include "trk_cluster.cuh"

static std::once_flag g_lut_once;
static int g_lut_rc = SDRB200_OK;

/* re-derive the carrier table exactly as mixcarr builds it (src/sdrcmn.c:641-646)
 * and compare with the constants compiled into the kernels */
static void lut_init(void)
{
    static const int8_t kc[32] = SDRB200_COST_INIT, ks[32] = SDRB200_SINT_INIT;
    const double DPI = 2.0 * 3.1415926535897932;
    for (int i = 0; i < 32; i++) {
        short c = (short)floor((cos(DPI / 32 * i) / SDRB200_CSCALE + 0.5));
        short s = (short)floor((sin(DPI / 32 * i) / SDRB200_CSCALE + 0.5));
        if (c != kc[i] || s != ks[i]) {
            sdrb200_set_error("carrier table mismatch at %d: libm gives (%d,%d)", i, c, s);
            g_lut_rc = SDRB200_EINVAL;
        }
    }
}
extern "C" int sdrb200_trk_lut_ready(void)
{
    std::call_once(g_lut_once, lut_init);
    return g_lut_rc;
}

/* ------------------------------------------------------------------------
 * single-call kernel: one CTA, one code period
 * out: II[ntap], QQ[ntap], remc, remp  (doubles)
 * ---------------------------------------------------------------------- */
template <int DTYPE>
__global__ void __launch_bounds__(TRK_THREADS, 1)
k_correlator(const int8_t *__restrict__ data, double ti, int n, double freq, double phi0,
             double crate, double coff, const int *__restrict__ taps_g, int ns,
             const int8_t *__restrict__ code_g, int coden, int qcap, double *__restrict__ out,
             int *__restrict__ err)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    trk_smem_t *sm = (trk_smem_t *)smem_raw;
    int *taps = (int *)(smem_raw + sizeof(trk_smem_t));
    const trk_work_t wk = trk_work_layout<DTYPE>((unsigned int *)(taps + SDRB200_MAXCORRN), n, qcap);
    const bool worker = threadIdx.x >= 64;

    trk_init_smem(sm);
    for (int i = threadIdx.x; i < coden; i += blockDim.x) sm->codetab[i] = code_g[i];
    if (threadIdx.x < ns) taps[threadIdx.x] = taps_g[threadIdx.x];
    __syncthreads();
    const int smax = taps[ns - 1];                                   /* sdrcmn.c:694 */

    trk_build_tables(sm, ti, n, freq, phi0, crate, coff, smax, coden);
    __syncthreads();
    if (worker) trk_mix_prefix<DTYPE>(sm, data, n, wk);
    else trk_carrier_rem(sm);
    __syncthreads();
    if (worker) trk_tap_sums<DTYPE>(sm, wk, data, n, taps, ns, smax);
    __syncthreads();

    const int ntap = 1 + 2 * ns;
    if (threadIdx.x < ntap) {
        out[threadIdx.x]        = (double)(int)sm->sum[threadIdx.x] * SDRB200_CSCALE;
        out[ntap + threadIdx.x] = (double)(int)sm->sum[SDRB200_MAXTAPS + threadIdx.x] * SDRB200_CSCALE;
    }
    if (threadIdx.x == 0) {
        out[2 * ntap] = sm->remc;
        out[2 * ntap + 1] = sm->remp;
        *err = sm->err;
    }
}

/* chip transitions inside n + 2*smax samples, with head room */
static int trk_qcap(int n, int smax, double chips_per_sample)
{
    return (int)((double)(n + 2 * smax) * chips_per_sample * 1.01) + 16;
}
static size_t corr_smem_bytes(int n, int dtype, int qcap)
{
    return sizeof(trk_smem_t) + sizeof(int) * SDRB200_MAXCORRN + sizeof(unsigned int) * TRK_WORK_WORDS(n, dtype, qcap);
}
#define TRK_SMEM_LIMIT (225 * 1024)

extern "C" int sdrb200_correlator(const int8_t *data, int dtype, double ti, int n, double freq,
                                  double phi0, double crate, double coff, const int *s, int ns,
                                  double *II, double *QQ, double *remc, double *remp,
                                  const short *code, int coden)
{
    if (!data || !s || !II || !QQ || !remc || !remp || !code || n < 1 || ns < 1 ||
        ns > SDRB200_MAXCORRN || coden < 1 || coden > 1024 || (dtype != 1 && dtype != 2)) {
        sdrb200_set_error("correlator: bad argument"); return SDRB200_EINVAL;
    }
    const int smax = s[ns - 1];
    if (smax < 0) { sdrb200_set_error("correlator: negative tap"); return SDRB200_EINVAL; }
    if (!(ti * crate > 0) || ti * crate > 0.5) { sdrb200_set_error("correlator: needs at least two samples per chip"); return SDRB200_EINVAL; }
    const int qcap = trk_qcap(n, smax, ti * crate);
    const size_t smem = corr_smem_bytes(n, dtype, qcap);
    if (smem > TRK_SMEM_LIMIT) { sdrb200_set_error("correlator: window of %d samples too long for one CTA", n); return SDRB200_EINVAL; }
    const int ntap = 1 + 2 * ns;
    const size_t nbytes = (size_t)n * dtype;
    /* device arena: [16 guard][data][32 guard] | code | taps | out | err */
    const size_t off_data = 256, off_code = off_data + ((nbytes + 64 + 255) & ~(size_t)255);
    const size_t off_taps = off_code + 1024, off_out = off_taps + 256, off_err = off_out + 1024;
    const size_t h_out = 2048;
    sdrb200_ws_t *w = sdrb200_ws(off_err + 256, h_out + 1024 + 256);
    if (!w) return SDRB200_ENODEV;
    int rc = sdrb200_trk_lut_ready();
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;

    int8_t *h_code = (int8_t *)(w->h_arena + h_out);
    for (int i = 0; i < coden; i++) {
        if (code[i] < -127 || code[i] > 127) { sdrb200_set_error("correlator: code value out of int8 range"); return SDRB200_EINVAL; }
        h_code[i] = (int8_t)code[i];
    }
    int *h_taps = (int *)(w->h_arena + h_out + 1024);
    memcpy(h_taps, s, sizeof(int) * ns);

    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena + off_data, data, nbytes, cudaMemcpyHostToDevice, st));
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena + off_code, h_code, coden, cudaMemcpyHostToDevice, st));
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena + off_taps, h_taps, sizeof(int) * ns, cudaMemcpyHostToDevice, st));

    double *d_out = (double *)(w->d_arena + off_out);
    int *d_err = (int *)(w->d_arena + off_err);
#define CORR_LAUNCH(DT)                                                                                   \
    do {                                                                                                  \
        SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_correlator<DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        k_correlator<DT><<<1, TRK_THREADS, smem, st>>>((const int8_t *)(w->d_arena + off_data), ti, n, freq, phi0, \
            crate, coff, (const int *)(w->d_arena + off_taps), ns, (const int8_t *)(w->d_arena + off_code), coden, qcap, d_out, d_err); \
    } while (0)
    if (dtype == 1) CORR_LAUNCH(1); else CORR_LAUNCH(2);
#undef CORR_LAUNCH
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    double *h = (double *)w->h_arena;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(h, d_out, sizeof(double) * (2 * ntap + 2), cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaMemcpyAsync(h + 2 * ntap + 2, d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    if (*(int *)(h + 2 * ntap + 2)) { sdrb200_set_error("correlator: NCO segment table overflow"); return SDRB200_ERANGE; }
    memcpy(II, h, sizeof(double) * ntap);
    memcpy(QQ, h + ntap, sizeof(double) * ntap);
    *remc = h[2 * ntap];
    *remp = h[2 * ntap + 1];
    return SDRB200_OK;
}

/* ========================================================================
 * Batched tracking engine: one CTA per channel, persistent over code periods.
 * The per-period scalar epilogue restates, on the device:
 *   sdrtracking bookkeeping      src/sdrtrk.c:31-38 (currnsamp, old* copies)
 *   sdrnavigation bit sync       src/sdrnav.c:15-41
 *   checksync / checkbit         src/sdrnav.c:198-233 / 241-282
 *   cumsumcorr / clearcumsumcorr src/sdrtrk.c:64-86
 *   pll / dll                    src/sdrtrk.c:94-150
 *   loop scheduling              src/sdrmain.c:269-312
 * ====================================================================== */
#define TRK_NAVSYNCTH 50                       /* NAVSYNCTH, src/sdr.h:157 */
#define TRK_PI 3.1415926535897932              /* PI, src/sdr.h:103        */

struct trk_chan_dev_t {
    double f_sf, ti, f_if, f_cf, foffset, crate, ctime, acqfreq;
    uint64_t buffloc0;
    int32_t dtype, clen, nsamp, prn, corrn, ne, nl, loopms, navrate, pad;
    int32_t corrp[SDRB200_MAXCORRN];
    sdrb200_loop_t prm1, prm2;
    int8_t code[1024];
};

struct trk_state_t {
    uint64_t buffloc, cnt;
    double codefreq, carrfreq, remcode, remcarr, oldremcode, oldremcarr;
    double codeNco, codeErr, carrNco, carrErr, freqErr;
    double II[SDRB200_MAXTAPS], QQ[SDRB200_MAXTAPS], oldI[SDRB200_MAXTAPS], oldQ[SDRB200_MAXTAPS];
    double sumI[SDRB200_MAXTAPS], sumQ[SDRB200_MAXTAPS], oldsumI[SDRB200_MAXTAPS], oldsumQ[SDRB200_MAXTAPS];
    /* nav (sdrnav_t subset, src/sdr.h:440-479) */
    double bitIP;
    int32_t biti, ocodei, synci, flagsync, swsync, swreset, swloop, navcnt, bit, flagloopfilter;
    int32_t currnsamp, pad;
    int32_t bitsync[32];
    /* observables: slot [0] of sdrtrk_t L/D/S/remcout/codei/cntout, Isum (src/sdr.h:384-400),
     * loopcnt of the channel thread (src/sdrmain.c:236,304) */
    double L0, D0, S0, remcout0, Isum;
    uint64_t obs_codei, obs_cnt;
    int64_t obs_dcnt;
    int32_t flagremcarradd, loopcnt;
    /* frame synchronisation (GPS L1CA): nav.fbits as a ring of 2 + 300 bits (fhead = oldest),
     * flagsyncf, flagtow, polarity, firstsf, firstsfcnt (src/sdr.h:440-479) */
    uint64_t firstsf, firstsfcnt;
    int32_t fhead, flagsyncf, flagtow, polarity, flagpolarityadd, flagpol;
    int8_t  fbits[1520];                     /* 302 (L1CA), 200 (G1) or 1512 (SBAS L1) entries in use */
};

/* checksync(), src/sdrnav.c:198-233 */
__device__ static int trk_checksync(double IP, double IPold, trk_state_t *st, int prn, int rate)
{
    if (prn > 5) {
        int corr = 0;
        for (int i = 0; i < rate - 1; i++) st->bitsync[i] = st->bitsync[i + 1];
        st->bitsync[rate - 1] = (IP < 0 ? -1 : 1);
        for (int i = 0; i < rate; i++) corr += st->bitsync[i];      /* overlay code is all ones */
        if (abs(corr) == rate) { st->synci = st->biti; return 1; }
    } else {
        if (IPold * IP < 0) {
            st->bitsync[st->biti] += 1;
            int maxv = st->bitsync[0], ind = 0;                     /* maxvi(), sdrcmn.c:417-431 */
            for (int i = 1; i < rate; i++) if (maxv < st->bitsync[i]) { maxv = st->bitsync[i]; ind = i; }
            st->synci = ind;
            if (maxv > TRK_NAVSYNCTH) {
                st->synci--;
                if (st->synci < 0) st->synci = rate - 1;
                return 1;
            }
        }
    }
    return 0;
}

/* length of nav.fbits (flen + addflen, src/sdr.h:159-174): 302 for GPS L1CA, 200 for GLONASS G1
 * 1512 for SBAS L1 (told apart by nav.rate: 20, 10, 2 code periods per unit) */
__device__ __forceinline__ int trk_fbits_len(int rate) { return rate == 10 ? 200 : (rate == 2 ? 1512 : 302); }

/* checkbit(), src/sdrnav.c:241-282; nav.fbits is a ring (fhead = oldest) */
__device__ static void trk_checkbit(double IP, int loopms, trk_state_t *st, int rate)
{
    int diffi = st->biti - st->synci;
    st->swreset = 0; st->swsync = 0;
    if (diffi == 1 || diffi == -rate + 1) {
        st->bitIP = IP; st->swreset = 1; st->navcnt = 1;
    } else {
        st->bitIP += IP;
    }
    st->swloop = (st->navcnt % loopms == 0) ? 1 : 0;
    if (diffi == 0) {
        const int pol = st->flagpol ? -1 : 1;                 /* set by a failed SBAS parity check, sdrnav.c:264-270, 405-407 */
        st->bit = (st->bitIP < 0) ? -pol : pol;
        st->fbits[st->fhead] = (int8_t)st->bit;               /* shift left + append, sdrnav.c:272-275 */
        st->fhead = (st->fhead + 1 == trk_fbits_len(rate)) ? 0 : st->fhead + 1;
        st->swsync = 1;
    }
    st->navcnt++;
}

/* paritycheck_l1ca(), src/sdrnav_gps.c:141-166 on +-1 values (b[0..1]: D29*, D30*) */
__device__ static int trk_parity_l1ca(const int *b)
{
    int p[6], stat = 0;
    p[0] = b[0]*b[2]*b[3]*b[4]*b[6]*b[7]*b[11]*b[12]*b[13]*b[14]*b[15]*b[18]*b[19]*b[21]*b[24];
    p[1] = b[1]*b[3]*b[4]*b[5]*b[7]*b[8]*b[12]*b[13]*b[14]*b[15]*b[16]*b[19]*b[20]*b[22]*b[25];
    p[2] = b[0]*b[2]*b[4]*b[5]*b[6]*b[8]*b[9]*b[13]*b[14]*b[15]*b[16]*b[17]*b[20]*b[21]*b[23];
    p[3] = b[1]*b[3]*b[5]*b[6]*b[7]*b[9]*b[10]*b[14]*b[15]*b[16]*b[17]*b[18]*b[21]*b[22]*b[24];
    p[4] = b[1]*b[2]*b[4]*b[6]*b[7]*b[8]*b[10]*b[11]*b[15]*b[16]*b[17]*b[18]*b[19]*b[22]*b[23]*b[25];
    p[5] = b[0]*b[4]*b[6]*b[7]*b[9]*b[10]*b[11]*b[12]*b[14]*b[16]*b[20]*b[23]*b[24]*b[25];
    for (int i = 0; i < 6; i++) stat += p[i] - b[26 + i];
    return stat == 0;
}

/* findpreamble() + paritycheck() for GPS L1CA, src/sdrnav.c:373-411, 324-345 (predecodefec is
 * a plain copy for L1CA, :296-299); called by one lane when a data bit has been decided     */
__device__ __noinline__ static int trk_frame_sync_l1ca(trk_state_t *st)
{
    const int pre[8] = {1, -1, -1, -1, 1, -1, 1, 1};                      /* src/sdrinit.c:492 */
    int corr = 0;
    for (int i = 0; i < 8; i++) { int k = st->fhead + 2 + i; if (k >= 302) k -= 302; corr += (int)st->fbits[k] * pre[i]; }
    if (abs(corr) != 8) return 0;
    st->polarity = corr > 0 ? 1 : -1;                                     /* kept even if the parity fails */
    int stat = 0, prev29 = 0, prev30 = 0;
    {
        int k0 = st->fhead, k1 = st->fhead + 1; if (k1 >= 302) k1 -= 302;
        prev29 = st->polarity * (int)st->fbits[k0]; prev30 = st->polarity * (int)st->fbits[k1];
    }
    for (int w = 0; w < 10; w++) {
        int b[32];
        b[0] = prev29; b[1] = prev30;
        for (int j = 0; j < 30; j++) {
            int k = st->fhead + 2 + w * 30 + j; if (k >= 302) k -= 302;
            b[2 + j] = st->polarity * (int)st->fbits[k];
        }
        /* the reference inverts bits 2..25 of its working copy in place; bits 30, 31 of this word
         * (the next word's D29*, D30*) are parity bits and stay as received                   */
        if (b[1] == -1) for (int j = 2; j < 26; j++) b[j] = -b[j];
        stat += trk_parity_l1ca(b);
        prev29 = b[30]; prev30 = b[31];
    }
    return stat == 10;
}

/* findpreamble() for GLONASS G1, src/sdrnav.c:394-400: the 30-unit time mark closes the string,
 * so it is looked for at the END of the 200-unit register (predecodefec is a plain copy, :296-299;
 * paritycheck() accepts every G1 string, :357-359)                                              */
__device__ __noinline__ static int trk_frame_sync_g1(trk_state_t *st)
{
    const unsigned int tm = 0x3E375096u;                  /* pre_g1, src/sdrinit.c:494-496: bit set = -1, first unit in bit 29 */
    int corr = 0;
    for (int i = 0; i < 30; i++) {
        int k = st->fhead + 170 + i; if (k >= 200) k -= 200;
        corr += (int)st->fbits[k] * (((tm >> (29 - i)) & 1u) ? -1 : 1);
    }
    if (abs(corr) != 30) return 0;
    st->polarity = corr > 0 ? 1 : -1;
    return 1;
}

/* SBAS L1 frame synchronisation by the whole epilogue warp: predecodefec() - the K = 7 rate-1/2 Viterbi
 * decoder of lib/ka9q-fec/viterbi27_port.c over the 1512 symbols of nav.fbits, started from scratch at
 * every symbol as the reference does (src/sdrnav.c:300-321) - then findpreamble() for two consecutive
 * preambles (:383-389) and paritycheck() with CRC-24Q (:347-356).
 * One lane per butterfly (viterbi27_port.c:127-141): lane i holds the path metrics of states i and
 * i + 32, produces those of 2i and 2i + 1 and their two decision bits; the metrics move to their
 * next owners with four shuffles, the decision words (states 0..31 / 32..63, decision_t.w[0..1] of
 * which only the low 32 bits are ever set) are OR-reduced into vdec[step][2].  unsigned arithmetic and
 * the signed comparison exactly as in the C source; no renormalisation there either (756 x 510 fits).
 * Chain-back from state 0 by lane 0 (:76-104).  Returns findpreamble()'s value; sets polarity, flagpol. */
__device__ __noinline__ static int trk_frame_sync_sbas(trk_state_t *st, unsigned int *vdec, int lane)
{
    constexpr int NSYM = 1512, NSTEP = NSYM / 2, NBITS = 750;
    const unsigned int b0 = (__popc((2u * lane) & 0x6Du) & 1) ? 255u : 0u;      /* set_viterbi27_polynomial_port */
    const unsigned int b1 = (__popc((2u * lane) & 0x4Fu) & 1) ? 255u : 0u;
    unsigned int lo = (lane == 0) ? 0u : 63u, hi = 63u;                         /* init_viterbi27_port, start state 0 */
    int k = st->fhead;
    for (int step = 0; step < NSTEP; step++) {
        const unsigned int s0 = (st->fbits[k] == 1) ? 0u : 255u;                /* enc[i] = fbits == 1 ? 0 : 255 */
        k = (k + 1 == NSYM) ? 0 : k + 1;
        const unsigned int s1 = (st->fbits[k] == 1) ? 0u : 255u;
        k = (k + 1 == NSYM) ? 0 : k + 1;
        const unsigned int metric = (b0 ^ s0) + (b1 ^ s1);
        unsigned int m0 = lo + metric, m1 = hi + (510u - metric);
        const unsigned int d0 = ((int)(m0 - m1) > 0) ? 1u : 0u;
        const unsigned int n0 = d0 ? m1 : m0;
        m0 -= (metric + metric - 510u);
        m1 += (metric + metric - 510u);
        const unsigned int d1 = ((int)(m0 - m1) > 0) ? 1u : 0u;
        const unsigned int n1 = d1 ? m1 : m0;
        const unsigned int x = (d0 << ((2 * lane) & 31)) | (d1 << ((2 * lane + 1) & 31));
        const unsigned int w0 = __reduce_or_sync(0xffffffffu, lane < 16 ? x : 0u);
        const unsigned int w1 = __reduce_or_sync(0xffffffffu, lane >= 16 ? x : 0u);
        if (lane == 0) { vdec[2 * step] = w0; vdec[2 * step + 1] = w1; }
        const int src = lane >> 1;
        const unsigned int a0 = __shfl_sync(0xffffffffu, n0, src), a1 = __shfl_sync(0xffffffffu, n1, src);
        const unsigned int c0 = __shfl_sync(0xffffffffu, n0, 16 + src), c1 = __shfl_sync(0xffffffffu, n1, 16 + src);
        lo = (lane & 1) ? a1 : a0;
        hi = (lane & 1) ? c1 : c0;
    }
    __syncwarp();
    int found = 0;
    if (lane == 0) {
        /* chainback_viterbi27_port(fec, dec, 750, 0): decisions are read 6 steps past the data bit */
        unsigned int dbits[24];                                                 /* decoded bit n at dbits[n >> 5] bit n & 31 */
        for (int i = 0; i < 24; i++) dbits[i] = 0u;
        unsigned int e = 0u;
        for (int n = NBITS - 1; n >= 0; n--) {
            const unsigned int sidx = e >> 2;
            const unsigned int kk = (vdec[2 * (n + 6) + (sidx >> 5)] >> (sidx & 31u)) & 1u;
            e = (e >> 1) | (kk << 7);
            dbits[n >> 5] |= kk << (n & 31);
        }
        auto bit = [&](int n) -> int { return (int)((dbits[n >> 5] >> (n & 31)) & 1u); };     /* fbitsdec = bit ? -1 : 1 */
        const int pre[16] = {1, -1, 1, -1, 1, 1, -1, -1, -1, 1, 1, -1, -1, 1, -1, 1};         /* pre_sbs[0..15], src/sdrinit.c:498-499 */
        int corr = 0;
        for (int i = 0; i < 8; i++) {
            corr += (bit(i) ? -1 : 1) * pre[i];
            corr += (bit(250 + i) ? -1 : 1) * pre[8 + i];
        }
        if (abs(corr) == 16) {
            st->polarity = corr > 0 ? 1 : -1;
            /* paritycheck(): the 226 message bits right-aligned in 29 bytes, CRC-24Q against the next 24;
             * a byte's bit is 1 where polarity * fbitsdec < 0 (bits2byte, src/sdrnav.c:154-171)         */
            const unsigned int flip = (st->polarity < 0) ? 1u : 0u;
            unsigned int crc = 0u, want = 0u;
            for (int i = 0; i < 6 + 226; i++) {
                const unsigned int b = (i < 6) ? 0u : ((unsigned int)bit(i - 6) ^ flip);
                crc ^= b << 23;
                crc = (crc & 0x800000u) ? ((crc << 1) ^ 0x1864CFBu) : (crc << 1);
                crc &= 0xFFFFFFu;
            }
            for (int i = 0; i < 24; i++) want = (want << 1) | ((unsigned int)bit(226 + i) ^ flip);
            if (crc == want) found = 1;
            else if (st->polarity == 1) st->flagpol = 1;
        }
    }
    return __shfl_sync(0xffffffffu, found, 0);
}

__device__ __forceinline__ long long trk_clock()
{
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory");
    return t;
}

/* pll() + dll(), src/sdrtrk.c:94-125 / 133-150, by one warp: the five libm calls
 * (atan2, 2 x atan, 2 x sqrt) are independent of each other and run on five
 * lanes at once; lane 0 then finishes the scalar updates in the reference's
 * order of operations.                                                        */
__device__ static void trk_loops_warp(trk_state_t *st, const sdrb200_loop_t *prm, double dt,
                                      const trk_chan_dev_t *ch, int lane)
{
    const double IP = st->sumI[0], QP = st->sumQ[0], oldIP = st->oldsumI[0], oldQP = st->oldsumQ[0];
    const double IE = st->sumI[ch->ne], IL = st->sumI[ch->nl], QE = st->sumQ[ch->ne], QL = st->sumQ[ch->nl];
    double r = 0.0;
    if (lane == 0)      r = (IP > 0) ? atan2(QP, IP) : atan2(-QP, -IP);
    else if (lane == 1) r = (IP == 0) ? TRK_PI / 2 : atan(QP / IP);
    else if (lane == 2) r = (oldIP == 0) ? TRK_PI / 2 : atan(oldQP / oldIP);
    else if (lane == 3) r = sqrt(IE * IE + QE * QE);
    else if (lane == 4) r = sqrt(IL * IL + QL * QL);
    const double f1 = __shfl_sync(0xffffffffu, r, 1), f2 = __shfl_sync(0xffffffffu, r, 2);
    const double e = __shfl_sync(0xffffffffu, r, 3), l = __shfl_sync(0xffffffffu, r, 4);
    if (lane == 0) {
        const double carrErr = r / TRK_PI;
        double freqErr = f1 - f2;
        if (freqErr > TRK_PI / 2) freqErr = TRK_PI - freqErr;
        if (freqErr < -TRK_PI / 2) freqErr = -TRK_PI - freqErr;
        st->carrNco += prm->pllaw * (carrErr - st->carrErr) + prm->pllw2 * dt * carrErr + prm->fllw * dt * freqErr;
        st->carrfreq = ch->acqfreq + st->carrNco;
        st->carrErr = carrErr;
        st->freqErr = freqErr;
        const double codeErr = (e - l) / (e + l);
        st->codeNco += prm->dllaw * (codeErr - st->codeErr) + prm->dllw2 * dt * codeErr;
        st->codefreq = ch->crate - st->codeNco + (st->carrfreq - ch->f_if - ch->foffset) / (ch->f_cf / ch->crate);
        st->codeErr = codeErr;
    }
}

/* the interpolation stocks of setobsdata (src/sdrtrk.c:160-209, src/sdr.h:384-391) per channel, in
 * global memory, as RINGS: the reference shifts 80-element arrays on every call, here the newest
 * element sits at `head` and a fetch unrolls the ring newest-first                            */
struct trk_obsstock_t {
    double   L[SDRB200_OBSINTERPN], D[SDRB200_OBSINTERPN], remcout[SDRB200_OBSINTERPN], S[SDRB200_OBSINTERPN];
    uint64_t codei[SDRB200_OBSINTERPN], cntout[SDRB200_OBSINTERPN], codeisum[SDRB200_OBSINTERPN];
    int32_t  head, shead;
};

struct trk_lf_t { double carrfreq, codefreq; unsigned long long cnt; int32_t navcnt, synci, flagsync, pad; };

/* epilogue of one code period, executed by ONE warp: everything sdrtracking and
 * the sdrthread loop body do around correlator()                              */
__device__ static void trk_epilogue(const trk_chan_dev_t *ch, trk_state_t *st, const unsigned int *sumI,
                                    const unsigned int *sumQ, double remc, double remp,
                                    sdrb200_trk_rec_t *rec, int lane, long long *ep = nullptr,
                                    trk_lf_t *lf = nullptr, volatile int *lf_flag = nullptr, int lf_val = 0,
                                    trk_obsstock_t *obs = nullptr, unsigned int *vdec = nullptr)
{
    long long e0 = 0;
#define TRK_EP(i) do { if (ep && lane == 0) { const long long e1 = trk_clock(); ep[i] += e1 - e0; e0 = e1; } } while (0)
    if (ep && lane == 0) e0 = trk_clock();
    const int ntap = 1 + 2 * ch->corrn;
    /* src/sdrtrk.c:35-38: the memcpy length 1+2*corrn*8 bytes covers 2*corrn doubles
     * (plus one stray byte of the next, which no consumer reads)               */
    for (int i = lane; i < 2 * ch->corrn; i += 32) { st->oldI[i] = st->II[i]; st->oldQ[i] = st->QQ[i]; }
    __syncwarp();
    /* src/sdrtrk.c:40-43: correlator's II (cos sums) lands in trk.QQ, QQ (sin sums) in trk.II */
    for (int i = lane; i < ntap; i += 32) {
        st->QQ[i] = (double)(int)sumI[i] * SDRB200_CSCALE;
        st->II[i] = (double)(int)sumQ[i] * SDRB200_CSCALE;
    }
    __syncwarp();
    if (lane == 0) {
        st->oldremcode = st->remcode;
        st->oldremcarr = st->remcarr;
        st->remcode = remc;
        st->remcarr = remp;
        /* sdrnavigation(), src/sdrnav.c:19-41 */
        const int rate = ch->navrate;
        /* cnt % rate: 32-bit while cnt fits (a 64-bit remainder is a subroutine) */
        st->biti = (st->cnt >> 32) ? (int)(st->cnt % (uint64_t)rate) : (int)((uint32_t)st->cnt % (uint32_t)rate);
        st->ocodei = st->biti - st->synci - 1;
        if (st->ocodei < 0) st->ocodei += rate;
        if (rate == 1 || !st->flagsync) {               /* the 2 s threshold matters only here */
            const double cntth = 2000 / (ch->ctime * 1000);
            if (rate == 1 && (double)st->cnt > cntth) { st->synci = 0; st->flagsync = 1; }
            if (!st->flagsync && (double)st->cnt > cntth)
                st->flagsync = trk_checksync(st->II[0], st->oldI[0], st, ch->prn, rate);
        }
        if (st->flagsync) {
            trk_checkbit(st->II[0], ch->loopms, st, rate);
            /* frame synchronisation, src/sdrnav.c:43-58 (GPS L1CA: 20 code periods per bit, subframe
             * preamble + parity; GLONASS G1: 10 code periods per unit, time mark of the string).  What
             * follows in the reference - decodenav and its reset when no TOW decodes, :60-83 -
             * is the host's: the record carries flagsyncf, polarity, firstsf, firstsfcnt       */
            if (st->swsync && (rate == 20 || rate == 10)) {
                if (!st->flagtow) st->flagsyncf = (rate == 20) ? trk_frame_sync_l1ca(st) : trk_frame_sync_g1(st);
                if (st->flagsyncf && !st->flagtow) { st->firstsf = st->buffloc; st->firstsfcnt = st->cnt; st->flagtow = 1; }
            }
        }
    }
    __syncwarp();
    /* SBAS L1 (2 code periods per symbol): the same place in sdrnavigation(), but the Viterbi decoder is work
     * for the whole warp */
    if (ch->navrate == 2 && vdec) {
        const int need = __shfl_sync(0xffffffffu, (st->flagsync && st->swsync && !st->flagtow) ? 1 : 0, 0);
        if (need) {
            const int r = trk_frame_sync_sbas(st, vdec, lane);
            if (lane == 0) {
                st->flagsyncf = r;
                if (r) { st->firstsf = st->buffloc; st->firstsfcnt = st->cnt; st->flagtow = 1; }
            }
            __syncwarp();
        }
    }
    TRK_EP(12);
    /* cumsumcorr(), src/sdrtrk.c:64-76 (overlay polarity is +1 for L1CA/SBAS/G1) */
    for (int i = lane; i < ntap; i += 32) {
        st->oldsumI[i] += st->oldI[i]; st->oldsumQ[i] += st->oldQ[i];
        st->sumI[i] += st->II[i];      st->sumQ[i] += st->QQ[i];
    }
    __syncwarp();
    /* src/sdrmain.c:271-280 */
    const int mode = !st->flagsync ? 1 : (st->swloop ? 2 : 0);
    if (mode == 1) trk_loops_warp(st, &ch->prm1, ch->ctime, ch, lane);
    else if (mode == 2) trk_loops_warp(st, &ch->prm2, (double)ch->loopms / 1000, ch, lane);
    __syncwarp();
    /* the NCO lanes wait for exactly this: new frequencies and the counters of the next period */
    if (lf && lane == 0) {
        lf->carrfreq = st->carrfreq; lf->codefreq = st->codefreq; lf->cnt = st->cnt + 1;
        lf->navcnt = st->navcnt; lf->synci = st->synci; lf->flagsync = st->flagsync;
        __threadfence_block();
        *lf_flag = lf_val;
    }
    TRK_EP(13);
    /* setobsdata(), src/sdrtrk.c:160-209, called after the loop filters of a synchronised
     * channel (src/sdrmain.c:285-304).  Only slot [0] of the interpolation stocks lives on
     * the device.                                                                        */
    int flagobs = 0;
    if (mode == 2 && lane == 0) {
        flagobs = (st->loopcnt % (100 / ch->loopms) == 0) ? 2 : 1;     /* SNSMOOTHMS, src/sdr.h:198 */
        st->obs_codei = st->buffloc;
        st->obs_cnt = st->cnt;
        st->obs_dcnt = (int64_t)(st->cnt - st->firstsfcnt);       /* (cnt - nav.firstsfcnt) of trk.tow[0] */
        st->remcout0 = st->oldremcode * ch->f_sf / st->codefreq;
        st->D0 = -(st->carrfreq - ch->f_if - ch->foffset);
        if (!st->flagremcarradd) { st->L0 -= st->remcarr / (2.0 * TRK_PI); st->flagremcarradd = 1; }
        if (st->flagsyncf && !st->flagpolarityadd) {            /* src/sdrtrk.c:186-195 */
            if (st->polarity == 1) st->L0 += 0.5;
            st->flagpolarityadd = 1;
        }
        st->L0 += st->D0 * ((double)(ch->loopms * st->currnsamp) / ch->f_sf);
        st->Isum += fabs(st->sumI[0]);
        if (flagobs == 2) {
            st->S0 = 10 * log(st->Isum / 100.0 / 100.0) + log(500.0) + 5;
            st->Isum = 0;
        }
        if (obs) {                                            /* shiftdata(...) of src/sdrtrk.c:163-168, 199-201 */
            const int h = (obs->head + SDRB200_OBSINTERPN - 1) % SDRB200_OBSINTERPN;
            obs->L[h] = st->L0; obs->D[h] = st->D0; obs->remcout[h] = st->remcout0;
            obs->codei[h] = st->obs_codei; obs->cntout[h] = st->obs_cnt;
            obs->head = h;
            if (flagobs == 2) {
                const int sh = (obs->shead + SDRB200_OBSINTERPN - 1) % SDRB200_OBSINTERPN;
                obs->S[sh] = st->S0; obs->codeisum[sh] = st->buffloc;
                obs->shead = sh;
            }
        }
        st->loopcnt++;
    }
    __syncwarp();
    if (rec) {
        for (int i = lane; i < ntap; i += 32) { rec->II[i] = st->II[i]; rec->QQ[i] = st->QQ[i]; }
        if (lane == 0) {
            rec->cnt = st->cnt; rec->buffloc = st->buffloc; rec->currnsamp = st->currnsamp;
            rec->flagsync = st->flagsync; rec->flagloopfilter = mode;
            rec->swsync = st->swsync; rec->swloop = st->swloop; rec->synci = st->synci;
            rec->biti = st->biti; rec->bit = st->bit;
            rec->sumI0 = st->sumI[0]; rec->sumQ0 = st->sumQ[0];
            rec->codefreq = st->codefreq; rec->codeErr = st->codeErr; rec->codeNco = st->codeNco;
            rec->carrfreq = st->carrfreq; rec->carrErr = st->carrErr; rec->carrNco = st->carrNco;
            rec->freqErr = st->freqErr; rec->remcode = st->remcode; rec->remcarr = st->remcarr;
            rec->L0 = st->L0; rec->D0 = st->D0; rec->S0 = st->S0; rec->remcout0 = st->remcout0;
            rec->Isum = st->Isum; rec->obs_codei = st->obs_codei; rec->obs_cnt = st->obs_cnt;
            rec->flagobs = flagobs; rec->flagsyncf = st->flagsyncf; rec->polarity = st->polarity; rec->flagtow = st->flagtow;
            rec->firstsf = st->firstsf; rec->firstsfcnt = st->firstsfcnt; rec->obs_dcnt = st->obs_dcnt;
        }
    }
    __syncwarp();
    /* clearcumsumcorr(), src/sdrtrk.c:77-86, src/sdrmain.c:310-312 */
    if (mode)
        for (int i = lane; i < ntap; i += 32) { st->oldsumI[i] = 0; st->oldsumQ[i] = 0; st->sumI[i] = 0; st->sumQ[i] = 0; }
    if (lane == 0) {
        st->flagloopfilter = mode;
        st->cnt++;
        st->buffloc += (uint64_t)st->currnsamp;
    }
    __syncwarp();
    TRK_EP(14);
#undef TRK_EP
}

#define TRK_NT_SMALL 384
#include "trk_run_kernel.cuh"

/* ------------------------------------------------------------- host side */
struct sdrb200_trk {
    int nch, dtype, maxn, smax, maxtap, qcap;
    int cl, ctrl_works;              /* cluster size per channel (0: not chosen yet); rank 0's team takes periods */
    cudaStream_t st;
    trk_chan_dev_t *h_chan, *d_chan;
    trk_state_t *h_state0, *d_state;
    sdrb200_trk_rec_t *d_rec; size_t rec_cap;       /* records */
    trk_obsstock_t *d_obs;                          /* [nch] interpolation stocks */
    sdrb200_trk_rec_t *h_rec; size_t h_rec_cap;     /* pinned staging: a device-to-host copy into pageable memory
                                                     * blocks inside the runtime and serialises the launches of
                                                     * other host threads (one engine per channel thread in the drop-in) */
    int *h_nrec;
    int *d_nrec, *d_err;
    long long *d_prof;               /* phase profile of CTA 0 (NULL: off) */
    int last_nperiods, last_with_rec;
};

static void trk_init_state(const trk_chan_dev_t *c, trk_state_t *s)
{
    memset(s, 0, sizeof(*s));
    s->buffloc = c->buffloc0;
    s->carrfreq = c->acqfreq;             /* src/sdracq.c:54-55 */
    s->codefreq = c->crate;
}

extern "C" int sdrb200_trk_create(sdrb200_trk_t **out, int nch, const sdrb200_trk_chan_t *chan)
{
    if (!out || nch < 1 || !chan) { sdrb200_set_error("trk_create: bad argument"); return SDRB200_EINVAL; }
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    int rc = sdrb200_trk_lut_ready();
    if (rc) return rc;
    sdrb200_trk_t *e = (sdrb200_trk_t *)calloc(1, sizeof(*e));
    if (!e) return SDRB200_ENOMEM;
    e->nch = nch; e->dtype = chan[0].dtype;
    e->h_chan = (trk_chan_dev_t *)calloc(nch, sizeof(trk_chan_dev_t));
    e->h_state0 = (trk_state_t *)calloc(nch, sizeof(trk_state_t));
    for (int i = 0; i < nch; i++) {
        const sdrb200_trk_chan_t *c = &chan[i];
        trk_chan_dev_t *d = &e->h_chan[i];
        if (c->dtype != e->dtype || (c->dtype != 1 && c->dtype != 2) || c->corrn < 1 ||
            c->corrn > SDRB200_MAXCORRN || c->clen < 1 || c->clen > 1024 || !c->code ||
            c->navrate < 1 || c->navrate > 32 || c->nsamp < 1 || c->loopms < 1 ||
            c->ne < 0 || c->ne > 2 * c->corrn || c->nl < 0 || c->nl > 2 * c->corrn) {
            sdrb200_set_error("trk_create: channel %d has unsupported parameters", i);
            free(e->h_chan); free(e->h_state0); free(e); return SDRB200_EINVAL;
        }
        d->f_sf = c->f_sf; d->ti = 1 / c->f_sf;                 /* sdr->ti=1/f_sf, sdrinit.c:598 */
        d->f_if = c->f_if; d->f_cf = c->f_cf; d->foffset = c->foffset;
        d->crate = c->crate; d->ctime = c->ctime; d->acqfreq = c->acqfreq; d->buffloc0 = c->buffloc;
        d->dtype = c->dtype; d->clen = c->clen; d->nsamp = c->nsamp; d->prn = c->prn;
        d->corrn = c->corrn; d->ne = c->ne; d->nl = c->nl; d->loopms = c->loopms; d->navrate = c->navrate;
        memcpy(d->corrp, c->corrp, sizeof(d->corrp));
        d->prm1 = c->prm1; d->prm2 = c->prm2;
        memcpy(d->code, c->code, c->clen);
        int maxn = c->nsamp + c->nsamp / 64 + 16;
        if (1 + 2 * c->corrn > e->maxtap) e->maxtap = 1 + 2 * c->corrn;
        if (maxn > e->maxn) e->maxn = maxn;
        if (c->corrp[c->corrn - 1] > e->smax) e->smax = c->corrp[c->corrn - 1];
        trk_init_state(d, &e->h_state0[i]);
        if (c->crate / c->f_sf > 0.5) {
            sdrb200_set_error("trk_create: channel %d has fewer than two samples per chip", i);
            free(e->h_chan); free(e->h_state0); free(e); return SDRB200_EINVAL;
        }
        { int q = trk_qcap(maxn, c->corrp[c->corrn - 1], 1.001 * c->crate / c->f_sf); if (q > e->qcap) e->qcap = q; }
    }
    if (cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc((void **)&e->d_chan, sizeof(trk_chan_dev_t) * nch) != cudaSuccess ||
        cudaMalloc((void **)&e->d_state, sizeof(trk_state_t) * nch) != cudaSuccess ||
        cudaMalloc((void **)&e->d_obs, sizeof(trk_obsstock_t) * nch) != cudaSuccess ||
        cudaMalloc((void **)&e->d_nrec, sizeof(int) * (nch + 1)) != cudaSuccess) {
        sdrb200_set_error("trk_create: device allocation failed");
        sdrb200_trk_destroy(e); return SDRB200_ENOMEM;
    }
    e->d_err = e->d_nrec + nch;
    cudaMemcpyAsync(e->d_chan, e->h_chan, sizeof(trk_chan_dev_t) * nch, cudaMemcpyHostToDevice, e->st);
    cudaMemcpyAsync(e->d_state, e->h_state0, sizeof(trk_state_t) * nch, cudaMemcpyHostToDevice, e->st);
    cudaMemsetAsync(e->d_nrec, 0, sizeof(int) * (nch + 1), e->st);
    cudaMemsetAsync(e->d_obs, 0, sizeof(trk_obsstock_t) * nch, e->st);
    SDRB200_CUDA_TRY(cudaStreamSynchronize(e->st));
    *out = e;
    return SDRB200_OK;
}

extern "C" void sdrb200_trk_destroy(sdrb200_trk_t *e)
{
    if (!e) return;
    if (e->h_rec) cudaFreeHost(e->h_rec);
    if (e->h_nrec) cudaFreeHost(e->h_nrec);
    cudaFree(e->d_chan); cudaFree(e->d_state); cudaFree(e->d_obs); cudaFree(e->d_rec); cudaFree(e->d_nrec); cudaFree(e->d_prof);
    if (e->st) cudaStreamDestroy(e->st);
    free(e->h_chan); free(e->h_state0); free(e);
}

/* phase profile of channel 0 over the last run (cycles): go, tables, convert,
 * mix+prefix, tap sums, epilogue, periods; on=1 enables, then read after sync */
extern "C" int sdrb200_trk_profile(sdrb200_trk_t *e, int on, long long *out7)
{
    if (!e) return SDRB200_EINVAL;
    if (on && !e->d_prof) {
        SDRB200_CUDA_TRY(cudaMalloc((void **)&e->d_prof, 32 * sizeof(long long)));
        SDRB200_CUDA_TRY(cudaMemset(e->d_prof, 0, 32 * sizeof(long long)));
    }
    if (out7 && e->d_prof) {
        SDRB200_CUDA_TRY(cudaStreamSynchronize(e->st));
        SDRB200_CUDA_TRY(cudaMemcpy(out7, e->d_prof, 32 * sizeof(long long), cudaMemcpyDeviceToHost));
        SDRB200_CUDA_TRY(cudaMemset(e->d_prof, 0, 32 * sizeof(long long)));
    }
    if (!on && e->d_prof) { cudaFree(e->d_prof); e->d_prof = NULL; }
    return SDRB200_OK;
}

#ifdef TRK_DEBUG
extern "C" int sdrb200_trk_dbg(long long *out) { return cudaMemcpyFromSymbol(out, g_dbg, sizeof(long long) * 4096) == cudaSuccess ? 0 : -4; }
#endif

extern "C" int sdrb200_trk_reset(sdrb200_trk_t *e)
{
    if (!e) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(e->d_state, e->h_state0, sizeof(trk_state_t) * e->nch, cudaMemcpyHostToDevice, e->st));
    SDRB200_CUDA_TRY(cudaMemsetAsync(e->d_obs, 0, sizeof(trk_obsstock_t) * e->nch, e->st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(e->st));
    return SDRB200_OK;
}

extern "C" void *sdrb200_trk_cuda_stream(sdrb200_trk_t *e) { return e ? (void *)e->st : NULL; }

static int trk_launch(sdrb200_trk_t *e, const sdrb200_stream_t *s, int nperiods, int with_rec)
{
    if (!e || !s || nperiods < 1) { sdrb200_set_error("trk_run: bad argument"); return SDRB200_EINVAL; }
    if (s->dtype != e->dtype) { sdrb200_set_error("trk_run: stream dtype mismatch"); return SDRB200_EINVAL; }
    if (with_rec) {
        size_t need = (size_t)e->nch * nperiods;
        if (need > e->rec_cap) {
            SDRB200_CUDA_TRY(cudaStreamSynchronize(e->st));
            cudaFree(e->d_rec); e->d_rec = NULL; e->rec_cap = 0;
            if (cudaMalloc((void **)&e->d_rec, need * sizeof(sdrb200_trk_rec_t)) != cudaSuccess) {
                cudaGetLastError(); sdrb200_set_error("trk_run: record buffer allocation failed"); return SDRB200_ENOMEM;
            }
            e->rec_cap = need;
        }
    }
    sdrb200_trk_rec_t *recs = with_rec ? e->d_rec : NULL;
    /* CTA size: the small instantiation when one round of its worker team covers the period, and - the
     * team takes a longer period in several rounds - when the channels need more than two waves of the
     * large one anyway: a channel is a latency chain, so two small CTAs per SM carry more channels at once
     * (2048 channels x 16 368 samples: 9.6 k -> 11.8 k channel.s/s; at 148 channels the large CTA wins,
     * 8.2 k against 5.9 k).  SDRB200_TRK_SMALL=0/1 overrides.                                         */
    int small = ((size_t)e->maxn * e->dtype + 3 <= (size_t)(TRK_NT_SMALL - TRK_T0) * TRK_SPAN) ? 1 : 0;
    {
        int nsm = 148, dev = 0;
        if (!small && cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && e->nch >= 2 * nsm) small = 1;
        cudaGetLastError();
        const char *env = getenv("SDRB200_TRK_SMALL"); if (env) small = atoi(env) ? 1 : 0;
    }
    const int nt = small ? TRK_NT_SMALL : TRK_THREADS;
    size_t smem = ((sizeof(trk_cl_smem_t) + 15) & ~(size_t)15) +
                  sizeof(unsigned int) * TRKD_WORK_WORDS(e->maxn, 2 * e->smax, (nt - TRK_T0) / 32);
    if (smem > TRK_SMEM_LIMIT) { sdrb200_set_error("trk_run: code period too long for one CTA"); return SDRB200_EINVAL; }
    const void *fn = (e->dtype == 1) ? (small ? (const void *)k_trk_cl<1, TRK_NT_SMALL> : (const void *)k_trk_cl<1, TRK_THREADS>)
                                     : (small ? (const void *)k_trk_cl<2, TRK_NT_SMALL> : (const void *)k_trk_cl<2, TRK_THREADS>);
    SDRB200_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[1];
    cfg.blockDim = dim3(nt); cfg.dynamicSmemBytes = smem; cfg.stream = e->st;
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (!e->cl) {
        /* cluster size: the largest of 8, 4, 2 whose clusters are all co-resident (so that every
         * channel advances at once); beyond that one CTA per channel, scheduled in waves */
        int cl = 1;
        const char *env = getenv("SDRB200_TRK_CLUSTER");
        if (env && atoi(env) >= 1 && atoi(env) <= 8) cl = atoi(env);
        else {
            for (int c = 8; c >= 2; c >>= 1) {
                int nact = 0;
                cfg.gridDim = dim3(e->nch * c); at[0].val.clusterDim.x = c;
                if (cudaOccupancyMaxActiveClusters(&nact, fn, &cfg) == cudaSuccess && nact >= e->nch) { cl = c; break; }
                cudaGetLastError();
            }
        }
        e->cl = cl;
        env = getenv("SDRB200_TRK_CTRL_WORKS");
        e->ctrl_works = env ? atoi(env) : 0;
    }
    cfg.gridDim = dim3(e->nch * e->cl); at[0].val.clusterDim.x = e->cl;
    if (getenv("SDRB200_VERBOSE")) {
        int nact = 0; cudaOccupancyMaxActiveClusters(&nact, fn, &cfg); cudaGetLastError();
        fprintf(stderr, "sdrb200: k_trk_cl: %d channels, %d threads, cluster %d, %zu B shared memory, %d clusters co-resident\n", e->nch, nt, e->cl, smem, nact);
    }
    {
        const trk_chan_dev_t *a_ch = e->d_chan; trk_state_t *a_st = e->d_state; const int8_t *a_if = s->d_buf;
        uint64_t a_cap = s->capacity, a_av = s->count; int a_np = nperiods, a_mx = e->maxn, a_qc = e->qcap;
        int *a_nrec = e->d_nrec, *a_err = e->d_err, a_cw = e->ctrl_works; long long *a_prof = e->d_prof;
        trk_obsstock_t *a_obs = e->d_obs;
        void *args[] = {&a_ch, &a_st, &a_if, &a_cap, &a_av, &a_np, &a_mx, &a_qc, &recs, &a_nrec, &a_err, &a_cw, &a_prof, &a_obs};
        SDRB200_CUDA_TRY(cudaLaunchKernelExC(&cfg, fn, args));
    }
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    e->last_nperiods = nperiods; e->last_with_rec = with_rec;
    return SDRB200_OK;
}

/* cluster size per channel (1, 2, 4 or 8; 0 = choose at the next launch from the co-residency of
 * this engine's channels alone) and whether rank 0's worker team takes periods too.  For callers
 * that run several engines at once and know the total number of channels on the GPU.        */
extern "C" int sdrb200_trk_set_cluster(sdrb200_trk_t *e, int cluster_size, int ctrl_works)
{
    if (!e || cluster_size < 0 || cluster_size > 8 || (cluster_size & (cluster_size - 1))) {
        sdrb200_set_error("trk_set_cluster: cluster size must be 0, 1, 2, 4 or 8"); return SDRB200_EINVAL;
    }
    e->cl = cluster_size; e->ctrl_works = ctrl_works ? 1 : 0;
    return SDRB200_OK;
}

extern "C" int sdrb200_trk_enqueue(sdrb200_trk_t *e, const sdrb200_stream_t *s, int nperiods)
{
    return trk_launch(e, s, nperiods, 0);
}

/* the stocks of one channel, newest first - the arrays sdrsync.c interpolates (src/sdrsync.c:96-121) */
extern "C" int sdrb200_trk_fetch_obs(sdrb200_trk_t *e, int ch, sdrb200_trk_obs_t *out)
{
    if (!e || !out || ch < 0 || ch >= e->nch) { sdrb200_set_error("trk_fetch_obs: bad argument"); return SDRB200_EINVAL; }
    trk_obsstock_t h;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(&h, e->d_obs + ch, sizeof(h), cudaMemcpyDeviceToHost, e->st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(e->st));
    for (int j = 0; j < SDRB200_OBSINTERPN; j++) {
        const int k = (h.head + j) % SDRB200_OBSINTERPN, ks = (h.shead + j) % SDRB200_OBSINTERPN;
        out->L[j] = h.L[k]; out->D[j] = h.D[k]; out->remcout[j] = h.remcout[k];
        out->codei[j] = h.codei[k]; out->cntout[j] = h.cntout[k];
        out->S[j] = h.S[ks]; out->codeisum[j] = h.codeisum[ks];
    }
    return SDRB200_OK;
}

extern "C" int sdrb200_trk_sync(sdrb200_trk_t *e)
{
    if (!e) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaStreamSynchronize(e->st));
    return SDRB200_OK;
}

extern "C" int sdrb200_trk_run(sdrb200_trk_t *e, const sdrb200_stream_t *s, int nperiods,
                               sdrb200_trk_rec_t *rec, int *nrec)
{
    int rc = trk_launch(e, s, nperiods, rec != NULL);
    if (rc) return rc;
    if (!e->h_nrec) SDRB200_CUDA_TRY(cudaMallocHost((void **)&e->h_nrec, sizeof(int) * (size_t)(e->nch + 1)));
    int *h = e->h_nrec;
    const size_t nrecs = (size_t)e->nch * nperiods;
    if (rec && nrecs > e->h_rec_cap) {
        if (e->h_rec) cudaFreeHost(e->h_rec);
        e->h_rec = NULL; e->h_rec_cap = 0;
        SDRB200_CUDA_TRY(cudaMallocHost((void **)&e->h_rec, sizeof(sdrb200_trk_rec_t) * nrecs));
        e->h_rec_cap = nrecs;
    }
    SDRB200_CUDA_TRY(cudaMemcpyAsync(h, e->d_nrec, sizeof(int) * (e->nch + 1), cudaMemcpyDeviceToHost, e->st));
    if (rec)
        SDRB200_CUDA_TRY(cudaMemcpyAsync(e->h_rec, e->d_rec, sizeof(sdrb200_trk_rec_t) * nrecs, cudaMemcpyDeviceToHost, e->st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(e->st));
    if (rec) memcpy(rec, e->h_rec, sizeof(sdrb200_trk_rec_t) * nrecs);
    if (h[e->nch]) {
        const int code = h[e->nch];
        SDRB200_CUDA_TRY(cudaMemsetAsync(e->d_err, 0, sizeof(int), e->st));
        sdrb200_set_error(code == 2 ? "trk_run: a channel lags by more than the stream's capacity (samples overwritten)"
                        : code == 3 ? "trk_run: code period outside the work area (currnsamp <= 0 or > nsamp + nsamp/64 + 16)"
                                    : "trk_run: NCO segment table overflow");
        if (nrec) memcpy(nrec, h, sizeof(int) * e->nch);
        return SDRB200_ERANGE;
    }
    if (nrec) memcpy(nrec, h, sizeof(int) * e->nch);
    return SDRB200_OK;
}
