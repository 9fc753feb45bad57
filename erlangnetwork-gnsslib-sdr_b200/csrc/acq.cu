/* acq.cu - batched FFT acquisition over the PRN x Doppler x code-phase grid.
 *
 * Stands in for sdracquisition() (src/sdracq.c:14-62), pcorrelator()
 * (src/sdrcmn.c:738-773), cpxconv() (src/sdrcmn.c:228-251) and
 * checkacquisition() (src/sdracq.c:71-95).
 *
 * Mathematics.  The reference correlates a window of 2n live samples with an
 * n-sample replica zero-padded to m = 2n and keeps lags 0..n-1 of the m-point
 * circular correlation (src/sdrinit.c:650-652, src/sdrcmn.c:244-246).  For
 * those lags no index wraps, so the result is the LINEAR correlation and any
 * transform length M >= 2n gives the same numbers; here M = 2^k >= 2n.
 * M is split as M = R * L, L <= 16384 handled by one CTA (fft_device.cuh):
 *   forward :  X[R j + q] = FFT_L{ W_M^(b q) * sum_a x[L a + b] W_R^(a q) }[j]
 *   inverse :  y[t] = sum_q W_M^(-t q) * IFFT_L{ Y[R j + q] }[t mod L]
 * so spectra are stored de-interleaved as [q][j] and never re-ordered.
 * Scaling follows the reference exactly: input scale (float)(CSCALE/m)
 * (src/sdrcmn.c:764), power (re^2+im^2)/M^2 (src/sdrcmn.c:245).
 *
 * Pipeline per integration step (all on one stream, no host round trip):
 *   k_fwd   int8 IF -> table carrier wipe-off -> forward FFT, one per Doppler bin
 *   k_row   spectrum x conj(code spectrum) -> inverse sub-FFTs, per (PRN, bin)
 *   k_comb  recombine, |.|^2, accumulate P, per-row max / masked mean / 2nd max
 *   k_final per-PRN argmax over bins, C/N0, peak ratio, early-exit freeze
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "sdrb200_internal.h"
#include "nco_exact.h"
#include "fft_device.cuh"
#include "trk_cluster.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

/* the carrier table (src/sdrcmn.c:641-646), float copies for the wipe-off */
__constant__ float c_cosf[32] = {32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6,0,6,12,18,23,27,30,31};
__constant__ float c_sinf[32] = {0,6,12,18,23,27,30,31,32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6};

struct acq_rowstat_t {
    float maxp;      /* row maximum                                         */
    int   codei;     /* its first index                                     */
    float max2;      /* maxvd() over the row outside the exclusion window   */
    int   cnt;       /* samples outside the window                          */
    double sum;      /* their sum                                           */
};

#define ACQ_MAXEV 256
#define ACQ_MAXLANES 8
struct sdrb200_acq {
    sdrb200_acq_cfg_t cfg;
    int n_eff, m_ref, M, L, R;
    int nfx;                         /* wipe-off grids held: nfreq, or nprn * nfreq when every replica has its own (cfg.grid_per_prn) */
    int chunk;                       /* PRNs per scratch chunk                */
    int fchunk;                      /* Doppler bins per scratch chunk (nfreq unless one PRN is too large) */
    int split_fwd;                   /* forward transform in two passes (k_fwd_split* + k_fft_rows)        */
    int comb_fft;                    /* scratch path: radix-R recombination + k_peak_hbm                   */
    double *h_freq;
    cudaStream_t st;
    int nlanes;                      /* two-pass path: chunks in flight at once (own streams and scratch slices) */
    cudaStream_t lane_st[ACQ_MAXLANES];
    cudaEvent_t lane_ev[ACQ_MAXLANES], ev_fork;
    float2 *d_tw;                    /* [M] forward twiddles                  */
    uint8_t *d_cidx;                 /* [nfreq][m_ref] carrier table indices  */
    float2 *d_Xs;                    /* [nfreq][R][L]                         */
    float2 *d_Cs;                    /* [nprn][R][L]                          */
    float2 *d_S;                     /* [chunk][nfreq][R][L] scratch          */
    float  *d_P;                     /* [nprn][nfreq][n_eff]                  */
    acq_rowstat_t *d_row;            /* [nprn][nfreq] (legacy path only)      */
    unsigned long long *d_key;       /* [nprn] argmax keys (fused path)       */
    int fused, keep_power;
    int push;                        /* k_rowd (bulk-copy exchange) instead of k_roww (remote loads)        */
    int warp32;                      /* M = 32 x 1024: one warp per sub-transform (k_roww)                  */
    int rowp_ncl;                    /* k_roww / k_rowd: clusters launched (co-resident on this GPU)         */
    sdrb200_acq_result_t *d_res, *h_res;
    sdrb200_acq_result_t *d_hist;    /* [intg][nprn] tuples after every integration step (Doppler-slab searches) */
    int *d_frozen;
    double *d_freq;
    int8_t *d_hostwin; size_t hostwin_cap;   /* staging for run_host          */
    int8_t *h_pin; size_t pin_cap;
    /* optional per-kernel timing (bench.py): events around every launch      */
    int timing, nev;
    cudaEvent_t ev0[ACQ_MAXEV], ev1[ACQ_MAXEV];
    int evkind[ACQ_MAXEV];
    double tot_ms[4]; long long tot_cnt[4];
};
#define ACQ_T0(a) do { if ((a)->timing && (a)->nev < ACQ_MAXEV) cudaEventRecord((a)->ev0[(a)->nev], (a)->st); } while (0)
#define ACQ_T1(a, kind) do { if ((a)->timing && (a)->nev < ACQ_MAXEV) { cudaEventRecord((a)->ev1[(a)->nev], (a)->st); (a)->evkind[(a)->nev++] = (kind); } } while (0)

/* ---------------------------------------------------------------- kernels */

/* forward transform of one (item, q).  SRC 0: IF window with carrier wipe-off
 * (mixcarr + cpxcpx, src/sdrcmn.c:650-662, 185-195); SRC 1: real int8 replica
 * (cpxcpx(rcode,NULL,1.0,...), src/sdrinit.c:652).                           */
template <int L, int SRC>
__global__ void __launch_bounds__(L / 16)
k_fwd(const int8_t *__restrict__ src, size_t item_stride, int dtype, int nlive,
      const uint8_t *__restrict__ cidx, int cidx_stride, float scale,
      const float2 *__restrict__ tw, int R, float2 *__restrict__ out)
{
    extern __shared__ float2 fsm[];
    constexpr int T = L / 16;
    const int q = blockIdx.x, item = blockIdx.y, tid = threadIdx.x;
    const int M = L * R;
    const int8_t *d = src + (SRC == 1 ? (size_t)item * item_stride : 0);
    const uint8_t *ci = (SRC == 0) ? cidx + (size_t)item * cidx_stride : nullptr;
    /* the 32-entry carrier table lives one entry per lane: a lookup is one shuffle
     * (a __constant__ table would serialise on the 32 different indices of a warp) */
    const float lut_c = c_cosf[tid & 31], lut_s = c_sinf[tid & 31];
    float2 v[16];
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = make_float2(0.f, 0.f);
    /* residue-class sums: outer loop over the R blocks (one uniform twiddle each),
     * inner loop unrolled so that the 16 (or 32) byte loads of a block are in flight
     * together                                                                       */
    for (int a = 0; a < R; a++) {
        const float2 wa = (a != 0 && q != 0) ? __ldg(&tw[(size_t)((a * q) % R) * L]) : make_float2(1.f, 0.f);
        /* this lane's entry of the block's table: carrier LUT x input scale x W_R^(a q) */
        const float2 g = cmul(make_float2(lut_c * scale, lut_s * scale), wa);
        int idx[16]; int s0[16], s1[16];
#pragma unroll
        for (int e = 0; e < 16; e++) {
            /* branch-free: clamp the address, zero the sample past the window, so that
             * all loads of the block are issued back to back                        */
            const int t = L * a + tid + e * T;
            const bool live = t < nlive;
            const int tc = live ? t : nlive - 1;
            idx[e] = (SRC == 0) ? (int)ci[tc] : 0;
            if (SRC == 0 && dtype != 1) {
                const int v0 = d[2 * tc], v1 = d[2 * tc + 1];
                s0[e] = live ? v0 : 0; s1[e] = live ? v1 : 0;
            } else {
                const int v0 = d[tc];
                s0[e] = live ? v0 : 0; s1[e] = 0;
            }
        }
#pragma unroll
        for (int e = 0; e < 16; e++) {
            if (SRC == 0) {
                const float gx = __shfl_sync(0xffffffffu, g.x, idx[e]), gy = __shfl_sync(0xffffffffu, g.y, idx[e]);
                const float f0 = (float)s0[e];
                v[e].x = fmaf(f0, gx, v[e].x); v[e].y = fmaf(f0, gy, v[e].y);
                if (dtype != 1) {                      /* (s0 + i s1) * g */
                    const float f1 = (float)s1[e];
                    v[e].x = fmaf(-f1, gy, v[e].x); v[e].y = fmaf(f1, gx, v[e].y);
                }
            } else {
                const float f0 = (float)s0[e];
                v[e].x = fmaf(f0, wa.x, v[e].x); v[e].y = fmaf(f0, wa.y, v[e].y);
            }
        }
    }
    if (q != 0) {
#pragma unroll
        for (int e = 0; e < 16; e++) v[e] = cmul(v[e], __ldg(&tw[(size_t)(tid + e * T) * q]));
    }
    cta_fft<L, -1>(v, fsm, tw, R);
    float2 *o = out + ((size_t)item * R + q) * L;
#pragma unroll
    for (int e = 0; e < 16; e++) o[tid + e * T] = v[e];
    (void)M;
}

/* forward transform of the IF window for one (Doppler bin, q): mixcarr + cpxcpx
 * (src/sdrcmn.c:650-662, 185-195) fused into the residue-class sums of the
 * four-step transform.  The window is walked in R blocks of L samples.  In this
 * phase a thread owns 16 CONSECUTIVE samples of the block: its samples and their
 * carrier-table indices arrive as 16-byte loads (fetched one block ahead) and are
 * picked apart in registers.  The 32-entry carrier table x input scale x W_R^(a q)
 * is held one entry per lane, so a lookup is a shuffle and a sample costs two
 * FMAs.  One pass through shared memory then hands the sums over in the order
 * the CTA transform wants (element e of thread tid <-> b = tid + e*T).          */
__device__ __forceinline__ int acq_sbyte(unsigned int w, int j)      /* sign-extended byte j */
{
    return (int)(w << (24 - 8 * j)) >> 24;
}

template <int L, int DT>
__global__ void __launch_bounds__(L / 16, (L <= 4096) ? 4 : (L <= 8192 ? 2 : 1))
k_fwd_if(const int8_t *__restrict__ win, int nlive, const uint8_t *__restrict__ cidx,
         int cidx_stride, float scale, const float2 *__restrict__ tw, int R, float2 *__restrict__ out)
{
    extern __shared__ float2 fsm[];
    constexpr int T = L / 16;
    constexpr int NW = 4 * DT;                            /* sample words per thread and block */
    const int q = blockIdx.x, f = blockIdx.y, tid = threadIdx.x;
    const size_t g = (size_t)win;
    const int off = (int)(g & 15), wo = off >> 2, bo8 = 8 * (off & 3);
    const uint4 *wal = (const uint4 *)(g - off);
    const uint4 *cal = (const uint4 *)(cidx + (size_t)f * cidx_stride);
    const int live_bytes = nlive * DT + off;              /* relative to wal */
    const float lut_c = c_cosf[tid & 31] * scale, lut_s = c_sinf[tid & 31] * scale;
    const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
    uint4 pd[DT + 1], pi;
    auto fetch = [&](int a) {
        /* chunks (a*T + tid)*DT + {0..DT}: the extra one covers the misalignment;
         * anything past the live window reads as zero                              */
        const int c0 = (a * T + tid) * DT;
#pragma unroll
        for (int i = 0; i <= DT; i++) pd[i] = ((c0 + i) * 16 < live_bytes) ? __ldg(&wal[c0 + i]) : zero4;
        pi = ((a * T + tid) * 16 < nlive) ? __ldg(&cal[a * T + tid]) : zero4;
    };
    float2 u[16];
#pragma unroll
    for (int k = 0; k < 16; k++) u[k] = make_float2(0.f, 0.f);
    const int nblk = (nlive + L - 1) / L;                  /* blocks that hold samples */
    fetch(0);
    for (int a = 0; a < nblk; a++) {
        /* realign: drop `off` bytes from the front of the DT+1 chunks */
        unsigned int w[4 * (DT + 1) + 1];
#pragma unroll
        for (int i = 0; i <= DT; i++) { w[4 * i] = pd[i].x; w[4 * i + 1] = pd[i].y; w[4 * i + 2] = pd[i].z; w[4 * i + 3] = pd[i].w; }
        w[4 * (DT + 1)] = 0u;
        unsigned int t2[NW + 2], t1[NW + 1], sw[NW];
#pragma unroll
        for (int i = 0; i < NW + 2; i++) t2[i] = (wo & 2) ? w[i + 2] : w[i];
#pragma unroll
        for (int i = 0; i < NW + 1; i++) t1[i] = (wo & 1) ? t2[i + 1] : t2[i];
#pragma unroll
        for (int i = 0; i < NW; i++) sw[i] = __funnelshift_r(t1[i], t1[i + 1], bo8);
        const unsigned int iw[4] = {pi.x, pi.y, pi.z, pi.w};
        const int tb = L * a + 16 * tid;                   /* first sample of this thread */
        if (a + 1 < nblk) fetch(a + 1);
        const float2 wa = (a != 0 && q != 0) ? __ldg(&tw[(size_t)((a * q) % R) * L]) : make_float2(1.f, 0.f);
        const float2 g2 = cmul(make_float2(lut_c, lut_s), wa);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const int idx = (int)((iw[k >> 2] >> (8 * (k & 3))) & 0xffu);
            const float gx = __shfl_sync(0xffffffffu, g2.x, idx), gy = __shfl_sync(0xffffffffu, g2.y, idx);
            const bool live = tb + k < nlive;
            if (DT == 1) {
                const float f0 = live ? (float)acq_sbyte(sw[k >> 2], k & 3) : 0.f;
                u[k].x = fmaf(f0, gx, u[k].x); u[k].y = fmaf(f0, gy, u[k].y);
            } else {                                       /* (s0 + i s1) * g */
                const unsigned int ww = sw[k >> 1];
                const float f0 = live ? (float)acq_sbyte(ww, 2 * (k & 1)) : 0.f;
                const float f1 = live ? (float)acq_sbyte(ww, 2 * (k & 1) + 1) : 0.f;
                u[k].x = fmaf(f0, gx, u[k].x); u[k].y = fmaf(f0, gy, u[k].y);
                u[k].x = fmaf(-f1, gy, u[k].x); u[k].y = fmaf(f1, gx, u[k].y);
            }
        }
    }
    /* hand over: u[k] <-> b = 16*tid + k  ->  v[e] <-> b = tid + e*T */
#pragma unroll
    for (int k = 0; k < 16; k++) fsm[17 * tid + k] = u[k];            /* FFT_PAD(16*tid + k) */
    __syncthreads();
    float2 v[16];
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = fsm[FFT_PAD(tid + e * T)];
    __syncthreads();
    if (q != 0) {
#pragma unroll
        for (int e = 0; e < 16; e++) v[e] = cmul(v[e], __ldg(&tw[(size_t)(tid + e * T) * q]));
    }
    cta_fft<L, -1>(v, fsm, tw, R);
    float2 *o = out + ((size_t)f * R + q) * L;
#pragma unroll
    for (int e = 0; e < 16; e++) o[tid + e * T] = v[e];
}

/* per-row statistics over a power row held in global memory; all threads of
 * the CTA call it.  Implements the three scans of checkacquisition() for this
 * row: maxvd (first index wins, src/sdrcmn.c:461-476), meanvd (487-497) and
 * the second maxvd with the exclusion window and its data[0] seed quirk.      */
__device__ void acq_row_stats(const float *Prow, int n, int nsampchip,
                              float tmax, int tidx, acq_rowstat_t *out)
{
    __shared__ float s_f[32];
    __shared__ int   s_i[32];
    __shared__ double s_d[32];
    __shared__ int   s_codei;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    /* block argmax: larger value, then smaller index */
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, tmax, o);
        int   oi = __shfl_xor_sync(0xffffffffu, tidx, o);
        if (ov > tmax || (ov == tmax && oi < tidx)) { tmax = ov; tidx = oi; }
    }
    if (lane == 0) { s_f[warp] = tmax; s_i[warp] = tidx; }
    __syncthreads();
    if (warp == 0) {
        tmax = (lane < nwarp) ? s_f[lane] : -1.f;
        tidx = (lane < nwarp) ? s_i[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, tmax, o);
            int   oi = __shfl_xor_sync(0xffffffffu, tidx, o);
            if (ov > tmax || (ov == tmax && oi < tidx)) { tmax = ov; tidx = oi; }
        }
        if (lane == 0) { s_f[0] = tmax; s_codei = tidx; }
    }
    __syncthreads();
    const float rowmax = s_f[0];
    const int codei = s_codei;
    /* exclusion window, src/sdracq.c:81-82 */
    int exs = codei - 2 * nsampchip; if (exs < 0) exs += n;
    int exe = codei + 2 * nsampchip; if (exe >= n) exe -= n;
    double sum = 0.0; int cnt = 0; float m2 = -1.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const bool inc = (exs <= exe) ? (i < exs || i > exe) : (i < exs && i > exe);
        if (inc) {
            const float p = Prow[i];
            sum += (double)p; cnt++;
            if (i >= 1 && p > m2) m2 = p;
        }
    }
    if (threadIdx.x == 0) { const float p0 = Prow[0]; if (p0 > m2) m2 = p0; }   /* max=data[0] seed */
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    }
    if (lane == 0) { s_d[warp] = sum; s_i[warp] = cnt; s_f[warp] = m2; }
    __syncthreads();
    if (warp == 0) {
        sum = (lane < nwarp) ? s_d[lane] : 0.0;
        cnt = (lane < nwarp) ? s_i[lane] : 0;
        m2  = (lane < nwarp) ? s_f[lane] : -1.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sum += __shfl_xor_sync(0xffffffffu, sum, o);
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
        }
        if (lane == 0) { out->maxp = rowmax; out->codei = codei; out->max2 = m2; out->cnt = cnt; out->sum = sum; }
    }
}

/* inverse sub-transform of one (q, bin, PRN): cpxconv()'s spectrum product and
 * inverse FFT (src/sdrcmn.c:236-242).  For R == 1 the power, accumulation and
 * row statistics are fused here and nothing but the P row is written.          */
template <int L>
__global__ void __launch_bounds__(L / 16)
k_row(const float2 *__restrict__ Xs, const float2 *__restrict__ Cs, const float2 *__restrict__ tw,
      int R, int nfreq, int prn0, int f0, int nfc, float2 *__restrict__ S, float *__restrict__ P, int n_eff,
      int nsampchip, int first_step, float invM2, acq_rowstat_t *__restrict__ rows,
      const int *__restrict__ frozen)
{
    extern __shared__ float2 fsm[];
    constexpr int T = L / 16;
    /* grid (PRNs of the chunk, R, bins): the CTAs that read the same slice of a spectrum - one per PRN - are
     * scheduled next to each other, so that slice comes from HBM once per chunk instead of once per PRN */
    const int pc = blockIdx.x, q = blockIdx.y, fl = blockIdx.z, f = f0 + fl, p = prn0 + pc, tid = threadIdx.x;
    if (frozen[p]) return;
    const float2 *x = Xs + ((size_t)f * R + q) * L;
    const float2 *c = Cs + ((size_t)p * R + q) * L;
    float2 v[16];
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = cmulc(__ldg(&x[tid + e * T]), __ldg(&c[tid + e * T]));
    cta_fft<L, +1>(v, fsm, tw, R);
    if (R > 1) {
        float2 *o = S + (((size_t)pc * nfc + fl) * R + q) * L;
#pragma unroll
        for (int e = 0; e < 16; e++) o[tid + e * T] = v[e];
        return;
    }
    /* R == 1: y[t] = v, t = tid + e*T */
    float *Prow = P + ((size_t)p * nfreq + f) * n_eff;
    float tmax = -1.f; int tidx = 0x7fffffff;
#pragma unroll
    for (int e = 0; e < 16; e++) {
        const int t = tid + e * T;
        if (t < n_eff) {
            float pw = (v[e].x * v[e].x + v[e].y * v[e].y) * invM2;
            if (!first_step) pw += Prow[t];
            Prow[t] = pw;
            if (pw > tmax) { tmax = pw; tidx = t; }
        }
    }
    __syncthreads();                      /* Prow visible to the whole CTA */
    acq_row_stats(Prow, n_eff, nsampchip, tmax, tidx, &rows[(size_t)p * nfreq + f]);
}

/* recombination of the R sub-transforms of one (bin, PRN) row + power +
 * accumulation (cpxconv flagsum, src/sdrcmn.c:244-246) + row statistics        */
__global__ void __launch_bounds__(1024)
k_comb(const float2 *__restrict__ S, const float2 *__restrict__ tw, int R, int L, int nfreq, int prn0,
       int f0, int nfc, float *__restrict__ P, int n_eff, int nsampchip, int first_step, float invM2,
       acq_rowstat_t *__restrict__ rows, const int *__restrict__ frozen)
{
    const int fl = blockIdx.x, f = f0 + fl, pc = blockIdx.y, p = prn0 + pc;
    if (frozen[p]) return;
    const int M = R * L;
    const float2 *s = S + ((size_t)pc * nfc + fl) * (size_t)M;
    float *Prow = P + ((size_t)p * nfreq + f) * n_eff;
    float tmax = -1.f; int tidx = 0x7fffffff;
    for (int t = threadIdx.x; t < n_eff; t += blockDim.x) {
        const int tl = t & (L - 1);
        float2 y = __ldg(&s[tl]);
        for (int q = 1; q < R; q++) {
            float2 w = __ldg(&tw[(size_t)(((long long)t * q) % M)]);
            y = cadd(y, cmulc(__ldg(&s[(size_t)q * L + tl]), w));      /* * conj(W_M^(t q)) */
        }
        float pw = (y.x * y.x + y.y * y.y) * invM2;
        if (!first_step) pw += Prow[t];
        Prow[t] = pw;
        if (pw > tmax) { tmax = pw; tidx = t; }
    }
    __syncthreads();
    acq_row_stats(Prow, n_eff, nsampchip, tmax, tidx, &rows[(size_t)p * nfreq + f]);
}

/* 5-point DFT, z[a] = sum_g z[g] exp(SIGN 2 pi i a g / 5): the two conjugate pairs share their cosine and sine sums */
template <int SIGN> __device__ __forceinline__ void fft5(float2 &x0, float2 &x1, float2 &x2, float2 &x3, float2 &x4)
{
    constexpr float C1 = 0.30901699437494742f, C2 = -0.80901699437494742f;      /* cos(2 pi / 5), cos(4 pi / 5) */
    constexpr float S1 = SIGN * 0.95105651629515357f, S2 = SIGN * 0.58778525229247313f;
    const float2 a1 = cadd(x1, x4), a2 = cadd(x2, x3), b1 = csub(x1, x4), b2 = csub(x2, x3);
    const float2 e1 = make_float2(fmaf(C2, a2.x, fmaf(C1, a1.x, x0.x)), fmaf(C2, a2.y, fmaf(C1, a1.y, x0.y)));
    const float2 e2 = make_float2(fmaf(C1, a2.x, fmaf(C2, a1.x, x0.x)), fmaf(C1, a2.y, fmaf(C2, a1.y, x0.y)));
    /* i * (S1 b1 + S2 b2) and i * (S2 b1 - S1 b2) */
    const float2 o1 = make_float2(-fmaf(S2, b2.y, S1 * b1.y), fmaf(S2, b2.x, S1 * b1.x));
    const float2 o2 = make_float2(-fmaf(-S1, b2.y, S2 * b1.y), fmaf(-S1, b2.x, S2 * b1.x));
    x0 = cadd(x0, cadd(a1, a2));
    x1 = cadd(e1, o1); x4 = csub(e1, o1);
    x2 = cadd(e2, o2); x3 = csub(e2, o2);
}

/* 3-point DFT, z[a] = sum_g z[g] exp(SIGN 2 pi i a g / 3) */
template <int SIGN> __device__ __forceinline__ void fft3(float2 &x0, float2 &x1, float2 &x2)
{
    constexpr float S = SIGN * 0.86602540378443865f;                            /* sin(2 pi / 3) */
    const float2 a = cadd(x1, x2), b = csub(x1, x2);
    const float2 e = make_float2(fmaf(-0.5f, a.x, x0.x), fmaf(-0.5f, a.y, x0.y));
    const float2 o = make_float2(-S * b.y, S * b.x);                            /* i S b */
    x0 = cadd(x0, a); x1 = cadd(e, o); x2 = csub(e, o);
}

/* G-point DFT in registers, natural order: z[a] = sum_g z[g] exp(SIGN 2 pi i a g / G) */
template <int G, int SIGN> __device__ __forceinline__ void fft_small(float2 (&z)[G])
{
    if constexpr (G == 3) {
        fft3<SIGN>(z[0], z[1], z[2]);
    } else if constexpr (G == 6) {
        /* g = 2 g1 + g2, a = a1 + 3 a2: two 3-point transforms, W_6^(a1), one radix-2 stage */
        float2 e0 = z[0], e1 = z[2], e2 = z[4], o0 = z[1], o1 = z[3], o2 = z[5];
        fft3<SIGN>(e0, e1, e2);
        fft3<SIGN>(o0, o1, o2);
        o1 = cmul(o1, make_float2(0.5f, SIGN * 0.86602540378443865f));
        o2 = cmul(o2, make_float2(-0.5f, SIGN * 0.86602540378443865f));
        z[0] = cadd(e0, o0); z[1] = cadd(e1, o1); z[2] = cadd(e2, o2);
        z[3] = csub(e0, o0); z[4] = csub(e1, o1); z[5] = csub(e2, o2);
    } else if constexpr (G == 12) {
        /* g = 4 g1 + g2, a = a1 + 3 a2: four 3-point transforms over g1, W_12^(g2 a1), three 4-point transforms over g2 */
        constexpr float H = 0.86602540378443865f;
        float2 f[4][3];
#pragma unroll
        for (int g2 = 0; g2 < 4; g2++) { f[g2][0] = z[g2]; f[g2][1] = z[4 + g2]; f[g2][2] = z[8 + g2]; fft3<SIGN>(f[g2][0], f[g2][1], f[g2][2]); }
        f[1][1] = cmul(f[1][1], make_float2(H, SIGN * 0.5f));                   /* W_12^1 */
        f[1][2] = cmul(f[1][2], make_float2(0.5f, SIGN * H));                   /* W_12^2 */
        f[2][1] = cmul(f[2][1], make_float2(0.5f, SIGN * H));                   /* W_12^2 */
        f[2][2] = cmul(f[2][2], make_float2(-0.5f, SIGN * H));                  /* W_12^4 */
        f[3][1] = make_float2(-SIGN * f[3][1].y, SIGN * f[3][1].x);             /* W_12^3 = SIGN i */
        f[3][2] = make_float2(-f[3][2].x, -f[3][2].y);                          /* W_12^6 = -1 */
#pragma unroll
        for (int a1 = 0; a1 < 3; a1++) {
            fft4<SIGN>(f[0][a1], f[1][a1], f[2][a1], f[3][a1]);
            z[a1] = f[0][a1]; z[a1 + 3] = f[1][a1]; z[a1 + 6] = f[2][a1]; z[a1 + 9] = f[3][a1];
        }
    } else if constexpr (G == 5) {
        fft5<SIGN>(z[0], z[1], z[2], z[3], z[4]);
    } else if constexpr (G == 10) {
        /* g = 2 g1 + g2, a = a1 + 5 a2: two 5-point transforms, W_10^(a1), one radix-2 stage */
        float2 e0 = z[0], e1 = z[2], e2 = z[4], e3 = z[6], e4 = z[8];
        float2 o0 = z[1], o1 = z[3], o2 = z[5], o3 = z[7], o4 = z[9];
        fft5<SIGN>(e0, e1, e2, e3, e4);
        fft5<SIGN>(o0, o1, o2, o3, o4);
        o1 = cmul(o1, make_float2(0.80901699437494742f, SIGN * 0.58778525229247313f));
        o2 = cmul(o2, make_float2(0.30901699437494742f, SIGN * 0.95105651629515357f));
        o3 = cmul(o3, make_float2(-0.30901699437494742f, SIGN * 0.95105651629515357f));
        o4 = cmul(o4, make_float2(-0.80901699437494742f, SIGN * 0.58778525229247313f));
        z[0] = cadd(e0, o0); z[1] = cadd(e1, o1); z[2] = cadd(e2, o2); z[3] = cadd(e3, o3); z[4] = cadd(e4, o4);
        z[5] = csub(e0, o0); z[6] = csub(e1, o1); z[7] = csub(e2, o2); z[8] = csub(e3, o3); z[9] = csub(e4, o4);
    } else if constexpr (G == 2) {
        const float2 a = z[0], b = z[1]; z[0] = cadd(a, b); z[1] = csub(a, b);
    } else if constexpr (G == 4) {
        fft4<SIGN>(z[0], z[1], z[2], z[3]);
    } else if constexpr (G == 8) {
        float2 e0 = z[0], e1 = z[2], e2 = z[4], e3 = z[6];
        float2 o0 = z[1], o1 = z[3], o2 = z[5], o3 = z[7];
        fft4<SIGN>(e0, e1, e2, e3);
        fft4<SIGN>(o0, o1, o2, o3);
        o1 = cw16<SIGN, 2>(o1); o2 = cw16<SIGN, 4>(o2); o3 = cw16<SIGN, 6>(o3);
        z[0] = cadd(e0, o0); z[1] = cadd(e1, o1); z[2] = cadd(e2, o2); z[3] = cadd(e3, o3);
        z[4] = csub(e0, o0); z[5] = csub(e1, o1); z[6] = csub(e2, o2); z[7] = csub(e3, o3);
    } else if constexpr (G == 32) {
        fft32<SIGN, false>(z);
    } else {
        static_assert(G == 16, "unsupported group count");
        fft16<SIGN>(z);
    }
}

/* Forward transform of long windows (R = 16 G sub-transforms, the scratch path) in two
 * passes, decimation in frequency: X[R j + q] = FFT_L over t of  W_M^(t q) * sum_a x[t + L a] W_R^(a q).
 * Pass 1 (k_fwd_split): carrier wipe-off (mixcarr table lookup, src/sdrcmn.c:650-662, and the
 * 1/(32 m) scale of pcorrelator, :764) fused into the R-point transform over a for 32
 * consecutive t per CTA - thread (warp g, lane) owns a = g + G j, 16 points in registers, then a
 * shared-memory transpose and the G-point stage - times W_M^(t q), written as row q of the
 * spectrum buffer.  Pass 2 (k_fft_rows): the L-point CTA transform of every row in place,
 * which leaves X[R j + q] at [q][j], the layout the row kernels read.  (k_fwd_if, used up to
 * R = 8, accumulates the R blocks per CTA: O(M R) per bin.)                                 */
template <int G, int DT>
__global__ void __launch_bounds__(32 * G)
k_fwd_split(const int8_t *__restrict__ win, int nlive, const uint8_t *__restrict__ cidx, int cidx_stride,
            float scale, const float2 *__restrict__ tw, int L, float2 *__restrict__ out)
{
    extern __shared__ float2 xsm[];                       /* [16 q2][G g][32 lanes] */
    __shared__ float2 s_lut[32];
    constexpr int R = 16 * G, A2 = (16 + G - 1) / G;      /* q2 values finished per warp (G = 5, 10: the last round is partial) */
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int t = blockIdx.x * 32 + lane, f = blockIdx.y;
    const int M = R * L;
    if (threadIdx.x < 32) s_lut[threadIdx.x] = make_float2(c_cosf[threadIdx.x] * scale, c_sinf[threadIdx.x] * scale);
    __syncthreads();
    const uint8_t *ci = cidx + (size_t)f * cidx_stride;
    float2 v[16];
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const int k = t + L * (g + G * j);
        v[j] = make_float2(0.f, 0.f);
        if (k < nlive) {
            const float2 lu = s_lut[ci[k] & 31];
            if (DT == 1) {
                const float x = (float)win[k];
                v[j] = make_float2(x * lu.x, x * lu.y);
            } else {
                const float x0 = (float)win[2 * k], x1 = (float)win[2 * k + 1];
                v[j] = make_float2(x0 * lu.x - x1 * lu.y, x0 * lu.y + x1 * lu.x);
            }
        }
    }
    fft16<-1>(v);                                         /* over j: v[q2] = U_g[q2] */
#pragma unroll
    for (int q2 = 0; q2 < 16; q2++) {
        float2 u = v[q2];
        if (q2 > 0) u = cmul(u, __ldg(&tw[(size_t)(q2 * g) * L]));          /* W_R^(q2 g) */
        xsm[(q2 * G + g) * 32 + lane] = u;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < A2; i++) {
        const int q2 = g + G * i;
        if (q2 >= 16) break;
        float2 z[G];
#pragma unroll
        for (int gg = 0; gg < G; gg++) z[gg] = xsm[(q2 * G + gg) * 32 + lane];
        fft_small<G, -1>(z);
#pragma unroll
        for (int q1 = 0; q1 < G; q1++) {
            const int q = q2 + 16 * q1;
            out[((size_t)f * R + q) * L + t] = cmul(z[q1], __ldg(&tw[t * q]));                 /* t q < L R = M */
        }
    }
}

/* the same first pass for the cluster path (R = 2, 4, 8 sub-transforms): one thread per t does
 * the wipe-off of its R samples x[t + L a] and the whole R-point stage in registers.  Together
 * with k_fft_rows it replaces k_fwd_if, whose CTAs each walk the entire window (R blocks, one
 * dependent load per block: 31 us for 71 bins of 32768 points, latency-bound).             */
template <int R, int DT>
__global__ void __launch_bounds__(256)
k_fwd_split_small(const int8_t *__restrict__ win, int nlive, const uint8_t *__restrict__ cidx, int cidx_stride,
                  float scale, const float2 *__restrict__ tw, int L, float2 *__restrict__ out)
{
    __shared__ float2 s_lut[32];
    const int t = blockIdx.x * 256 + threadIdx.x, f = blockIdx.y;
    const int M = R * L;
    if (threadIdx.x < 32) s_lut[threadIdx.x] = make_float2(c_cosf[threadIdx.x] * scale, c_sinf[threadIdx.x] * scale);
    __syncthreads();
    if (t >= L) return;
    const uint8_t *ci = cidx + (size_t)f * cidx_stride;
    float2 z[R];
#pragma unroll
    for (int a = 0; a < R; a++) {
        const int k = t + L * a;
        z[a] = make_float2(0.f, 0.f);
        if (k < nlive) {
            const float2 lu = s_lut[ci[k] & 31];
            if (DT == 1) {
                const float x = (float)win[k];
                z[a] = make_float2(x * lu.x, x * lu.y);
            } else {
                const float x0 = (float)win[2 * k], x1 = (float)win[2 * k + 1];
                z[a] = make_float2(x0 * lu.x - x1 * lu.y, x0 * lu.y + x1 * lu.x);
            }
        }
    }
    fft_small<R, -1>(z);
#pragma unroll
    for (int q = 0; q < R; q++)
        out[((size_t)f * R + q) * L + t] = (q == 0) ? z[0] : cmul(z[q], __ldg(&tw[t * q]));
}

/* operand order of the warp kernels (k_roww / k_rowd): lane l of a warp owns the points l + 32 e;
 * points e = 2h and 2h + 1 of a lane are stored next to each other so that one 16-byte load
 * fetches both: position of point j = (j >> 6) * 64 + (j & 31) * 2 + ((j >> 5) & 1)            */
__host__ __device__ __forceinline__ int acq_pair_pos(int j) { return (j >> 6) * 64 + (j & 31) * 2 + ((j >> 5) & 1); }

/* one-time reordering of the code-spectrum bank into that order (rows of 1024 points) */
__global__ void __launch_bounds__(1024)
k_pair_order(float2 *__restrict__ X)
{
    float2 *row = X + (size_t)blockIdx.x * 1024;
    const float2 v = row[threadIdx.x];
    __syncthreads();
    row[acq_pair_pos(threadIdx.x)] = v;
}

template <int L>
__global__ void __launch_bounds__(L / 16)
k_fft_rows(float2 *__restrict__ X, const float2 *__restrict__ tw, int R, int pair_order)
{
    extern __shared__ float2 fsm[];
    constexpr int T = L / 16;
    float2 *row = X + ((size_t)blockIdx.y * R + blockIdx.x) * L;
    float2 v[16];
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = row[threadIdx.x + e * T];
    cta_fft<L, -1>(v, fsm, tw, R);
    if (pair_order) {                                   /* layout of the warp kernels, see acq_pair_pos */
        __syncthreads();                                /* in place: every thread has loaded its points */
#pragma unroll
        for (int e = 0; e < 16; e++) row[acq_pair_pos(threadIdx.x + e * T)] = v[e];
        return;
    }
#pragma unroll
    for (int e = 0; e < 16; e++) row[threadIdx.x + e * T] = v[e];
}

/* recombination of the R = 16*G sub-transforms of a row as a real radix-R butterfly
 * (replaces the direct O(R) sum per lag of k_comb): one CTA of G warps takes 32 consecutive
 * t' of one (bin, PRN) row.  Thread (warp g, lane) loads y_q[t'] for q = g + G*j (coalesced:
 * a warp reads 256 contiguous bytes per q), applies conj(W_M^(t' q)), transforms over j in
 * registers (16 points), applies W_R^(a2 g), and after one shared-memory transpose finishes
 * the G-point transform over g for its share of a2.  Output a = a2 + 16*a1 is lag
 * t = t' + L*a: power, accumulation into P (cpxconv flagsum, src/sdrcmn.c:244-246) and the
 * PRN's argmax key (power bits : ~flat index, as in k_rowc) follow.  Only lags < n_eff are
 * formed.                                                                                */
template <int G>
__global__ void __launch_bounds__(32 * G)
k_comb_fft(const float2 *__restrict__ S, const float2 *__restrict__ tw, int L, int nfreq, int prn0, int f0,
           int nfc, float *__restrict__ P, int n_eff, int first_step, float invM2,
           unsigned long long *__restrict__ key, const int *__restrict__ frozen)
{
    extern __shared__ float2 xsm[];                       /* [16 a2][G g][32 lanes] */
    __shared__ unsigned long long s_k[32];
    constexpr int R = 16 * G, A2 = (16 + G - 1) / G;      /* a2 values finished per warp */
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int tp = blockIdx.x * 32 + lane, fl = blockIdx.y, f = f0 + fl, pc = blockIdx.z, p = prn0 + pc;
    if (frozen[p]) return;
    const int M = R * L;
    const float2 *s = S + ((size_t)pc * nfc + fl) * (size_t)M;
    float2 v[16];
#pragma unroll
    for (int j = 0; j < 16; j++) v[j] = __ldg(&s[(size_t)(g + G * j) * L + tp]);
    {
        /* conj(W_M^(t' (g + G j))) = wg * ws^j ; powers of ws by the doubling scheme of the
         * CTA transform (depth 4), so the rounding error does not grow with j            */
        float2 wg = __ldg(&tw[tp * g]); wg.y = -wg.y;                   /* tp g, tp G < L G <= M */
        float2 w1 = __ldg(&tw[tp * G]); w1.y = -w1.y;
        float2 w2 = cmul(w1, w1), w3 = cmul(w2, w1), w4 = cmul(w2, w2);
        float2 w5 = cmul(w4, w1), w6 = cmul(w3, w3), w7 = cmul(w4, w3), w8 = cmul(w4, w4);
        float2 w9 = cmul(w8, w1), w10 = cmul(w5, w5), w11 = cmul(w8, w3), w12 = cmul(w6, w6);
        float2 w13 = cmul(w8, w5), w14 = cmul(w7, w7), w15 = cmul(w8, w7);
        v[0] = cmul(v[0], wg);
        v[1] = cmul(v[1], cmul(w1, wg));   v[2] = cmul(v[2], cmul(w2, wg));   v[3] = cmul(v[3], cmul(w3, wg));
        v[4] = cmul(v[4], cmul(w4, wg));   v[5] = cmul(v[5], cmul(w5, wg));   v[6] = cmul(v[6], cmul(w6, wg));
        v[7] = cmul(v[7], cmul(w7, wg));   v[8] = cmul(v[8], cmul(w8, wg));   v[9] = cmul(v[9], cmul(w9, wg));
        v[10] = cmul(v[10], cmul(w10, wg)); v[11] = cmul(v[11], cmul(w11, wg)); v[12] = cmul(v[12], cmul(w12, wg));
        v[13] = cmul(v[13], cmul(w13, wg)); v[14] = cmul(v[14], cmul(w14, wg)); v[15] = cmul(v[15], cmul(w15, wg));
    }
    fft16<+1>(v);                                         /* over j: v[a2] = U_g[a2] */
#pragma unroll
    for (int a2 = 0; a2 < 16; a2++) {
        float2 u = v[a2];
        if (a2 > 0) { float2 w = __ldg(&tw[(size_t)(a2 * g) * L]); w.y = -w.y; u = cmul(u, w); }   /* W_R^(a2 g) */
        xsm[(a2 * G + g) * 32 + lane] = u;
    }
    __syncthreads();
    float *Prow = P + ((size_t)p * nfreq + f) * n_eff;
    float bv = -1.f; int bt = 0;
#pragma unroll
    for (int i = 0; i < A2; i++) {
        const int a2 = g + G * i;
        if (a2 >= 16) break;
        float2 z[G];
#pragma unroll
        for (int gg = 0; gg < G; gg++) z[gg] = xsm[(a2 * G + gg) * 32 + lane];
        fft_small<G, +1>(z);
        /* lags < n <= M/2 = 8 G L: only a = a2 + 16 a1 < 8 G can be one, i.e. a1 < G/2 - the other half of the
         * G-point stage is never formed (the compiler drops its butterflies with the loop)                    */
#pragma unroll
        for (int a1 = 0; a1 < (G + 1) / 2; a1++) {
            const int t = tp + L * (a2 + 16 * a1);
            if (t < n_eff) {
                float pw = (z[a1].x * z[a1].x + z[a1].y * z[a1].y) * invM2;
                if (!first_step) pw += Prow[t];
                Prow[t] = pw;
                if (pw > bv || (pw == bv && t < bt)) { bv = pw; bt = t; }
            }
        }
    }
    unsigned long long b = 0ull;
    if (bv >= 0.f)
        b = ((unsigned long long)__float_as_uint(bv) << 32) |
            (unsigned long long)(0xffffffffu - ((unsigned int)f * (unsigned int)n_eff + (unsigned int)bt));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long ob = __shfl_xor_sync(0xffffffffu, b, o);
        if (ob > b) b = ob;
    }
    if (lane == 0) s_k[g] = b;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < G; w++) if (s_k[w] > b) b = s_k[w];
        if (b) atomicMax(&key[p], b);
    }
}

/* checkacquisition() on the scratch path: the winning row of every PRN is read back from
 * the power matrix (masked mean, second maximum with the data[0] seed, C/N0, peak ratio,
 * decision, early-exit freeze: src/sdracq.c:71-95, 38-42)                               */
__global__ void __launch_bounds__(1024)
k_peak_hbm(const float *__restrict__ P, unsigned long long *__restrict__ key, int nfreq, int n_eff, int nsampchip,
           const double *__restrict__ freq, double ctime, double thresh, int step,
           sdrb200_acq_result_t *__restrict__ res, int *__restrict__ frozen)
{
    __shared__ acq_rowstat_t s_row;
    const int p = blockIdx.x;
    if (frozen[p]) return;
    const unsigned long long k = key[p];
    const unsigned int flat = k ? 0xffffffffu - (unsigned int)(k & 0xffffffffu) : 0u;
    const int f = (int)(flat / (unsigned int)n_eff), codei = (int)(flat - (unsigned int)f * (unsigned int)n_eff);
    const float maxp = __uint_as_float((unsigned int)(k >> 32));
    /* every thread proposes the known peak: the block argmax of acq_row_stats returns it */
    acq_row_stats(P + ((size_t)p * nfreq + f) * n_eff, n_eff, nsampchip, maxp, codei, &s_row);
    __syncthreads();
    if (threadIdx.x == 0) {
        const acq_rowstat_t r = s_row;
        sdrb200_acq_result_t o;
        o.maxp = (double)r.maxp; o.maxp2 = (double)r.max2;
        o.meanp = r.sum / (double)r.cnt;
        o.cn0 = 10 * log10(o.maxp / o.meanp / ctime);
        o.peakr = o.maxp / o.maxp2;
        o.codei = r.codei; o.freqi = f; o.acqfreq = freq[f];
        o.acquired = (o.peakr > thresh) ? 1 : 0;
        o.nintg = step + 1;
        res[p] = o;
        if (o.acquired) frozen[p] = 1;
        key[p] = 0ull;                                             /* next integration step */
    }
}

/* checkacquisition() for one PRN from its row statistics (src/sdracq.c:71-95)
 * plus the early exit of sdracquisition() (src/sdracq.c:38-42)                 */
__global__ void k_final(const acq_rowstat_t *__restrict__ rows, int nfreq, const double *__restrict__ freq,
                        double ctime, double thresh, int step, sdrb200_acq_result_t *__restrict__ res,
                        int *__restrict__ frozen)
{
    const int p = blockIdx.x;
    if (frozen[p]) return;
    __shared__ float s_v[32]; __shared__ int s_f[32];
    float best = -1.f; int bf = 0x7fffffff;
    for (int f = threadIdx.x; f < nfreq; f += blockDim.x) {
        float v = rows[(size_t)p * nfreq + f].maxp;
        if (v > best) { best = v; bf = f; }                       /* lowest bin wins ties */
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, best, o);
        int   of = __shfl_xor_sync(0xffffffffu, bf, o);
        if (ov > best || (ov == best && of < bf)) { best = ov; bf = of; }
    }
    if (lane == 0) { s_v[warp] = best; s_f[warp] = bf; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < nwarp; w++)
            if (s_v[w] > best || (s_v[w] == best && s_f[w] < bf)) { best = s_v[w]; bf = s_f[w]; }
        const acq_rowstat_t r = rows[(size_t)p * nfreq + bf];
        sdrb200_acq_result_t o;
        o.maxp = (double)r.maxp; o.maxp2 = (double)r.max2;
        o.meanp = r.sum / (double)r.cnt;
        o.cn0 = 10 * log10(o.maxp / o.meanp / ctime);
        o.peakr = o.maxp / o.maxp2;
        o.codei = r.codei; o.freqi = bf; o.acqfreq = freq[bf];
        o.acquired = (o.peakr > thresh) ? 1 : 0;
        o.nintg = step + 1;
        res[p] = o;
        if (o.acquired) frozen[p] = 1;
    }
}

/* ---------------------------------------------------- fused cluster kernel
 * One thread-block CLUSTER of R CTAs per (PRN, bin) row.  CTA q forms the
 * spectrum product and the inverse L-point sub-transform of residue class q
 * (cpxconv, src/sdrcmn.c:236-242), scales by W_M^(-t q) and leaves the result in
 * its own shared memory.  After a cluster barrier every CTA gathers an L/R-wide
 * slice of all R sub-transforms through distributed shared memory and finishes
 * the radix-R butterfly for the lag blocks that are kept (lags < n <= M/2 need
 * only R/2 of the R outputs), so neither the sub-transforms nor - for a single
 * integration - the power matrix ever go to HBM.
 *
 * mode 0 (search): power (+ optional accumulation into P, cpxconv flagsum,
 *   src/sdrcmn.c:244-246) and the PRN's global argmax as one 64-bit atomicMax
 *   key = power bits : ~flat index, i.e. maxvd()'s first-index-wins over the
 *   reference's [bin][lag] layout (src/sdracq.c:76-77, sdrcmn.c:461-476).
 * mode 1 (peak): only the winning row of every PRN (recomputed, or read back
 *   from P when P is kept): masked mean and second maximum around the known
 *   peak, C/N0, peak ratio, decision, early-exit freeze (src/sdracq.c:79-94,
 *   38-42).                                                                   */
struct acq_rowc_args {
    const float2 *Xs, *Cs, *tw;
    float *P;                               /* NULL: power matrix not kept       */
    unsigned long long *key;                /* [nprn]                            */
    int *frozen;
    sdrb200_acq_result_t *res;
    const double *freq;
    double ctime, thresh;
    int nfreq, n_eff, nsampchip, first_step, step;
    float invM2;
    int xs;                                 /* spectra of replica p start at bin p * xs (0: one grid for all) */
};

struct acq_part_t { double sum; int cnt; float max2; };

__device__ __forceinline__ void cl_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cl_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cl_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

template <int L, int R, int MODE>
__global__ void __launch_bounds__(L / 16, (L <= 4096) ? 4 : (L <= 8192 ? 2 : 1))
k_rowc(const acq_rowc_args A)
{
    extern __shared__ float2 fsm[];
    __shared__ unsigned long long s_key[32];
    __shared__ acq_part_t s_part[32];
    __shared__ acq_part_t s_cta;
    constexpr int T = L / 16;
    constexpr int NA = (R >= 2) ? (R + 1) / 2 : 1;    /* lag blocks kept (lags < n <= M/2) */
    constexpr int J = (16 + R - 1) / R;               /* slice columns per thread */
    constexpr int CW = (L + R - 1) / R;               /* slice columns per CTA    */
    constexpr bool EVEN = (16 % R) == 0;              /* R = 5, 10 (M = 5 * 2^k): the last round of columns is partial */
    constexpr int NOUT = (R == 1) ? 16 : NA * J;
    const int q = blockIdx.x, p = blockIdx.z, tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    cg::cluster_group cluster = cg::this_cluster();
    if (A.frozen[p]) return;                          /* whole cluster leaves     */
    int f = blockIdx.y, codei = 0;
    float maxp = 0.f;
    if (MODE == 1) {
        const unsigned long long k = A.key[p];
        const unsigned int flat = k ? 0xffffffffu - (unsigned int)(k & 0xffffffffu) : 0u;
        f = (int)(flat / (unsigned int)A.n_eff);
        codei = (int)(flat - (unsigned int)f * (unsigned int)A.n_eff);
        maxp = __uint_as_float((unsigned int)(k >> 32));
    }
    /* outputs of this thread in increasing lag order: pw[i] <-> lag tout(i) */
    float pw[NOUT];
    auto tout = [&](int i) -> int {
        if (R == 1) return tid + i * T;
        const int c = tid + (i % J) * T;
        if (!EVEN && (c >= CW || q * CW + c >= L)) return 0x7fffffff;        /* no such column: never < n_eff */
        return q * CW + c + L * (i / J);
    };
    bool from_hbm = (MODE == 1 && A.P != nullptr);
    if (from_hbm) {
        /* the accumulated row is in HBM already: read this CTA's share of it */
        const float *Prow = A.P + ((size_t)p * A.nfreq + f) * A.n_eff;
#pragma unroll
        for (int i = 0; i < NOUT; i++) { const int t = tout(i); pw[i] = (t < A.n_eff) ? Prow[t] : 0.f; }
    } else {
        const float2 *x = A.Xs + ((size_t)(p * A.xs + f) * R + q) * L;
        const float2 *c = A.Cs + ((size_t)p * R + q) * L;
        float2 v[16];
        float2 w[fft_plan_t<L>::N16];
        cta_fft_twiddles<L>(A.tw, R, w);
#pragma unroll
        for (int e = 0; e < 16; e++) v[e] = cmulc(__ldg(&x[tid + e * T]), __ldg(&c[tid + e * T]));
        cta_fft_pre<L, +1>(v, fsm, w);
        if constexpr (R == 1) {
#pragma unroll
            for (int e = 0; e < 16; e++) pw[e] = (v[e].x * v[e].x + v[e].y * v[e].y) * A.invM2;
        } else {
            /* y_q[t'] into own shared memory; W_M^(-t' q) is applied by the reader */
#pragma unroll
            for (int e = 0; e < 16; e++) fsm[tid + e * T] = v[e];
            /* publish: CTA barrier, then ONE thread issues the cluster-scope release fence
             * (cumulative over the stores the barrier made visible to it; every thread doing
             * it costs a memory barrier per warp), relaxed arrive by all.  Leaving the fence
             * out is 5 % faster and NOT safe: remote reads were seen to overtake the stores
             * (tests/test_gpu_acquisition.py::test_acq_cluster_exchange_repeatable failed once
             * in four runs of 60 repeats without it).                                      */
            __syncthreads();
            if (tid == 0) asm volatile("fence.acq_rel.cluster;" ::: "memory");
            cl_arrive_relaxed();
            float2 wj[J];
#pragma unroll
            for (int j = 0; j < J; j++) wj[j] = __ldg(&A.tw[q * CW + tid + j * T]);
            cl_wait();
            const float2 *zs[R];
#pragma unroll
            for (int r = 0; r < R; r++) zs[r] = cluster.map_shared_rank(fsm, r);
#pragma unroll
            for (int j = 0; j < J; j++) {
                const int tp = q * CW + tid + j * T;
                if (!EVEN && (tid + j * T >= CW || tp >= L)) {
#pragma unroll
                    for (int a = 0; a < NA; a++) pw[a * J + j] = 0.f;
                    continue;
                }
                float2 z[R];
#pragma unroll
                for (int r = 0; r < R; r++) z[r] = zs[r][tp];
                /* z_r *= W_M^(-t' r) = conj(w)^r */
                float2 wp = wj[j];
#pragma unroll
                for (int r = 1; r < R; r++) {
                    z[r] = cmulc(z[r], wp);
                    if (r + 1 < R) wp = cmul(wp, wj[j]);
                }
                float2 y[NA];
                if constexpr (R == 2) {
                    y[0] = cadd(z[0], z[1]);
                } else if constexpr (R == 4) {
                    y[0] = cadd(cadd(z[0], z[2]), cadd(z[1], z[3]));
                    y[1] = cadd(csub(z[0], z[2]), crot<+1>(csub(z[1], z[3])));
                } else if constexpr (R == 3 || R == 5 || R == 6 || R == 10 || R == 12) {
                    fft_small<R, +1>(z);
#pragma unroll
                    for (int a = 0; a < NA; a++) y[a] = z[a];
                } else if constexpr (R == 16) {
                    /* 16-point inverse transform; only y[0..7] are used (lags < n <= M/2) */
                    fft16<+1>(z);
#pragma unroll
                    for (int a = 0; a < NA; a++) y[a] = z[a];
                } else {
                    /* R == 8: even / odd 4-point transforms, then W_8^(-a) */
                    float2 e0 = z[0], e1 = z[2], e2 = z[4], e3 = z[6];
                    float2 o0 = z[1], o1 = z[3], o2 = z[5], o3 = z[7];
                    fft4<+1>(e0, e1, e2, e3);
                    fft4<+1>(o0, o1, o2, o3);
                    y[0] = cadd(e0, o0);
                    y[1] = cadd(e1, cw16<+1, 2>(o1));
                    y[2] = cadd(e2, cw16<+1, 4>(o2));
                    y[3] = cadd(e3, cw16<+1, 6>(o3));
                }
#pragma unroll
                for (int a = 0; a < NA; a++) pw[a * J + j] = (y[a].x * y[a].x + y[a].y * y[a].y) * A.invM2;
            }
            cl_arrive_relaxed();                      /* done with the peers' shared memory */
        }
    }

    if (MODE == 0) {
        float *Prow = A.P ? A.P + ((size_t)p * A.nfreq + f) * A.n_eff : nullptr;
        float bv = -1.f; int bt = 0;
#pragma unroll
        for (int i = 0; i < NOUT; i++) {
            const int t = tout(i);
            if (t < A.n_eff) {
                float x = pw[i];
                if (Prow) { if (!A.first_step) x += Prow[t]; Prow[t] = x; }     /* cpxconv flagsum */
                if (x > bv) { bv = x; bt = t; }
            }
        }
        unsigned long long b = 0ull;
        if (bv >= 0.f)
            b = ((unsigned long long)__float_as_uint(bv) << 32) |
                (unsigned long long)(0xffffffffu - ((unsigned int)f * (unsigned int)A.n_eff + (unsigned int)bt));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long ob = __shfl_xor_sync(0xffffffffu, b, o);
            if (ob > b) b = ob;
        }
        if (lane == 0) s_key[warp] = b;
        __syncthreads();
        if (warp == 0) {
            b = (lane < T / 32) ? s_key[lane] : 0ull;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long ob = __shfl_xor_sync(0xffffffffu, b, o);
                if (ob > b) b = ob;
            }
            if (lane == 0 && b) atomicMax(&A.key[p], b);
        }
        if (R > 1) cl_wait();                         /* peers may still be reading fsm */
        return;
    }

    /* MODE 1: masked statistics around the known peak (src/sdracq.c:79-88) */
    int exs = codei - 2 * A.nsampchip; if (exs < 0) exs += A.n_eff;
    int exe = codei + 2 * A.nsampchip; if (exe >= A.n_eff) exe -= A.n_eff;
    double sum = 0.0; int cnt = 0; float m2 = -1.f;
#pragma unroll
    for (int i = 0; i < NOUT; i++) {
        const int t = tout(i);
        if (t < A.n_eff) {
            const bool inc = (exs <= exe) ? (t < exs || t > exe) : (t < exs && t > exe);
            if (inc) { sum += (double)pw[i]; cnt++; }
            if ((inc || t == 0) && pw[i] > m2) m2 = pw[i];          /* maxvd seeds with data[0] */
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    }
    if (lane == 0) { s_part[warp].sum = sum; s_part[warp].cnt = cnt; s_part[warp].max2 = m2; }
    __syncthreads();
    if (tid == 0) {
        const int nw = (T + 31) / 32;
        for (int w = 1; w < nw; w++) { sum += s_part[w].sum; cnt += s_part[w].cnt; m2 = fmaxf(m2, s_part[w].max2); }
        s_cta.sum = sum; s_cta.cnt = cnt; s_cta.max2 = m2;
    }
    if (R > 1 && !from_hbm) cl_wait();                /* closes the arrive.relaxed above */
    cl_arrive_release();
    cl_wait();
    if (q == 0 && tid == 0) {
        for (int r = 1; r < R; r++) {
            const acq_part_t *o = cluster.map_shared_rank(&s_cta, r);
            sum += o->sum; cnt += o->cnt; m2 = fmaxf(m2, o->max2);
        }
        sdrb200_acq_result_t o;
        o.maxp = (double)maxp; o.maxp2 = (double)m2;
        o.meanp = sum / (double)cnt;                               /* meanvd, sdrcmn.c:487-497 */
        o.cn0 = 10 * log10(o.maxp / o.meanp / A.ctime);            /* sdracq.c:84 */
        o.peakr = o.maxp / o.maxp2;                                /* sdracq.c:88 */
        o.codei = codei; o.freqi = f; o.acqfreq = A.freq[p * A.xs + f];
        o.acquired = (o.peakr > A.thresh) ? 1 : 0;
        o.nintg = A.step + 1;
        A.res[p] = o;
        if (o.acquired) A.frozen[p] = 1;
        A.key[p] = 0ull;                                           /* next integration step */
    }
    cl_arrive_relaxed();
    cl_wait();
}

#define ROWP_MAXPRN 1024

/* ------------------------------------------------ one warp per sub-transform
 * k_roww: M = 32 x 1024.  A cluster of 8 CTAs x 4 warps owns a row; every WARP transforms one of
 * the 32 residue classes by itself (warp_fft1024: 32 points per lane, one transpose through the
 * warp's own shared-memory tile, no CTA barrier inside the transform), and every thread finishes
 * ONE column t' of the radix-32 recombination - 32 remote 8-byte loads, twiddles, a 32-point
 * transform of which only the 16 outputs below M/2 are formed (lags < n <= M/2).
 * Against the 8 x 4096 split this trades one shared-memory exchange per point for a longer
 * register butterfly: 32 instead of 48 bytes of shared-memory traffic per point, a third fewer
 * instructions (the pruned stage is 5 of 15 radix-2 levels instead of 3), two CTA barriers per row
 * instead of six.  The clusters are persistent (as many as the GPU holds at once, each walking a
 * contiguous run of rows numbered over the PRNs still searched, src/sdracq.c:38-42) and rows are
 * pipelined over two tiles per warp:
 *   warp, row i (tile b = i & 1): operands -> product -> first 32-point pass in registers,
 *        wait FREE[b]  (every warp of the cluster has gathered row i-2 out of the tiles b),
 *        transpose + second pass through tile b, results left there, release-arrive READY[b] at all
 *        eight CTAs; wait READY[1-b], gather column t' of row i-1 from the 32 tiles (1-b), recombine,
 *        power, argmax in registers until the PRN changes, arrive FREE[1-b] at all eight CTAs.
 * READY/FREE are mbarriers counting one arrival per warp of the cluster; a warp only waits for what
 * its peers did a whole sub-transform ago, and there is no CTA barrier in the loop.
 * MODE 1 is checkacquisition for the winning row of each PRN (one cluster per PRN).            */
template <int MODE>
__global__ void __launch_bounds__(128, 3)
k_roww(const acq_rowc_args A, int nprn)
{
    extern __shared__ __align__(16) float2 psm[];         /* [2 buffers][4 warps][WFFT_TILE] */
    __shared__ __align__(8) unsigned long long s_ready[2], s_free[2];
    __shared__ unsigned short s_act[ROWP_MAXPRN];
    __shared__ int s_na;
    __shared__ acq_part_t s_part[4];
    __shared__ acq_part_t s_cta;
    constexpr int CS = 8, WPC = 4, R = CS * WPC, L = 1024, NT = 32 * WPC, NA = R / 2, BUF = WPC * WFFT_TILE;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t c = cl_rank();
    const int cid = blockIdx.x / CS, ncl = gridDim.x / CS;
    const int q = (int)c * WPC + warp;                    /* residue class of this warp */
    const int tcol = (int)c * (L / CS) + tid;             /* column t' of this thread   */

    if (MODE == 1) { if (A.frozen[cid]) return; }         /* whole cluster leaves */
    if (tid < 32) {
        int na = 0;
        if (MODE == 0) {
            for (int base = 0; base < nprn; base += 32) {
                const int p = base + lane;
                const bool on = (p < nprn) && !A.frozen[p];
                const unsigned m = __ballot_sync(0xffffffffu, on);
                if (on) s_act[na + __popc(m & ((1u << lane) - 1u))] = (unsigned short)p;
                na += __popc(m);
            }
        }
        if (lane == 0) {
            s_na = na;
            cl_mbar_init(&s_ready[0], R); cl_mbar_init(&s_ready[1], R);      /* one arrival per WARP of the cluster */
            cl_mbar_init(&s_free[0], R);  cl_mbar_init(&s_free[1], R);
            cl_fence_mbar_init();
        }
    }
    __syncthreads();
    cl_sync_all();

    int r0, r1, codei = 0, fpk = 0;
    float maxp = 0.f;
    bool from_hbm = false;
    if (MODE == 0) {
        const int nrows = s_na * A.nfreq;
        const int rpc = (nrows + ncl - 1) / ncl;
        r0 = cid * rpc; r1 = min(r0 + rpc, nrows);
    } else {
        const unsigned long long k = A.key[cid];
        const unsigned int flat = k ? 0xffffffffu - (unsigned int)(k & 0xffffffffu) : 0u;
        fpk = (int)(flat / (unsigned int)A.n_eff);
        codei = (int)(flat - (unsigned int)fpk * (unsigned int)A.n_eff);
        maxp = __uint_as_float((unsigned int)(k >> 32));
        from_hbm = (A.P != nullptr);
        r0 = 0; r1 = from_hbm ? 0 : 1;
    }

    float2 w1 = __ldg(&A.tw[lane * R]); w1.y = -w1.y;     /* exp(+2 pi i lane / 1024)      */
    const float2 wc = __ldg(&A.tw[tcol]);                 /* W_M^(t'), used conjugated     */

    float bv = -1.f; unsigned int bflat = 0; int pcur = -1;
    auto flush = [&]() {
        unsigned long long b = 0ull;
        if (bv >= 0.f) b = ((unsigned long long)__float_as_uint(bv) << 32) | (unsigned long long)(0xffffffffu - bflat);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long ob = __shfl_xor_sync(0xffffffffu, b, o);
            if (ob > b) b = ob;
        }
        if (lane == 0 && b && pcur >= 0) atomicMax(&A.key[pcur], b);
        bv = -1.f; bflat = 0;
    };

    float pw[NA];                                         /* MODE 1 keeps the row's powers */
    int pp = 0, pf = 0, i = 0;
    for (int r = r0; r <= r1 && r0 < r1; r++, i++) {
        const int b = i & 1;
        int p = 0, f = 0;
        if (r < r1) {
            if (MODE == 0) { const int pa = r / A.nfreq; f = r - pa * A.nfreq; p = s_act[pa]; }
            else { p = cid; f = fpk; }
            float2 *tile = psm + (b * WPC + warp) * WFFT_TILE;
            const float2 *x = A.Xs + ((size_t)f * R + q) * L;
            const float2 *cc = A.Cs + ((size_t)p * R + q) * L;
            float2 v[32];
#pragma unroll
            for (int h = 0; h < 16; h++) {                /* pair order: points 2h, 2h+1 of this lane in 16 bytes */
                const float4 xa = __ldg((const float4 *)x + 32 * h + lane), ca = __ldg((const float4 *)cc + 32 * h + lane);
                v[2 * h] = cmulc(make_float2(xa.x, xa.y), make_float2(ca.x, ca.y));
                v[2 * h + 1] = cmulc(make_float2(xa.z, xa.w), make_float2(ca.z, ca.w));
            }
            warp_fft1024<+1>(v, tile, w1, [&]() { if (i >= 2) cl_mbar_wait(&s_free[b], (uint32_t)(((i >> 1) - 1) & 1)); });
#pragma unroll
            for (int e = 0; e < 32; e++) tile[lane + 32 * e] = v[e];
            /* the warp publishes its own tile: no CTA barrier anywhere in the row loop */
            __syncwarp();
            if (lane < CS) cl_mbar_arrive_remote(cl_map(&s_ready[b], (uint32_t)lane));
        }
        if (i >= 1) {
            const int pb = b ^ 1;
            cl_mbar_wait(&s_ready[pb], (uint32_t)(((i - 1) >> 1) & 1));
            float2 z[32];
            const uint32_t mine = cl_smem_u32(psm + pb * BUF + tcol);
#pragma unroll
            for (int rr = 0; rr < CS; rr++) {
                uint32_t base;
                asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(base) : "r"(mine), "r"(rr));
#pragma unroll
                for (int ww = 0; ww < WPC; ww++) {
                    float2 t;
                    asm volatile("ld.shared::cluster.v2.f32 {%0, %1}, [%2];" : "=f"(t.x), "=f"(t.y)
                                 : "r"(base + (uint32_t)(ww * WFFT_TILE * sizeof(float2))) : "memory");
                    z[rr * WPC + ww] = t;
                }
            }
            twiddle32<true>(z, wc);                       /* z_q *= conj(W_M^(t'))^q */
            fft32<+1, true>(z);
#pragma unroll
            for (int a = 0; a < NA; a++) pw[a] = (z[a].x * z[a].x + z[a].y * z[a].y) * A.invM2;
            /* this warp's share of FREE for all eight owners.  Reads-before-arrive only, so the light
             * arrive (release at CTA scope) is enough - but the loads must have RETURNED in every lane,
             * and ptxas hoists an arrive that does not depend on them right behind the loads (repeated
             * searches then differed).  pw[0] depends on all 32 loads of a lane and the vote on every
             * lane's pw[0]; the arrival count 1 + (a power is negative: never) is computed from it.    */
            const uint32_t one = 1u + (__ballot_sync(0xffffffffu, pw[0] < 0.f) != 0u);
            if (lane < CS) cl_mbar_arrive_remote_light(cl_map(&s_free[pb], (uint32_t)lane), one);

            if (MODE == 0) {
                if (pp != pcur) { flush(); pcur = pp; }
                /* branch-free: lags at or beyond n_eff get power -1 and never win; the accumulating path
                 * (cpxconv flagsum, src/sdrcmn.c:244-246) is a uniform branch around the whole loop     */
                if (A.P) {
                    float *Prow = A.P + ((size_t)pp * A.nfreq + pf) * A.n_eff;
#pragma unroll
                    for (int a = 0; a < NA; a++) {
                        const int t = tcol + L * a;
                        if (t < A.n_eff) { float xw = pw[a]; if (!A.first_step) xw += Prow[t]; Prow[t] = xw; pw[a] = xw; }
                        else pw[a] = -1.f;
                    }
                } else {
#pragma unroll
                    for (int a = NA / 2; a < NA; a++) pw[a] = (tcol + L * a < A.n_eff) ? pw[a] : -1.f;   /* n_eff > M/4 */
                }
                float lv = -1.f; int la = 0;
#pragma unroll
                for (int a = 0; a < NA; a++) { const bool up = pw[a] > lv; lv = up ? pw[a] : lv; la = up ? a : la; }
                if (lv > bv) { bv = lv; bflat = (unsigned int)pf * (unsigned int)A.n_eff + (unsigned int)(tcol + L * la); }
            }
        }
        pp = p; pf = f;
    }
    if (MODE == 0) {
        flush();
        cl_sync_all();                                    /* nobody leaves while a peer may still read */
        return;
    }

    /* MODE 1: masked statistics around the known peak (src/sdracq.c:79-88) */
    if (from_hbm) {
        const float *Prow = A.P + ((size_t)cid * A.nfreq + fpk) * A.n_eff;
#pragma unroll
        for (int a = 0; a < NA; a++) { const int t = tcol + L * a; pw[a] = (t < A.n_eff) ? Prow[t] : 0.f; }
    }
    int exs = codei - 2 * A.nsampchip; if (exs < 0) exs += A.n_eff;
    int exe = codei + 2 * A.nsampchip; if (exe >= A.n_eff) exe -= A.n_eff;
    double sum = 0.0; int cnt = 0; float m2 = -1.f;
#pragma unroll
    for (int a = 0; a < NA; a++) {
        const int t = tcol + L * a;
        if (t < A.n_eff) {
            const bool inc = (exs <= exe) ? (t < exs || t > exe) : (t < exs && t > exe);
            if (inc) { sum += (double)pw[a]; cnt++; }
            if ((inc || t == 0) && pw[a] > m2) m2 = pw[a];            /* maxvd seeds with data[0] */
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    }
    if (lane == 0) { s_part[warp].sum = sum; s_part[warp].cnt = cnt; s_part[warp].max2 = m2; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < WPC; w++) { sum += s_part[w].sum; cnt += s_part[w].cnt; m2 = fmaxf(m2, s_part[w].max2); }
        s_cta.sum = sum; s_cta.cnt = cnt; s_cta.max2 = m2;
    }
    cl_sync_all();
    if (c == 0 && tid == 0) {
        for (uint32_t rr = 1; rr < CS; rr++) {
            const uint32_t ad = cl_map(&s_cta, rr);
            double os; int oc; float om;
            asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(os) : "r"(ad) : "memory");
            asm volatile("ld.shared::cluster.s32 %0, [%1];" : "=r"(oc) : "r"(ad + 8) : "memory");
            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(om) : "r"(ad + 12) : "memory");
            sum += os; cnt += oc; m2 = fmaxf(m2, om);
        }
        sdrb200_acq_result_t o;
        o.maxp = (double)maxp; o.maxp2 = (double)m2;
        o.meanp = sum / (double)cnt;                               /* meanvd, sdrcmn.c:487-497 */
        o.cn0 = 10 * log10(o.maxp / o.meanp / A.ctime);            /* sdracq.c:84 */
        o.peakr = o.maxp / o.maxp2;                                /* sdracq.c:88 */
        o.codei = codei; o.freqi = fpk; o.acqfreq = A.freq[fpk];
        o.acquired = (o.peakr > A.thresh) ? 1 : 0;
        o.nintg = A.step + 1;
        A.res[cid] = o;
        if (o.acquired) A.frozen[cid] = 1;
        A.key[cid] = 0ull;                                         /* next integration step */
    }
    cl_sync_all();
}

/* ------------------------------------------------ bulk-copy exchange
 * k_rowd: the same warp-per-sub-transform arithmetic as k_roww, but the sub-transforms are PUSHED:
 * a cluster is 8 CTAs x 4 warps (two CTAs per SM); a warp leaves its 1024
 * results in its tile and one lane sends the eight 1 KB slices to the eight CTAs of the cluster with
 * cp.async.bulk (shared::cta -> shared::cluster, UBLKCP), each copy completing bytes on the
 * receiver's READY mbarrier.  No cluster-scope fence (the release of a generic store to remote
 * readers is a MEMBAR.GPU per warp and row: 13 % of k_roww's stall samples), no remote loads
 * (8 % more): a thread reads its 32 inputs of the recombination from LOCAL shared memory.  Two
 * receive buffers per CTA pipeline the rows; FREE mbarriers (32 warp arrivals) hand a buffer back to
 * the producers.  With eight warps per SM the registers allow the operands of row i+1 to be loaded
 * while row i-1 is recombined.                                                                  */
__device__ __forceinline__ void bulk_push(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t mbar_cluster)
{
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(mbar_cluster) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(void *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(cl_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_cta(void *bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(cl_smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}

#define ROWD_CS 8
#define ROWD_WPC 4
#define ROWD_SMEM (sizeof(float2) * (ROWD_WPC * WFFT_TILE + 2 * 32 * (1024 / ROWD_CS)))
template <int MODE>
__global__ void __launch_bounds__(32 * ROWD_WPC, 2)
k_rowd(const acq_rowc_args A, int nprn)
{
    extern __shared__ __align__(16) float2 psm[];         /* [8 warps][WFFT_TILE] tiles, then recv[2][32][256] */
    __shared__ __align__(8) unsigned long long s_ready[2], s_free[2];
    __shared__ unsigned short s_act[ROWP_MAXPRN];
    __shared__ int s_na;
    __shared__ acq_part_t s_part[ROWD_WPC];
    __shared__ acq_part_t s_cta;
    constexpr int CS = ROWD_CS, WPC = ROWD_WPC, R = CS * WPC, L = 1024, NA = R / 2, NCOL = L / CS;
    constexpr uint32_t SLICE = NCOL * sizeof(float2);     /* bytes one warp sends to one CTA */
    static_assert(R == 32 && NCOL == 32 * WPC, "one column per thread");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t c = cl_rank();
    const int cid = blockIdx.x / CS, ncl = gridDim.x / CS;
    const int q = (int)c * WPC + warp;                    /* residue class of this warp */
    const int tcol = (int)c * NCOL + tid;                 /* column t' of this thread   */
    float2 *tile = psm + warp * WFFT_TILE;
    float2 *recv = psm + WPC * WFFT_TILE;

    if (MODE == 1) { if (A.frozen[cid]) return; }         /* whole cluster leaves */
    if (tid < 32) {
        int na = 0;
        if (MODE == 0) {
            for (int base = 0; base < nprn; base += 32) {
                const int p = base + lane;
                const bool on = (p < nprn) && !A.frozen[p];
                const unsigned m = __ballot_sync(0xffffffffu, on);
                if (on) s_act[na + __popc(m & ((1u << lane) - 1u))] = (unsigned short)p;
                na += __popc(m);
            }
        }
        if (lane == 0) {
            s_na = na;
            cl_mbar_init(&s_ready[0], 1); cl_mbar_init(&s_ready[1], 1);      /* one arming arrival + 32 KB of copies */
            cl_mbar_init(&s_free[0], R);  cl_mbar_init(&s_free[1], R);       /* one arrival per warp of the cluster  */
            cl_fence_mbar_init();
            mbar_expect_tx(&s_ready[0], R * SLICE);
            mbar_expect_tx(&s_ready[1], R * SLICE);
        }
    }
    __syncthreads();
    cl_sync_all();

    int r0, r1, codei = 0, fpk = 0;
    float maxp = 0.f;
    bool from_hbm = false;
    if (MODE == 0) {
        const int nrows = s_na * A.nfreq;
        const int rpc = (nrows + ncl - 1) / ncl;
        r0 = cid * rpc; r1 = min(r0 + rpc, nrows);
    } else {
        const unsigned long long k = A.key[cid];
        const unsigned int flat = k ? 0xffffffffu - (unsigned int)(k & 0xffffffffu) : 0u;
        fpk = (int)(flat / (unsigned int)A.n_eff);
        codei = (int)(flat - (unsigned int)fpk * (unsigned int)A.n_eff);
        maxp = __uint_as_float((unsigned int)(k >> 32));
        from_hbm = (A.P != nullptr);
        r0 = 0; r1 = from_hbm ? 0 : 1;
    }

    float2 w1 = __ldg(&A.tw[lane * R]); w1.y = -w1.y;     /* exp(+2 pi i lane / 1024)      */
    const float2 wc = __ldg(&A.tw[tcol]);                 /* W_M^(t'), used conjugated     */

    float bv = -1.f; unsigned int bflat = 0; int pcur = -1;
    auto flush = [&]() {
        unsigned long long b = 0ull;
        if (bv >= 0.f) b = ((unsigned long long)__float_as_uint(bv) << 32) | (unsigned long long)(0xffffffffu - bflat);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long ob = __shfl_xor_sync(0xffffffffu, b, o);
            if (ob > b) b = ob;
        }
        if (lane == 0 && b && pcur >= 0) atomicMax(&A.key[pcur], b);
        bv = -1.f; bflat = 0;
    };
    auto row_of = [&](int r, int &p, int &f) {
        if (MODE == 0) { const int pa = r / A.nfreq; f = r - pa * A.nfreq; p = s_act[pa]; }
        else { p = cid; f = fpk; }
    };

    float pw[NA];                                         /* MODE 1 keeps the row's powers */
    float4 xv[16], cv[16];                                /* operands of the coming row, pair order */
    int pn = 0, fn = 0, pc = 0, fc = 0, pq = 0, fq = 0, i = 0;
    if (r0 < r1) {
        row_of(r0, pn, fn);
        const float2 *x = A.Xs + ((size_t)fn * R + q) * L, *cc = A.Cs + ((size_t)pn * R + q) * L;
#pragma unroll
        for (int h = 0; h < 16; h++) { xv[h] = __ldg((const float4 *)x + 32 * h + lane); cv[h] = __ldg((const float4 *)cc + 32 * h + lane); }
    }
    for (int r = r0; r <= r1 && r0 < r1; r++, i++) {
        const int b = i & 1;
        pq = pc; fq = fc; pc = pn; fc = fn;
        if (r < r1) {
            float2 v[32];
#pragma unroll
            for (int h = 0; h < 16; h++) {
                v[2 * h] = cmulc(make_float2(xv[h].x, xv[h].y), make_float2(cv[h].x, cv[h].y));
                v[2 * h + 1] = cmulc(make_float2(xv[h].z, xv[h].w), make_float2(cv[h].z, cv[h].w));
            }
            /* the tile is the source of the previous row's copies: they must have read it */
            warp_fft1024<+1>(v, tile, w1, [&]() { if (lane == 0) bulk_wait_read0(); __syncwarp(); });
#pragma unroll
            for (int e = 0; e < 32; e++) tile[lane + 32 * e] = v[e];
            fence_proxy_async_smem();                     /* generic-proxy stores -> visible to the copy engine */
            __syncwarp();
            if (i >= 2) mbar_wait_cta(&s_free[b], (uint32_t)(((i >> 1) - 1) & 1));  /* receivers are done with buffer b (no data rides on FREE: CTA-scope wait) */
            if (lane == 0) {
                const uint32_t src = cl_smem_u32(tile);
                const uint32_t dst = cl_smem_u32(recv + ((size_t)(b * R + q)) * NCOL);
                const uint32_t bar = cl_smem_u32(&s_ready[b]);
#pragma unroll
                for (int c2 = 0; c2 < CS; c2++) {
                    uint32_t d, m;
                    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(dst), "r"(c2));
                    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(m) : "r"(bar), "r"(c2));
                    bulk_push(d, src + c2 * SLICE, SLICE, m);
                }
                bulk_commit();
            }
            if (r + 1 < r1) {                             /* operands of the next row fly during the recombination */
                row_of(r + 1, pn, fn);
                const float2 *x = A.Xs + ((size_t)fn * R + q) * L, *cc = A.Cs + ((size_t)pn * R + q) * L;
#pragma unroll
                for (int h = 0; h < 16; h++) { xv[h] = __ldg((const float4 *)x + 32 * h + lane); cv[h] = __ldg((const float4 *)cc + 32 * h + lane); }
            }
        }
        if (i >= 1) {                                     /* row i-1 = (pq, fq), receive buffer pb */
            const int pb = b ^ 1;
            mbar_wait_cta(&s_ready[pb], (uint32_t)(((i - 1) >> 1) & 1));
            float2 z[32];
            const float2 *rb = recv + (size_t)(pb * R) * NCOL + tid;
#pragma unroll
            for (int qq = 0; qq < R; qq++) z[qq] = rb[qq * NCOL];
            twiddle32<true>(z, wc);                       /* z_q *= conj(W_M^(t'))^q */
            fft32<+1, true>(z);
#pragma unroll
            for (int a = 0; a < NA; a++) pw[a] = (z[a].x * z[a].x + z[a].y * z[a].y) * A.invM2;
            /* this warp has read its part of buffer pb: tell the producer warps of all CTAs, and arm the
             * buffer for the row after next (its copies start once FREE completes).  The arrival count
             * is computed from every lane's pw[0] so that the arrive cannot be scheduled before the
             * loads have returned (see k_roww).                                                      */
            const uint32_t one = 1u + (__ballot_sync(0xffffffffu, pw[0] < 0.f) != 0u);
            if (lane < CS) cl_mbar_arrive_remote_light(cl_map(&s_free[pb], (uint32_t)lane), one);
            if (tid == 0 && r + 1 < r1) mbar_expect_tx(&s_ready[pb], R * SLICE);
            if (MODE == 0) {
                if (pq != pcur) { flush(); pcur = pq; }
                /* branch-free: lags at or beyond n_eff get power -1 and never win; the accumulating path
                 * (cpxconv flagsum, src/sdrcmn.c:244-246) is a uniform branch around the whole loop     */
                if (A.P) {
                    float *Prow = A.P + ((size_t)pq * A.nfreq + fq) * A.n_eff;
#pragma unroll
                    for (int a = 0; a < NA; a++) {
                        const int t = tcol + L * a;
                        if (t < A.n_eff) { float xw = pw[a]; if (!A.first_step) xw += Prow[t]; Prow[t] = xw; pw[a] = xw; }
                        else pw[a] = -1.f;
                    }
                } else {
#pragma unroll
                    for (int a = NA / 2; a < NA; a++) pw[a] = (tcol + L * a < A.n_eff) ? pw[a] : -1.f;   /* n_eff > M/4 */
                }
                float lv = -1.f; int la = 0;
#pragma unroll
                for (int a = 0; a < NA; a++) { const bool up = pw[a] > lv; lv = up ? pw[a] : lv; la = up ? a : la; }
                if (lv > bv) { bv = lv; bflat = (unsigned int)fq * (unsigned int)A.n_eff + (unsigned int)(tcol + L * la); }
            }
        }
    }
    if (lane == 0) bulk_wait0();                          /* my copies have landed before my shared memory goes away */
    if (MODE == 0) {
        flush();
        cl_sync_all();
        return;
    }

    /* MODE 1: masked statistics around the known peak (src/sdracq.c:79-88) */
    if (from_hbm) {
        const float *Prow = A.P + ((size_t)cid * A.nfreq + fpk) * A.n_eff;
#pragma unroll
        for (int a = 0; a < NA; a++) { const int t = tcol + L * a; pw[a] = (t < A.n_eff) ? Prow[t] : 0.f; }
    }
    int exs = codei - 2 * A.nsampchip; if (exs < 0) exs += A.n_eff;
    int exe = codei + 2 * A.nsampchip; if (exe >= A.n_eff) exe -= A.n_eff;
    double sum = 0.0; int cnt = 0; float m2 = -1.f;
#pragma unroll
    for (int a = 0; a < NA; a++) {
        const int t = tcol + L * a;
        if (t < A.n_eff) {
            const bool inc = (exs <= exe) ? (t < exs || t > exe) : (t < exs && t > exe);
            if (inc) { sum += (double)pw[a]; cnt++; }
            if ((inc || t == 0) && pw[a] > m2) m2 = pw[a];            /* maxvd seeds with data[0] */
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    }
    if (lane == 0) { s_part[warp].sum = sum; s_part[warp].cnt = cnt; s_part[warp].max2 = m2; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < WPC; w++) { sum += s_part[w].sum; cnt += s_part[w].cnt; m2 = fmaxf(m2, s_part[w].max2); }
        s_cta.sum = sum; s_cta.cnt = cnt; s_cta.max2 = m2;
    }
    cl_sync_all();
    if (c == 0 && tid == 0) {
        for (uint32_t rr = 1; rr < CS; rr++) {
            const uint32_t ad = cl_map(&s_cta, rr);
            double os; int oc; float om;
            asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(os) : "r"(ad) : "memory");
            asm volatile("ld.shared::cluster.s32 %0, [%1];" : "=r"(oc) : "r"(ad + 8) : "memory");
            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(om) : "r"(ad + 12) : "memory");
            sum += os; cnt += oc; m2 = fmaxf(m2, om);
        }
        sdrb200_acq_result_t o;
        o.maxp = (double)maxp; o.maxp2 = (double)m2;
        o.meanp = sum / (double)cnt;                               /* meanvd, sdrcmn.c:487-497 */
        o.cn0 = 10 * log10(o.maxp / o.meanp / A.ctime);            /* sdracq.c:84 */
        o.peakr = o.maxp / o.maxp2;                                /* sdracq.c:88 */
        o.codei = codei; o.freqi = fpk; o.acqfreq = A.freq[fpk];
        o.acquired = (o.peakr > A.thresh) ? 1 : 0;
        o.nintg = A.step + 1;
        A.res[cid] = o;
        if (o.acquired) A.frozen[cid] = 1;
        A.key[cid] = 0ull;                                         /* next integration step */
    }
    cl_sync_all();
}

/* ------------------------------------------------------------- host side */
template <int L> static int launch_fwd(int src_kind, cudaStream_t st, int R, int nitems, const int8_t *src,
                                       size_t item_stride, int dtype, int nlive, const uint8_t *cidx,
                                       int cidx_stride, float scale, const float2 *tw, float2 *out)
{
    const size_t smem = sizeof(float2) * FFT_SMEM_ELEMS(L);
    dim3 grid(R, nitems);
    if (src_kind == 0) {
        if (dtype == 1) {
            SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fwd_if<L, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_fwd_if<L, 1><<<grid, L / 16, smem, st>>>(src, nlive, cidx, cidx_stride, scale, tw, R, out);
        } else {
            SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fwd_if<L, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_fwd_if<L, 2><<<grid, L / 16, smem, st>>>(src, nlive, cidx, cidx_stride, scale, tw, R, out);
        }
    } else {
        SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fwd<L, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_fwd<L, 1><<<grid, L / 16, smem, st>>>(src, item_stride, dtype, nlive, cidx, cidx_stride, scale, tw, R, out);
    }
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    return SDRB200_OK;
}

template <int L> static int launch_row(cudaStream_t st, const sdrb200_acq *a, float2 *S, int prn0, int nprn, int f0, int nfc,
                                       int first_step)
{
    const size_t smem = sizeof(float2) * FFT_SMEM_ELEMS(L);
    dim3 grid(nprn, a->R, nfc);
    const float invM2 = 1.0f / ((float)a->M * (float)a->M);
    SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_row<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_row<L><<<grid, L / 16, smem, st>>>(a->d_Xs, a->d_Cs, a->d_tw, a->R, a->cfg.nfreq, prn0, f0, nfc, S, a->d_P,
                                         a->n_eff, a->cfg.nsampchip, first_step, invM2, a->d_row, a->d_frozen);
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    return SDRB200_OK;
}

/* two-pass forward transform of the IF window for R = 16 G */
template <int G> static int launch_fwd_split_g(cudaStream_t st, const sdrb200_acq *a, const int8_t *win, float scale, int f0, int nfc)
{
    const size_t smem = sizeof(float2) * 16 * G * 32;
    dim3 grid(a->L / 32, nfc);
    const uint8_t *cidx = a->d_cidx + (size_t)f0 * a->M;
    float2 *xs = a->d_Xs + (size_t)f0 * a->M;
    if (a->cfg.dtype == 1) {
        SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fwd_split<G, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_fwd_split<G, 1><<<grid, 32 * G, smem, st>>>(win, a->m_ref, cidx, a->M, scale, a->d_tw, a->L, xs);
    } else {
        SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fwd_split<G, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_fwd_split<G, 2><<<grid, 32 * G, smem, st>>>(win, a->m_ref, cidx, a->M, scale, a->d_tw, a->L, xs);
    }
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    return SDRB200_OK;
}
template <int L> static int launch_fft_rows(cudaStream_t st, const sdrb200_acq *a, int f0, int nfc)
{
    const size_t smem = sizeof(float2) * FFT_SMEM_ELEMS(L);
    SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fft_rows<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_fft_rows<L><<<dim3(a->R, nfc), L / 16, smem, st>>>(a->d_Xs + (size_t)f0 * a->M, a->d_tw, a->R, a->warp32 ? 1 : 0);
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    return SDRB200_OK;
}
template <int R> static int launch_fwd_split_small_r(cudaStream_t st, const sdrb200_acq *a, const int8_t *win, float scale, int f0, int nfc)
{
    dim3 grid((a->L + 255) / 256, nfc);
    const uint8_t *cidx = a->d_cidx + (size_t)f0 * a->M;
    float2 *xs = a->d_Xs + (size_t)f0 * a->M;
    if (a->cfg.dtype == 1)
        k_fwd_split_small<R, 1><<<grid, 256, 0, st>>>(win, a->m_ref, cidx, a->M, scale, a->d_tw, a->L, xs);
    else
        k_fwd_split_small<R, 2><<<grid, 256, 0, st>>>(win, a->m_ref, cidx, a->M, scale, a->d_tw, a->L, xs);
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    return SDRB200_OK;
}
static int launch_fwd_split(cudaStream_t st, const sdrb200_acq *a, const int8_t *win, float scale, int f0, int nfc)
{
    switch (a->R) {
    case 2:   return launch_fwd_split_small_r<2>(st, a, win, scale, f0, nfc);
    case 4:   return launch_fwd_split_small_r<4>(st, a, win, scale, f0, nfc);
    case 8:   return launch_fwd_split_small_r<8>(st, a, win, scale, f0, nfc);
    case 16:  return launch_fwd_split_small_r<16>(st, a, win, scale, f0, nfc);
    case 3:   return launch_fwd_split_small_r<3>(st, a, win, scale, f0, nfc);
    case 6:   return launch_fwd_split_small_r<6>(st, a, win, scale, f0, nfc);
    case 12:  return launch_fwd_split_small_r<12>(st, a, win, scale, f0, nfc);
    case 5:   return launch_fwd_split_small_r<5>(st, a, win, scale, f0, nfc);
    case 10:  return launch_fwd_split_small_r<10>(st, a, win, scale, f0, nfc);
    case 32:  return launch_fwd_split_g<2>(st, a, win, scale, f0, nfc);   /* one thread per t with a 32-point register stage: 38 vs 34 us */
    case 64:  return launch_fwd_split_g<4>(st, a, win, scale, f0, nfc);
    case 128: return launch_fwd_split_g<8>(st, a, win, scale, f0, nfc);
    case 256: return launch_fwd_split_g<16>(st, a, win, scale, f0, nfc);
    case 48:  return launch_fwd_split_g<3>(st, a, win, scale, f0, nfc);   /* M = 3 * 2^k */
    case 96:  return launch_fwd_split_g<6>(st, a, win, scale, f0, nfc);
    case 192: return launch_fwd_split_g<12>(st, a, win, scale, f0, nfc);
    case 80:  return launch_fwd_split_g<5>(st, a, win, scale, f0, nfc);   /* M = 5 * 2^k */
    case 160: return launch_fwd_split_g<10>(st, a, win, scale, f0, nfc);
    }
    return SDRB200_EINVAL;
}

/* radix-R recombination (R = 16 G, G = 2, 4, 8, 16 and 5, 10); 0 if R has no such kernel */
static int comb_fft_supported(int R) { return R == 32 || R == 64 || R == 128 || R == 256 || R == 80 || R == 160 || R == 48 || R == 96 || R == 192; }
template <int G> static int launch_comb_fft_g(cudaStream_t st, const sdrb200_acq *a, const float2 *S, int prn0, int nprn, int f0, int nfc,
                                              int first_step)
{
    const size_t smem = sizeof(float2) * 16 * G * 32;
    dim3 grid(a->L / 32, nfc, nprn);
    const float invM2 = 1.0f / ((float)a->M * (float)a->M);
    SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_comb_fft<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_comb_fft<G><<<grid, 32 * G, smem, st>>>(S, a->d_tw, a->L, a->cfg.nfreq, prn0, f0, nfc, a->d_P, a->n_eff,
                                              first_step, invM2, a->d_key, a->d_frozen);
    sdrb200_count_launch(1);
    SDRB200_CUDA_TRY(cudaGetLastError());
    return SDRB200_OK;
}
static int launch_comb_fft(cudaStream_t st, const sdrb200_acq *a, const float2 *S, int prn0, int nprn, int f0, int nfc, int first_step)
{
    switch (a->R) {
    case 32:  return launch_comb_fft_g<2>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 64:  return launch_comb_fft_g<4>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 128: return launch_comb_fft_g<8>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 256: return launch_comb_fft_g<16>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 48:  return launch_comb_fft_g<3>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 96:  return launch_comb_fft_g<6>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 192: return launch_comb_fft_g<12>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 80:  return launch_comb_fft_g<5>(st, a, S, prn0, nprn, f0, nfc, first_step);
    case 160: return launch_comb_fft_g<10>(st, a, S, prn0, nprn, f0, nfc, first_step);
    }
    return SDRB200_EINVAL;
}

template <int L, int R, int MODE> static int launch_rowc(cudaStream_t st, const acq_rowc_args &A, int nfreq_grid, int nprn)
{
    const size_t smem = sizeof(float2) * FFT_SMEM_ELEMS(L);
    SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_rowc<L, R, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (R > 8) SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_rowc<L, R, MODE>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(R, nfreq_grid, nprn);
    cfg.blockDim = dim3(L / 16, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = R; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    SDRB200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_rowc<L, R, MODE>, A));
    sdrb200_count_launch(1);
    return SDRB200_OK;
}

/* clusters of more than 8 CTAs are a non-portable size: how many of them the device holds at once (0: none) */
template <int L, int R> static int rowc_big_clusters(void)
{
    const size_t smem = sizeof(float2) * FFT_SMEM_ELEMS(L);
    if (cudaFuncSetAttribute(k_rowc<L, R, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
        cudaFuncSetAttribute(k_rowc<L, R, 0>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); return 0; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(R, 16, 1); cfg.blockDim = dim3(L / 16, 1, 1); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = R; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, k_rowc<L, R, 0>, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

template <int MODE> static int roww_config(cudaLaunchConfig_t *cfg, cudaLaunchAttribute *at, cudaStream_t st, int ncl)
{
    const size_t smem = 2 * 4 * sizeof(float2) * WFFT_TILE;
    SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_roww<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    *cfg = cudaLaunchConfig_t{};
    cfg->gridDim = dim3(8 * ncl, 1, 1);
    cfg->blockDim = dim3(128, 1, 1);
    cfg->dynamicSmemBytes = smem;
    cfg->stream = st;
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 8; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg->attrs = at; cfg->numAttrs = 1;
    return SDRB200_OK;
}
static int roww_max_clusters(int *ncl)
{
    cudaLaunchConfig_t cfg; cudaLaunchAttribute at[1];
    int rc = roww_config<0>(&cfg, at, 0, 1024);
    if (rc) return rc;
    int n = 0;
    SDRB200_CUDA_TRY(cudaOccupancyMaxActiveClusters(&n, k_roww<0>, &cfg));
    *ncl = n;
    return SDRB200_OK;
}
/* mode 0: the search over ncl persistent clusters; mode 1: one cluster per PRN */
static int launch_roww(int mode, cudaStream_t st, const acq_rowc_args &A, int nprn, int ncl)
{
    cudaLaunchConfig_t cfg; cudaLaunchAttribute at[1];
    int rc = mode ? roww_config<1>(&cfg, at, st, nprn) : roww_config<0>(&cfg, at, st, ncl);
    if (rc) return rc;
    if (mode) SDRB200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_roww<1>, A, nprn));
    else SDRB200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_roww<0>, A, nprn));
    sdrb200_count_launch(1);
    return SDRB200_OK;
}

template <int MODE> static int rowd_config(cudaLaunchConfig_t *cfg, cudaLaunchAttribute *at, cudaStream_t st, int ncl)
{
    const size_t smem = ROWD_SMEM;
    SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_rowd<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    *cfg = cudaLaunchConfig_t{};
    cfg->gridDim = dim3(ROWD_CS * ncl, 1, 1);
    cfg->blockDim = dim3(32 * ROWD_WPC, 1, 1);
    cfg->dynamicSmemBytes = smem;
    cfg->stream = st;
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = ROWD_CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg->attrs = at; cfg->numAttrs = 1;
    return SDRB200_OK;
}
static int rowd_max_clusters(int *ncl)
{
    cudaLaunchConfig_t cfg; cudaLaunchAttribute at[1];
    int rc = rowd_config<0>(&cfg, at, 0, 1024);
    if (rc) return rc;
    int n = 0;
    SDRB200_CUDA_TRY(cudaOccupancyMaxActiveClusters(&n, k_rowd<0>, &cfg));
    *ncl = n;
    return SDRB200_OK;
}
static int launch_rowd(int mode, cudaStream_t st, const acq_rowc_args &A, int nprn, int ncl)
{
    cudaLaunchConfig_t cfg; cudaLaunchAttribute at[1];
    int rc = mode ? rowd_config<1>(&cfg, at, st, nprn) : rowd_config<0>(&cfg, at, st, ncl);
    if (rc) return rc;
    if (mode) SDRB200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_rowd<1>, A, nprn));
    else SDRB200_CUDA_TRY(cudaLaunchKernelEx(&cfg, k_rowd<0>, A, nprn));
    sdrb200_count_launch(1);
    return SDRB200_OK;
}

static int launch_rowc_any(int L, int R, int mode, cudaStream_t st, const acq_rowc_args &A, int nfreq_grid, int nprn)
{
#define RC(LL, RR) if (L == LL && R == RR) return mode ? launch_rowc<LL, RR, 1>(st, A, nfreq_grid, nprn) \
                                                       : launch_rowc<LL, RR, 0>(st, A, nfreq_grid, nprn);
    RC(512, 1) RC(1024, 1) RC(2048, 1) RC(4096, 1)
    RC(1024, 2) RC(2048, 2) RC(4096, 2) RC(8192, 2) RC(16384, 2)
    RC(1024, 4) RC(2048, 4) RC(4096, 4) RC(8192, 4) RC(16384, 4)
    RC(1024, 8) RC(2048, 8) RC(4096, 8) RC(8192, 8) RC(16384, 8)
    RC(4096, 16)
    RC(1024, 5) RC(2048, 5) RC(4096, 5) RC(8192, 5) RC(4096, 10)        /* M = 5 * 2^k */
    RC(1024, 3) RC(2048, 3) RC(4096, 3) RC(4096, 6) RC(8192, 6) RC(4096, 12)  /* M = 3 * 2^k */
#undef RC
    sdrb200_set_error("acquisition: no fused kernel for L=%d R=%d", L, R);
    return SDRB200_EINVAL;
}

#define ACQ_DISPATCH_L(Lval, CALL)                                             \
    switch (Lval) {                                                            \
    case 256:   { constexpr int LL = 256;   rc = CALL; } break;                \
    case 512:   { constexpr int LL = 512;   rc = CALL; } break;                \
    case 1024:  { constexpr int LL = 1024;  rc = CALL; } break;                \
    case 2048:  { constexpr int LL = 2048;  rc = CALL; } break;                \
    case 4096:  { constexpr int LL = 4096;  rc = CALL; } break;                \
    case 8192:  { constexpr int LL = 8192;  rc = CALL; } break;                \
    case 16384: { constexpr int LL = 16384; rc = CALL; } break;                \
    default: sdrb200_set_error("unsupported sub-FFT length %d", Lval); rc = SDRB200_EINVAL; }

extern "C" void sdrb200_acq_destroy(sdrb200_acq_t *a)
{
    if (!a) return;
    cudaFree(a->d_tw); cudaFree(a->d_cidx); cudaFree(a->d_Xs); cudaFree(a->d_Cs); cudaFree(a->d_S);
    cudaFree(a->d_P); cudaFree(a->d_row); cudaFree(a->d_key); cudaFree(a->d_res); cudaFree(a->d_hist); cudaFree(a->d_frozen); cudaFree(a->d_freq);
    cudaFree(a->d_hostwin);
    if (a->h_res) cudaFreeHost(a->h_res);
    if (a->h_pin) cudaFreeHost(a->h_pin);
    if (a->ev0[0]) for (int i = 0; i < ACQ_MAXEV; i++) { cudaEventDestroy(a->ev0[i]); cudaEventDestroy(a->ev1[i]); }
    for (int l = 0; l < ACQ_MAXLANES; l++) { if (a->lane_st[l]) cudaStreamDestroy(a->lane_st[l]); if (a->lane_ev[l]) cudaEventDestroy(a->lane_ev[l]); }
    if (a->ev_fork) cudaEventDestroy(a->ev_fork);
    if (a->st) cudaStreamDestroy(a->st);
    free(a->h_freq);
    free(a);
}

extern "C" void *sdrb200_acq_cuda_stream(sdrb200_acq_t *a) { return a ? (void *)a->st : NULL; }

/* M = R * L.  Up to R = 8 the R sub-transforms of a row are finished inside one
 * thread-block cluster (k_rowc); longer windows (20 ms coherent blocks) use the
 * scratch path k_row + k_comb.  SDRB200_LSUB overrides L for experiments.       */
static int acq_odd_shape(int m, int odd, int grid_per_prn, int *L, int *R, int *fused);
static void acq_choose(int M, int allow_warp, int *L, int *R, int *fused, int *warp32)
{
    if (M % 5 == 0 || M % 3 == 0) {                      /* 5 * 2^k or 3 * 2^k, see acq_choose_len */
        *warp32 = 0;
        if (!acq_odd_shape(M, (M % 5 == 0) ? 5 : 3, 0, L, R, fused)) { *L = 4096; *R = M / 4096; *fused = 0; }
        return;
    }
    int lsub = (M <= 4096) ? M : (M / 8 > 4096 ? M / 8 : 4096);
    /* two-pass path: 256-thread CTAs (L = 4096) as long as the recombination has a kernel for M / 4096 sub-transforms
     * (config 4, M = 2^20 = 256 x 4096: 706 -> 619 ms per search against 128 x 8192; 64 x 16384: 820 ms) */
    if (lsub > 16384) lsub = (M / 4096 <= 256) ? 4096 : (M / 8192 <= 256 ? 8192 : 16384);
    const char *env = getenv("SDRB200_LSUB");
    if (env) { int v = atoi(env); if ((v == 1024 || v == 2048 || v == 4096 || v == 8192 || v == 16384) && v <= M) lsub = v; }
    /* M = 65536 (GLONASS at 20 MS/s): 16 x 4096 in a cluster of 16 CTAs (non-portable size) keeps the
     * 256-thread CTAs and four of them per SM; SDRB200_ACQ_R8 selects 8 x 8192 (512 threads, two per SM) */
    if (M == 65536 && !env && !getenv("SDRB200_ACQ_R8")) lsub = 4096;
    *L = lsub; *R = M / lsub;
    *fused = (*R <= 16 && (*R <= 8 || *L == 4096) && !getenv("SDRB200_ACQ_SCRATCH")) ? 1 : 0;
    /* M = 32768 as 32 sub-transforms of 1024 points, one WARP each (k_roww / k_rowd): measured
     * alternatives to the 8 x 4096 cluster kernel, selected with SDRB200_ACQ_WARP=pull|push      */
    *warp32 = 0;
    const char *wsel = getenv("SDRB200_ACQ_WARP");
    if (allow_warp && M == 32768 && *fused && !env && wsel && (!strcmp(wsel, "pull") || !strcmp(wsel, "push"))) {
        *L = 1024; *R = 32; *warp32 = (wsel[1] == 'u' && wsel[2] == 's') ? 2 : 1;
    }
}

/* transform length for a window of m_ref = 2 n samples.  The correlation is linear for lags < n whenever the
 * transform holds the window, so the length need not be a power of two: M = 5 * 2^k or 3 * 2^k (0.625 / 0.75 of the
 * next power of two) when the window fits it and there are kernels for it - 80 or 160 sub-transforms of 4096 .. 16384 points on the
 * two-pass path (config 4: 2 n = 654 720 <= 655 360 = 160 x 4096 instead of 2^20, 37.5 % less arithmetic and traffic
 * in every pass), 5 x 1024 .. 5 x 4096 and 10 x 4096 on the cluster path (GLONASS at 20 MS/s: 2 n = 40 000 <= 40 960).
 * SDRB200_ACQ_POW2 keeps the power of two.                                                                        */
/* 1 when length m = odd * 2^k (odd = 3 or 5) has kernels: *L, *R, *fused as acq_choose returns them */
static int acq_odd_shape(int m, int odd, int grid_per_prn, int *L, int *R, int *fused)
{
    if (m % odd) return 0;
    const int base = m / odd;                            /* a power of two */
    if (!getenv("SDRB200_ACQ_SCRATCH")) {
        /* cluster kernel: odd x 1024 .. odd x 4096, then 2 odd x 4096 and (factor 3) 12 x 4096 */
        if (base == 1024 || base == 2048 || base == 4096) { *L = base; *R = odd; *fused = 1; return 1; }
        if (base == 8192) {
            *L = 4096; *R = 2 * odd; *fused = 1;
            if ((odd == 5 && getenv("SDRB200_ACQ_R5")) ) { *L = 8192; *R = 5; }
            return 1;
        }
        if (base == 16384 && odd == 3) { *L = 4096; *R = 12; *fused = 1; return 1; }
    }
    if (grid_per_prn) return 0;
    /* two-pass path: 16 x odd or 32 x odd sub-transforms (and 64 x 3) of 4096 .. 16384 points */
    const int rmax = (odd == 3) ? 192 : 160;
    int l = 4096;
    const char *e = getenv("SDRB200_LSUB");
    if (e && (atoi(e) == 4096 || atoi(e) == 8192 || atoi(e) == 16384)) l = atoi(e);
    else while (l < 16384 && m / l > rmax) l <<= 1;
    const int r = m / l;
    if (m % l == 0 && (r == 16 * odd || r == 32 * odd || (odd == 3 && r == 64 * odd))) { *L = l; *R = r; *fused = 0; return 1; }
    return 0;
}

static int acq_choose_len(int m_ref, int grid_per_prn)
{
    int M = 512;
    while (M < m_ref) M <<= 1;
    if (getenv("SDRB200_ACQ_POW2")) return M;
    int l, r, f;
    const int m5 = M / 8 * 5, m3 = M / 4 * 3;
    if (m5 >= m_ref && acq_odd_shape(m5, 5, grid_per_prn, &l, &r, &f)) return m5;
    if (m3 >= m_ref && acq_odd_shape(m3, 3, grid_per_prn, &l, &r, &f)) return m3;
    return M;
}

/* the decomposition a plan for nsamp x ncoh samples per window will use, without creating one (no device needed;
 * on a device that cannot hold clusters of 10 / 16 CTAs the plan itself takes 5 x 8192 / 8 x 8192 instead) */
extern "C" int sdrb200_acq_plan_shape(int nsamp, int ncoh, int grid_per_prn, int *M, int *L, int *R, int *fused)
{
    if (nsamp < 64 || !M || !L || !R || !fused) { sdrb200_set_error("acq_plan_shape: bad argument"); return SDRB200_EINVAL; }
    if (ncoh < 1) ncoh = 1;
    int w32 = 0;
    *M = acq_choose_len(2 * nsamp * ncoh, grid_per_prn);
    acq_choose(*M, 0, L, R, fused, &w32);
    return SDRB200_OK;
}

extern "C" int sdrb200_acq_create(sdrb200_acq_t **out, const sdrb200_acq_cfg_t *cfg)
{
    if (!out || !cfg || !cfg->freq || !cfg->codes || cfg->nsamp < 64 || cfg->nfreq < 1 || cfg->nprn < 1 ||
        cfg->clen < 1 || cfg->intg < 1 || (cfg->dtype != 1 && cfg->dtype != 2) || cfg->f_sf <= 0) {
        sdrb200_set_error("acq_create: bad argument"); return SDRB200_EINVAL;
    }
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    sdrb200_acq *a = (sdrb200_acq *)calloc(1, sizeof(*a));
    if (!a) return SDRB200_ENOMEM;
    a->cfg = *cfg;
    if (a->cfg.ncoh < 1) a->cfg.ncoh = 1;
    if (a->cfg.threshold <= 0) a->cfg.threshold = 3.0;           /* ACQTH */
    a->n_eff = cfg->nsamp * a->cfg.ncoh;
    a->m_ref = 2 * a->n_eff;                                      /* acq.nfft = 2*nsamp, sdrinit.c:625 */
    a->M = acq_choose_len(a->m_ref, cfg->grid_per_prn);
    acq_choose(a->M, !cfg->grid_per_prn, &a->L, &a->R, &a->fused, &a->warp32);   /* the warp kernels share one grid */
    /* a device (or partition) that cannot hold a cluster of 10 / 16 CTAs gets the portable split of the same length */
    if (a->fused && a->R == 10 && a->L == 4096 && rowc_big_clusters<4096, 10>() < 1) { a->L = 8192; a->R = 5; }
    if (a->fused && a->R == 12 && a->L == 4096 && rowc_big_clusters<4096, 12>() < 1) { a->L = 8192; a->R = 6; }
    if (a->fused && a->R == 16 && a->L == 4096 && rowc_big_clusters<4096, 16>() < 1) { a->L = 8192; a->R = 8; }
    const int nfreq = cfg->nfreq, nprn = cfg->nprn, M = a->M, n_eff = a->n_eff;
    /* one wipe-off grid per replica (GLONASS FDMA: every channel has its own carrier, src/sdrinit.c:633-635):
     * the rows of replica p read the spectra of bins p * nfreq .. ; fused cluster kernel only */
    a->nfx = cfg->grid_per_prn ? nprn * nfreq : nfreq;
    if (cfg->grid_per_prn && !a->fused) {
        sdrb200_set_error("acq_create: grid_per_prn needs the cluster kernel (window of at most 16 sub-transforms)");
        free(a); return SDRB200_EINVAL;
    }
    if (a->R > 256) { sdrb200_set_error("acq_create: window too long"); free(a); return SDRB200_EINVAL; }
    a->h_freq = (double *)malloc(sizeof(double) * a->nfx);
    memcpy(a->h_freq, cfg->freq, sizeof(double) * a->nfx);
    a->cfg.freq = a->h_freq; a->cfg.codes = NULL;
    /* scratch chunking.  Round 1 kept S within 64 MB "so that it stays L2-resident"; measured at config-4 size
     * the opposite holds - 8 rows per launch pair 618 ms, 16: 564, 36: 530, 141 (a whole PRN, 1.2 GB of scratch in
     * HBM): 493 ms - large grids matter more than where the scratch lives.  With the PRNs of a chunk adjacent in the grid of k_row (a spectrum slice comes from HBM once per chunk) two PRNs
     * per chunk give another 3-5 %: 3 GB budget, SDRB200_SCRATCH_MB overrides.                             */
    {
        size_t per_prn = (size_t)nfreq * M * sizeof(float2);
        size_t budget = (size_t)3072 << 20;
        { const char *env = getenv("SDRB200_SCRATCH_MB"); if (env && atoi(env) >= 8) budget = (size_t)atoi(env) << 20; }
        a->chunk = (int)(budget / per_prn); if (a->chunk < 1) a->chunk = 1; if (a->chunk > nprn) a->chunk = nprn;
        /* one PRN's bins alone exceed the budget (20 ms coherent blocks: 8.4 MB per row):
         * the chunk becomes a group of bins of one PRN                                  */
        a->fchunk = nfreq;
        if (per_prn > budget) {
            a->fchunk = (int)(budget / ((size_t)M * sizeof(float2)));
            if (a->fchunk < 1) a->fchunk = 1;
            if (a->fchunk > nfreq) a->fchunk = nfreq;
        }
        /* experiment switches, read once per plan: SDRB200_ACQ_COMB_NAIVE (direct-sum combine and
         * the walking forward kernel), SDRB200_FWD_WALK (walking forward kernel on the cluster path) */
        const int naive = getenv("SDRB200_ACQ_COMB_NAIVE") != NULL;
        a->comb_fft = (!a->fused && a->R > 1 && comb_fft_supported(a->R) && !naive) ? 1 : 0;
        a->split_fwd = (a->comb_fft || a->warp32 || (a->fused && (a->R == 2 || a->R == 4 || a->R == 8 || a->R == 16 || a->R == 5 || a->R == 10 || a->R == 3 || a->R == 6 || a->R == 12) &&
                                        !getenv("SDRB200_FWD_WALK") && !naive)) ? 1 : 0;
        const char *env = getenv("SDRB200_FCHUNK");
        if (env && atoi(env) >= 1 && atoi(env) <= nfreq) { a->fchunk = atoi(env); a->chunk = 1; }
        a->nlanes = 1;
        env = getenv("SDRB200_ACQ_LANES");
        if (env && atoi(env) >= 1 && atoi(env) <= ACQ_MAXLANES) a->nlanes = atoi(env);
        if (!a->comb_fft) a->nlanes = 1;
    }
    if (getenv("SDRB200_VERBOSE"))
        fprintf(stderr, "sdrb200: acquisition plan: window %d, transform %d = %d x %d, %s, %d PRN x %d bins per chunk, %d lane(s)\n",
                a->m_ref, a->M, a->R, a->L, a->fused ? "cluster kernel" : "two-pass path", a->chunk, a->fchunk, a->nlanes);
    int rc = SDRB200_OK;
    /* the warp-per-sub-transform kernels are persistent: as many clusters as the GPU holds at once */
    a->rowp_ncl = 0;
    if (a->warp32 && nprn > ROWP_MAXPRN) { sdrb200_set_error("acq_create: more than %d PRNs in one plan", ROWP_MAXPRN); free(a->h_freq); free(a); return SDRB200_EINVAL; }
    if (a->warp32) {
        int ncl = 0;
        a->push = (a->warp32 == 2);
        int rc2 = a->push ? rowd_max_clusters(&ncl) : roww_max_clusters(&ncl);
        if (rc2 || ncl < 1) { sdrb200_set_error("acq_create: k_roww cannot be resident"); free(a->h_freq); free(a); return SDRB200_ECUDA; }
        const int rows = nprn * nfreq;
        a->rowp_ncl = ncl < rows ? ncl : rows;
        const char *env = getenv("SDRB200_ROWP_NCL");
        if (env && atoi(env) >= 1 && atoi(env) <= ncl) a->rowp_ncl = atoi(env);
        if (getenv("SDRB200_VERBOSE")) fprintf(stderr, "sdrb200: k_roww/k_rowd: %d co-resident clusters, %d used for %d rows\n", ncl, a->rowp_ncl, rows);
    }
#define ACQ_ALLOC(ptr, bytes) if (rc == SDRB200_OK && cudaMalloc((void **)&(ptr), (bytes)) != cudaSuccess) { \
        cudaGetLastError(); sdrb200_set_error("acq_create: cudaMalloc(%zu) failed", (size_t)(bytes)); rc = SDRB200_ENOMEM; }
    if (cudaStreamCreateWithFlags(&a->st, cudaStreamNonBlocking) != cudaSuccess) rc = SDRB200_ECUDA;
    ACQ_ALLOC(a->d_tw, sizeof(float2) * M);
    ACQ_ALLOC(a->d_cidx, (size_t)a->nfx * M);                   /* row stride M: 16-byte aligned rows */
    ACQ_ALLOC(a->d_Xs, sizeof(float2) * (size_t)a->nfx * M);
    ACQ_ALLOC(a->d_Cs, sizeof(float2) * (size_t)nprn * M);
    if (!a->fused && a->R > 1) ACQ_ALLOC(a->d_S, sizeof(float2) * (size_t)a->nlanes * a->chunk * a->fchunk * M);
    if (rc == SDRB200_OK && a->nlanes > 1) {
        if (cudaEventCreateWithFlags(&a->ev_fork, cudaEventDisableTiming) != cudaSuccess) rc = SDRB200_ECUDA;
        for (int l = 0; l < a->nlanes && rc == SDRB200_OK; l++)
            if (cudaStreamCreateWithFlags(&a->lane_st[l], cudaStreamNonBlocking) != cudaSuccess ||
                cudaEventCreateWithFlags(&a->lane_ev[l], cudaEventDisableTiming) != cudaSuccess) rc = SDRB200_ECUDA;
    }
    /* the power matrix is kept in HBM only when it has to be: accumulation over
     * several integrations, the scratch path, or on request (sdrb200_acq_keep_power) */
    if (!a->fused || a->cfg.intg > 1) ACQ_ALLOC(a->d_P, sizeof(float) * (size_t)nprn * nfreq * n_eff);
    if (!a->fused) ACQ_ALLOC(a->d_row, sizeof(acq_rowstat_t) * (size_t)nprn * nfreq);
    ACQ_ALLOC(a->d_key, sizeof(unsigned long long) * nprn);
    ACQ_ALLOC(a->d_res, sizeof(sdrb200_acq_result_t) * nprn);
    ACQ_ALLOC(a->d_frozen, sizeof(int) * nprn);
    ACQ_ALLOC(a->d_freq, sizeof(double) * a->nfx);
    if (rc == SDRB200_OK && cudaMallocHost((void **)&a->h_res, sizeof(sdrb200_acq_result_t) * nprn) != cudaSuccess) rc = SDRB200_ENOMEM;
    if (rc) { sdrb200_acq_destroy(a); return rc; }

    /* twiddles exp(-2*pi*i*k/M), computed in double */
    {
        float2 *h = (float2 *)malloc(sizeof(float2) * M);
        for (int k = 0; k < M; k++) {
            double ang = -6.283185307179586476925286766559 * (double)k / (double)M;
            h[k] = make_float2((float)cos(ang), (float)sin(ang));
        }
        cudaMemcpyAsync(a->d_tw, h, sizeof(float2) * M, cudaMemcpyHostToDevice, a->st);
        cudaStreamSynchronize(a->st);
        free(h);
    }
    /* carrier table indices per Doppler bin: mixcarr(dataR,dtype,ti,m,freq[i],0.0,...)
     * (src/sdrcmn.c:761) has phi0 = 0 and a fixed grid, so the index sequence is a
     * constant of the configuration; it is generated with the exact NCO replay. */
    {
        uint8_t *h = (uint8_t *)calloc((size_t)a->nfx, (size_t)M);
        nco_table_t *tb = (nco_table_t *)malloc(sizeof(nco_table_t));
        const double ti = 1 / cfg->f_sf;
        for (int f = 0; f < a->nfx && rc == SDRB200_OK; f++) {
            double phi, ps;
            nco_carrier_init(a->h_freq[f], 0.0, ti, &phi, &ps);
            if (nco_walk(phi, ps, a->m_ref, 0.0, tb)) { sdrb200_set_error("acq_create: NCO table overflow"); rc = SDRB200_ERANGE; break; }
            int32_t hint = 0;
            uint8_t *row = h + (size_t)f * M;
            for (int t = 0; t < a->m_ref; t++) row[t] = (uint8_t)(((int)nco_value(tb, t, &hint)) & SDRB200_CMASK);
        }
        if (rc == SDRB200_OK) {
            cudaMemcpyAsync(a->d_cidx, h, (size_t)a->nfx * M, cudaMemcpyHostToDevice, a->st);
            cudaMemcpyAsync(a->d_freq, a->h_freq, sizeof(double) * a->nfx, cudaMemcpyHostToDevice, a->st);
            cudaStreamSynchronize(a->st);
        }
        free(h); free(tb);
        if (rc) { sdrb200_acq_destroy(a); return rc; }
    }
    /* code-spectrum bank: rescode(code,clen,0,0,ci,nsamp,rcode); cpxcpx; cpxfft
     * (src/sdrinit.c:650-653), at the padded length, resident in HBM            */
    {
        int8_t *h = (int8_t *)malloc((size_t)nprn * n_eff);
        const double ti = 1 / cfg->f_sf, ci = ti * cfg->crate;       /* sdr->ci, sdrinit.c:606 */
        if (rc == SDRB200_OK) {
            /* the reference's own sequential loop (one-time, host): coff += ci with
             * the wrap test, src/sdrcmn.c:614-617; ncoh > 1 simply keeps walking    */
            int *chip = (int *)malloc(sizeof(int) * n_eff);
            double coff = nco_code_init(0.0, 0, ci, cfg->clen);
            for (int t = 0; t < n_eff; t++, coff += ci) {
                if (coff >= cfg->clen) coff -= cfg->clen;
                chip[t] = (int)coff;
            }
            for (int p = 0; p < nprn; p++)
                for (int t = 0; t < n_eff; t++) h[(size_t)p * n_eff + t] = cfg->codes[(size_t)p * cfg->clen + chip[t]];
            free(chip);
            int8_t *d_rc = NULL;
            if (cudaMalloc((void **)&d_rc, (size_t)nprn * n_eff) != cudaSuccess) rc = SDRB200_ENOMEM;
            else {
                cudaMemcpyAsync(d_rc, h, (size_t)nprn * n_eff, cudaMemcpyHostToDevice, a->st);
                ACQ_DISPATCH_L(a->L, (launch_fwd<LL>(1, a->st, a->R, nprn, d_rc, (size_t)n_eff, 1, n_eff, NULL, 0, 1.0f, a->d_tw, a->d_Cs)));
                if (rc == SDRB200_OK && a->warp32) {
                    k_pair_order<<<nprn * a->R, 1024, 0, a->st>>>(a->d_Cs);
                    if (cudaGetLastError() != cudaSuccess) rc = SDRB200_ECUDA;
                }
                cudaStreamSynchronize(a->st);
                cudaFree(d_rc);
            }
        }
        free(h);
        if (rc) { sdrb200_acq_destroy(a); return rc; }
    }
    *out = a;
    return SDRB200_OK;
}

/* fold the events of the previous enqueue into the per-kernel totals */
static int acq_timing_collect(sdrb200_acq *a)
{
    if (a->nev == 0) return SDRB200_OK;
    SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
    for (int i = 0; i < a->nev; i++) {
        float ms = 0.f;
        SDRB200_CUDA_TRY(cudaEventElapsedTime(&ms, a->ev0[i], a->ev1[i]));
        a->tot_ms[a->evkind[i]] += ms; a->tot_cnt[a->evkind[i]]++;
    }
    a->nev = 0;
    return SDRB200_OK;
}

/* kernel classes: 0 k_fwd, 1 k_row, 2 k_comb, 3 k_final.  on=1 starts (and
 * clears) the totals, on=0 stops.                                             */
extern "C" int sdrb200_acq_set_timing(sdrb200_acq_t *a, int on)
{
    if (!a) return SDRB200_EINVAL;
    if (on && !a->ev0[0])
        for (int i = 0; i < ACQ_MAXEV; i++) {
            SDRB200_CUDA_TRY(cudaEventCreate(&a->ev0[i]));
            SDRB200_CUDA_TRY(cudaEventCreate(&a->ev1[i]));
        }
    if (on) { memset(a->tot_ms, 0, sizeof(a->tot_ms)); memset(a->tot_cnt, 0, sizeof(a->tot_cnt)); a->nev = 0; }
    a->timing = on ? 1 : 0;
    return SDRB200_OK;
}
extern "C" int sdrb200_acq_kernel_ms(sdrb200_acq_t *a, double *ms4, long long *cnt4)
{
    if (!a || !ms4 || !cnt4) return SDRB200_EINVAL;
    int rc = acq_timing_collect(a);
    if (rc) return rc;
    for (int i = 0; i < 4; i++) { ms4[i] = a->tot_ms[i]; cnt4[i] = a->tot_cnt[i]; }
    return SDRB200_OK;
}

/* all integration steps of sdracquisition()'s loop (src/sdracq.c:29-43) for every
 * PRN, enqueued on the plan's stream                                            */
static int acq_steps(sdrb200_acq *a, const int8_t *base, uint64_t capacity, uint64_t buffloc0)
{
    const sdrb200_acq_cfg_t *c = &a->cfg;
    int rc = SDRB200_OK;
    const float scale = (float)(SDRB200_CSCALE / a->m_ref);            /* sdrcmn.c:764 */
    if (a->timing) { int rc2 = acq_timing_collect(a); if (rc2) return rc2; }
    SDRB200_CUDA_TRY(cudaMemsetAsync(a->d_frozen, 0, sizeof(int) * c->nprn, a->st));
    SDRB200_CUDA_TRY(cudaMemsetAsync(a->d_res, 0, sizeof(sdrb200_acq_result_t) * c->nprn, a->st));
    SDRB200_CUDA_TRY(cudaMemsetAsync(a->d_key, 0, sizeof(unsigned long long) * c->nprn, a->st));
    for (int step = 0; step < c->intg; step++) {
        const uint64_t pos = (buffloc0 + (uint64_t)step * a->n_eff) % capacity;
        const int8_t *win = base + pos * c->dtype;
        ACQ_T0(a);
        if (a->split_fwd) {
            /* (tried: the two passes over groups of bins that fit the L2, so that the rows pass overwrites
             * the first pass's output on chip - GLONASS forward 5.4 -> 6.7 ms per search, the smaller grids
             * cost more than the write traffic they save)                                               */
            rc = launch_fwd_split(a->st, a, win, scale, 0, a->nfx);
            if (rc) return rc;
            ACQ_DISPATCH_L(a->L, (launch_fft_rows<LL>(a->st, a, 0, a->nfx)));
        } else {
            ACQ_DISPATCH_L(a->L, (launch_fwd<LL>(0, a->st, a->R, a->nfx, win, 0, c->dtype, a->m_ref, a->d_cidx,
                                                 a->M, scale, a->d_tw, a->d_Xs)));
        }
        if (rc) return rc;
        ACQ_T1(a, 0);
        if (a->fused) {
            acq_rowc_args A;
            A.Xs = a->d_Xs; A.Cs = a->d_Cs; A.tw = a->d_tw; A.P = a->d_P; A.key = a->d_key;
            A.frozen = a->d_frozen; A.res = a->d_res; A.freq = a->d_freq;
            A.ctime = c->ctime; A.thresh = c->threshold; A.nfreq = c->nfreq; A.n_eff = a->n_eff;
            A.xs = c->grid_per_prn ? c->nfreq : 0;
            A.nsampchip = c->nsampchip; A.first_step = (step == 0); A.step = step;
            A.invM2 = 1.0f / ((float)a->M * (float)a->M);
            ACQ_T0(a);
            if (a->warp32) rc = a->push ? launch_rowd(0, a->st, A, c->nprn, a->rowp_ncl) : launch_roww(0, a->st, A, c->nprn, a->rowp_ncl);
            else rc = launch_rowc_any(a->L, a->R, 0, a->st, A, c->nfreq, c->nprn);
            if (rc) return rc;
            ACQ_T1(a, 1);
            ACQ_T0(a);
            if (a->warp32) rc = a->push ? launch_rowd(1, a->st, A, c->nprn, 0) : launch_roww(1, a->st, A, c->nprn, 0);
            else rc = launch_rowc_any(a->L, a->R, 1, a->st, A, 1, c->nprn);
            if (rc) return rc;
            ACQ_T1(a, 3);
            if (a->d_hist)
                SDRB200_CUDA_TRY(cudaMemcpyAsync(a->d_hist + (size_t)step * c->nprn, a->d_res, sizeof(sdrb200_acq_result_t) * c->nprn,
                                                 cudaMemcpyDeviceToDevice, a->st));
            continue;
        }
        /* scratch path: sub-transforms of a chunk of rows go through d_S (sized to stay in L2),
         * then the radix-R recombination; SDRB200_ACQ_COMB_NAIVE selects the direct-sum
         * combine with per-row statistics (kept as a cross-check)                        */
        const int use_fft = a->comb_fft;
        /* nlanes > 1: consecutive chunks go to different streams, each with its own slice of the scratch, so the
         * sub-transform pass of one chunk fills the SMs the combine pass of another leaves idle in its tail
         * (fork after the forward pass, join before the peak statistics)                                   */
        const int lanes = (use_fft && !a->timing) ? a->nlanes : 1;
        const size_t lane_elems = (size_t)a->chunk * a->fchunk * a->M;
        int ci = 0;
        if (lanes > 1) {
            SDRB200_CUDA_TRY(cudaEventRecord(a->ev_fork, a->st));
            for (int l = 0; l < lanes; l++) SDRB200_CUDA_TRY(cudaStreamWaitEvent(a->lane_st[l], a->ev_fork, 0));
        }
        for (int p0 = 0; p0 < c->nprn; p0 += a->chunk) {
            const int np = (c->nprn - p0 < a->chunk) ? c->nprn - p0 : a->chunk;
            for (int f0 = 0; f0 < c->nfreq; f0 += a->fchunk, ci++) {
                const int nfc = (c->nfreq - f0 < a->fchunk) ? c->nfreq - f0 : a->fchunk;
                cudaStream_t cst = (lanes > 1) ? a->lane_st[ci % lanes] : a->st;
                float2 *S = a->d_S + ((lanes > 1) ? (size_t)(ci % lanes) * lane_elems : 0);
                ACQ_T0(a);
                ACQ_DISPATCH_L(a->L, (launch_row<LL>(cst, a, S, p0, np, f0, nfc, step == 0)));
                if (rc) return rc;
                ACQ_T1(a, 1);
                if (a->R > 1) {
                    ACQ_T0(a);
                    if (use_fft) {
                        rc = launch_comb_fft(cst, a, S, p0, np, f0, nfc, step == 0);
                        if (rc) return rc;
                    } else {
                        dim3 grid(nfc, np);
                        const float invM2 = 1.0f / ((float)a->M * (float)a->M);
                        k_comb<<<grid, 1024, 0, a->st>>>(a->d_S, a->d_tw, a->R, a->L, c->nfreq, p0, f0, nfc, a->d_P, a->n_eff,
                                                         c->nsampchip, step == 0, invM2, a->d_row, a->d_frozen);
                        sdrb200_count_launch(1);
                        SDRB200_CUDA_TRY(cudaGetLastError());
                    }
                    ACQ_T1(a, 2);
                }
            }
        }
        if (lanes > 1)
            for (int l = 0; l < lanes; l++) {
                SDRB200_CUDA_TRY(cudaEventRecord(a->lane_ev[l], a->lane_st[l]));
                SDRB200_CUDA_TRY(cudaStreamWaitEvent(a->st, a->lane_ev[l], 0));
            }
        ACQ_T0(a);
        if (use_fft) {
            k_peak_hbm<<<c->nprn, 1024, 0, a->st>>>(a->d_P, a->d_key, c->nfreq, a->n_eff, c->nsampchip, a->d_freq,
                                                    c->ctime, c->threshold, step, a->d_res, a->d_frozen);
        } else {
            k_final<<<c->nprn, 128, 0, a->st>>>(a->d_row, c->nfreq, a->d_freq, c->ctime, c->threshold, step,
                                               a->d_res, a->d_frozen);
        }
        sdrb200_count_launch(1);
        SDRB200_CUDA_TRY(cudaGetLastError());
        ACQ_T1(a, 3);
        if (a->d_hist)
            SDRB200_CUDA_TRY(cudaMemcpyAsync(a->d_hist + (size_t)step * c->nprn, a->d_res, sizeof(sdrb200_acq_result_t) * c->nprn,
                                             cudaMemcpyDeviceToDevice, a->st));
    }
    return SDRB200_OK;
}

extern "C" int sdrb200_acq_enqueue(sdrb200_acq_t *a, const sdrb200_stream_t *s, uint64_t buffloc0)
{
    if (!a || !s) { sdrb200_set_error("acq_run: bad argument"); return SDRB200_EINVAL; }
    if (s->dtype != a->cfg.dtype) { sdrb200_set_error("acq_run: stream dtype mismatch"); return SDRB200_EINVAL; }
    const uint64_t need = (uint64_t)(a->cfg.intg + 1) * a->n_eff;
    if (buffloc0 + need > s->count || s->count - buffloc0 > s->capacity) {
        sdrb200_set_error("acq_run: window [%llu,+%llu) not held by the stream (count %llu)",
                          (unsigned long long)buffloc0, (unsigned long long)need, (unsigned long long)s->count);
        return SDRB200_ERANGE;
    }
    if ((size_t)a->m_ref * a->cfg.dtype + 32 > SDRB200_STREAM_GUARD) { sdrb200_set_error("acq_run: window exceeds the ring guard"); return SDRB200_EINVAL; }
    return acq_steps(a, s->d_buf, s->capacity, buffloc0);
}

extern "C" int sdrb200_acq_fetch(sdrb200_acq_t *a, sdrb200_acq_result_t *results)
{
    if (!a || !results) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(a->h_res, a->d_res, sizeof(sdrb200_acq_result_t) * a->cfg.nprn,
                                     cudaMemcpyDeviceToHost, a->st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
    memcpy(results, a->h_res, sizeof(sdrb200_acq_result_t) * a->cfg.nprn);
    return SDRB200_OK;
}

/* ---- multi-GPU plumbing: results stay on the device until ONE collective gathers them ---- */
extern "C" const void *sdrb200_acq_results_device(sdrb200_acq_t *a) { return a ? (const void *)a->d_res : NULL; }

extern "C" int sdrb200_acq_copy_results_device(sdrb200_acq_t *a, void *d_dst)
{
    if (!a || !d_dst) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(d_dst, a->d_res, sizeof(sdrb200_acq_result_t) * a->cfg.nprn, cudaMemcpyDeviceToDevice, a->st));
    return SDRB200_OK;
}

extern "C" int sdrb200_acq_sync(sdrb200_acq_t *a)
{
    if (!a) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
    return SDRB200_OK;
}

/* per-step history: a Doppler SLAB of a search cannot take the reference's early exit by itself
 * (src/sdracq.c:29-43 tests the peak of the WHOLE grid after every integration), so a slab plan
 * runs every step (threshold set out of reach by the caller) and keeps the tuple of each step;
 * the owner of the search replays the early exit over the slabs' histories (shard.merge_doppler_slabs) */
extern "C" int sdrb200_acq_set_history(sdrb200_acq_t *a, int on)
{
    if (!a) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
    if (on && !a->d_hist) {
        const size_t bytes = sizeof(sdrb200_acq_result_t) * (size_t)a->cfg.intg * a->cfg.nprn;
        if (cudaMalloc((void **)&a->d_hist, bytes) != cudaSuccess) { cudaGetLastError(); sdrb200_set_error("acq_set_history: cudaMalloc failed"); return SDRB200_ENOMEM; }
        SDRB200_CUDA_TRY(cudaMemset(a->d_hist, 0, bytes));
    } else if (!on && a->d_hist) { cudaFree(a->d_hist); a->d_hist = NULL; }
    return SDRB200_OK;
}
extern "C" const void *sdrb200_acq_history_device(sdrb200_acq_t *a) { return a ? (const void *)a->d_hist : NULL; }
extern "C" int sdrb200_acq_fetch_history(sdrb200_acq_t *a, sdrb200_acq_result_t *hist)
{
    if (!a || !hist) return SDRB200_EINVAL;
    if (!a->d_hist) { sdrb200_set_error("acq_fetch_history: history is off (sdrb200_acq_set_history)"); return SDRB200_EINVAL; }
    SDRB200_CUDA_TRY(cudaMemcpyAsync(hist, a->d_hist, sizeof(sdrb200_acq_result_t) * (size_t)a->cfg.intg * a->cfg.nprn,
                                     cudaMemcpyDeviceToHost, a->st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
    return SDRB200_OK;
}

extern "C" int sdrb200_acq_run(sdrb200_acq_t *a, const sdrb200_stream_t *s, uint64_t buffloc0,
                               sdrb200_acq_result_t *results)
{
    int rc = sdrb200_acq_enqueue(a, s, buffloc0);
    if (rc) return rc;
    return sdrb200_acq_fetch(a, results);
}

/* the first half of sdrb200_acq_run_host: H2D copy and all integration steps enqueued on the plan's
 * stream, no wait - a caller with several plans (the reference runs one thread per channel) keeps the
 * next search's copy and launches behind the current one and collects with sdrb200_acq_fetch()     */
extern "C" int sdrb200_acq_enqueue_host(sdrb200_acq_t *a, const int8_t *host_if, uint64_t nsamples)
{
    if (!a || !host_if) { sdrb200_set_error("acq_enqueue_host: bad argument"); return SDRB200_EINVAL; }
    const uint64_t need = (uint64_t)(a->cfg.intg + 1) * a->n_eff;
    if (nsamples < need) { sdrb200_set_error("acq_enqueue_host: need %llu samples", (unsigned long long)need); return SDRB200_ERANGE; }
    const size_t bytes = (size_t)need * a->cfg.dtype;
    if (bytes + 512 > a->hostwin_cap) {
        SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
        cudaFree(a->d_hostwin); a->d_hostwin = NULL; a->hostwin_cap = 0;
        if (cudaMalloc((void **)&a->d_hostwin, bytes + 512) != cudaSuccess) { cudaGetLastError(); return SDRB200_ENOMEM; }
        a->hostwin_cap = bytes + 512;
    }
    SDRB200_CUDA_TRY(cudaMemcpyAsync(a->d_hostwin + 256, host_if, bytes, cudaMemcpyHostToDevice, a->st));
    return acq_steps(a, a->d_hostwin + 256, ~(uint64_t)0, 0);
}

extern "C" int sdrb200_acq_run_host(sdrb200_acq_t *a, const int8_t *host_if, uint64_t nsamples,
                                    sdrb200_acq_result_t *results)
{
    if (!a || !host_if || !results) { sdrb200_set_error("acq_run_host: bad argument"); return SDRB200_EINVAL; }
    const uint64_t need = (uint64_t)(a->cfg.intg + 1) * a->n_eff;
    if (nsamples < need) { sdrb200_set_error("acq_run_host: need %llu samples", (unsigned long long)need); return SDRB200_ERANGE; }
    const size_t bytes = (size_t)need * a->cfg.dtype;
    if (bytes + 512 > a->hostwin_cap) {
        SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
        cudaFree(a->d_hostwin); a->d_hostwin = NULL; a->hostwin_cap = 0;
        if (cudaMalloc((void **)&a->d_hostwin, bytes + 512) != cudaSuccess) { cudaGetLastError(); return SDRB200_ENOMEM; }
        a->hostwin_cap = bytes + 512;
    }
    SDRB200_CUDA_TRY(cudaMemcpyAsync(a->d_hostwin + 256, host_if, bytes, cudaMemcpyHostToDevice, a->st));
    int rc = acq_steps(a, a->d_hostwin + 256, ~(uint64_t)0, 0);
    if (rc) return rc;
    return sdrb200_acq_fetch(a, results);
}

extern "C" int sdrb200_acq_keep_power(sdrb200_acq_t *a, int on)
{
    if (!a) return SDRB200_EINVAL;
    if (on && !a->d_P) {
        SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
        const size_t bytes = sizeof(float) * (size_t)a->cfg.nprn * a->cfg.nfreq * a->n_eff;
        if (cudaMalloc((void **)&a->d_P, bytes) != cudaSuccess) {
            cudaGetLastError(); sdrb200_set_error("acq_keep_power: cudaMalloc(%zu) failed", bytes); return SDRB200_ENOMEM;
        }
    } else if (!on && a->d_P && a->fused && a->cfg.intg == 1) {
        SDRB200_CUDA_TRY(cudaStreamSynchronize(a->st));
        cudaFree(a->d_P); a->d_P = NULL;
    }
    return SDRB200_OK;
}

extern "C" int sdrb200_acq_power(sdrb200_acq_t *a, int prn_index, double *power)
{
    if (!a || !power || prn_index < 0 || prn_index >= a->cfg.nprn) return SDRB200_EINVAL;
    if (!a->d_P) { sdrb200_set_error("acq_power: the power matrix is not kept (call sdrb200_acq_keep_power first)"); return SDRB200_EINVAL; }
    const size_t n = (size_t)a->cfg.nfreq * a->n_eff;
    float *h = (float *)malloc(sizeof(float) * n);
    if (!h) return SDRB200_ENOMEM;
    cudaError_t e = cudaMemcpyAsync(h, a->d_P + (size_t)prn_index * n, sizeof(float) * n, cudaMemcpyDeviceToHost, a->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(a->st);
    if (e != cudaSuccess) { free(h); sdrb200_set_error("acq_power: %s", cudaGetErrorString(e)); return SDRB200_ECUDA; }
    for (size_t i = 0; i < n; i++) power[i] = (double)h[i];
    free(h);
    return SDRB200_OK;
}
