/* trk_device.cuh - one code period of the tracking correlator, executed by a
 * whole CTA.  Device-side counterpart of correlator() (src/sdrcmn.c:687-722):
 *   mixcarr  (src/sdrcmn.c:633-669)  carrier wipe-off with the 32-entry table
 *   rescode  (src/sdrcmn.c:608-621)  nearest-chip replica, n + 2*smax samples
 *   dot_23 / dot_22 (src/sdrcmn.c:310-354)  1+2*ns tap sums
 * All sums are exact integers (the reference accumulates the same integers in
 * double), the NCO index sequences are reproduced exactly via nco_exact.h, so
 * II/QQ are bit-identical to the reference and remc/remp are identical doubles.
 */
#ifndef SDRB200_TRK_DEVICE_CUH
#define SDRB200_TRK_DEVICE_CUH

#include <cuda_runtime.h>
#include <stdint.h>
#include "nco_exact.h"
#include "sdrb200_internal.h"

/* carrier table: cost[i]=floor(cos(2*pi*i/32)*32+0.5), sint likewise
 * (src/sdrcmn.c:641-646); trk.cu re-derives the table with libm at start-up
 * and refuses to run if it differs                 */
static __constant__ int8_t c_cost[32] = {32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6,0,6,12,18,23,27,30,31};
static __constant__ int8_t c_sint[32] = {0,6,12,18,23,27,30,31,32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6};
/* the same values for the host-side self check in trk.cu */
#define SDRB200_COST_INIT {32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6,0,6,12,18,23,27,30,31}
#define SDRB200_SINT_INIT {0,6,12,18,23,27,30,31,32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6}

struct trk_smem_t {
    nco_table_t carr;                       /* carrier NCO segments             */
    nco_table_t code;                       /* code NCO segments                */
    unsigned long long sumI[SDRB200_MAXTAPS];
    unsigned long long sumQ[SDRB200_MAXTAPS];
    double   remc, remp;
    int32_t  err;
    int8_t   codetab[1024];                 /* +-1 chips                        */
};

/* binary search: last segment with k0 <= k */
__device__ __forceinline__ int trk_seg_find(const nco_table_t *t, int k)
{
    int lo = 0, hi = t->nseg - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (t->seg[mid].k0 <= k) lo = mid; else hi = mid - 1;
    }
    return lo;
}

/* value at index k given a segment cursor that is advanced monotonically */
__device__ __forceinline__ double trk_seg_value(const nco_table_t *t, int k, int &s)
{
    while (s + 1 < t->nseg && t->seg[s + 1].k0 <= k) s++;
    return __dadd_rn(t->seg[s].x0, __dmul_rn((double)(k - t->seg[s].k0), t->seg[s].d));
}

/* Phase 1 (two threads): build both NCO tables and the remainders.
 * Must be followed by __syncthreads().                                         */
__device__ __forceinline__ void trk_build_tables(trk_smem_t *sm, int dtype, double ti, int n,
                                                 double freq, double phi0, double crate,
                                                 double coff, int smax, int coden)
{
    if (threadIdx.x == 0) {
        double phi, ps;
        nco_carrier_init(freq, phi0, ti, &phi, &ps);
        if (nco_walk(phi, ps, n, 0.0, &sm->carr)) sm->err = 1;
        sm->remp = nco_carrier_rem(sm->carr.x_end);
    }
    if (threadIdx.x == 32) {
        double ci = __dmul_rn(ti, crate);                   /* ti*crate, sdrcmn.c:709 */
        double c0 = nco_code_init(coff, smax, ci, coden);
        if (nco_walk(c0, ci, n + 2 * smax, (double)coden, &sm->code)) sm->err = 1;
        sm->remc = nco_code_rem(sm->code.x_end, smax, ci);
    }
    if (threadIdx.x >= 64 && threadIdx.x < 64 + SDRB200_MAXTAPS) {
        sm->sumI[threadIdx.x - 64] = 0ULL;
        sm->sumQ[threadIdx.x - 64] = 0ULL;
    }
    (void)dtype;
}

/* Phase 2: resampled replica rc[p], p in [0, n+2*smax) as +-1 bytes.
 * Must be followed by __syncthreads().                                         */
__device__ __forceinline__ void trk_fill_replica(const trk_smem_t *sm, int8_t *rc, int n, int smax)
{
    const int N = n + 2 * smax;
    for (int base = threadIdx.x * 16; base < N; base += blockDim.x * 16) {
        int s = trk_seg_find(&sm->code, base);
        int end = min(base + 16, N);
        for (int p = base; p < end; p++) {
            double v = trk_seg_value(&sm->code, p, s);
            rc[p] = sm->codetab[__double2int_rz(v)];
        }
    }
}

/* Phase 3: wipe-off + tap sums.  data points at sample 0 of the window in
 * global memory (any alignment; the 16 bytes before and after the window must
 * be readable, which the IF stream's layout guarantees).
 * taps[i] (i < ns) are the correlator offsets in samples (trk.corrp).
 * Output order as the reference: {P, E1, L1, E2, L2, ...} (sdrcmn.c:682-683),
 * E_i uses rc shifted by -taps[i], L_i by +taps[i] (sdrcmn.c:712-715).
 * Ends with the block-level sums complete in sm->sumI/Q after a __syncthreads().*/
template <int DTYPE>
__device__ __forceinline__ void trk_accumulate(trk_smem_t *sm, const int8_t *rc,
                                               const int8_t *__restrict__ data, int n,
                                               const int *taps, int ns, int smax)
{
    constexpr int SPV = 16 / DTYPE;                     /* samples per 16-byte vector */
    const uintptr_t addr = (uintptr_t)data;
    const int a = (int)(addr & 15);                     /* bytes of misalignment      */
    const int4 *vbase = (const int4 *)(addr - a);
    const int nvec = (n * DTYPE + a + 15) >> 4;
    const int ntap = 1 + 2 * ns;
    const int lane = threadIdx.x & 31;

    for (int v0 = 0; v0 < nvec; v0 += blockDim.x) {     /* uniform trip count         */
        const int v = v0 + threadIdx.x;
        int I[SPV], Q[SPV];
        int kfirst = 0;
        bool active = v < nvec;
        if (active) {
            int4 w = __ldg(vbase + v);
            const int8_t *b = (const int8_t *)&w;
            /* first sample index covered by this vector (may be < 0 or >= n) */
            kfirst = (v * 16 - a) / DTYPE;              /* a is even for IQ streams   */
            if (v * 16 - a < 0) kfirst = -((a - v * 16) / DTYPE);
            int s = trk_seg_find(&sm->carr, max(kfirst, 0));
#pragma unroll
            for (int j = 0; j < SPV; j++) {
                int k = kfirst + j;
                I[j] = 0; Q[j] = 0;
                if (k >= 0 && k < n) {
                    double phi = trk_seg_value(&sm->carr, k, s);
                    int idx = __double2int_rz(phi) & SDRB200_CMASK;
                    int c = c_cost[idx], sn = c_sint[idx];
                    if (DTYPE == 1) {
                        int x = b[j];
                        I[j] = c * x; Q[j] = sn * x;                   /* sdrcmn.c:660-661 */
                    } else {
                        int x0 = b[2 * j], x1 = b[2 * j + 1];
                        I[j] = c * x0 - sn * x1;                       /* sdrcmn.c:654-655 */
                        Q[j] = sn * x0 + c * x1;
                    }
                }
            }
        }
        for (int t = 0; t < ntap; t++) {
            int off = 0;
            if (t > 0) { int i = (t - 1) >> 1; off = ((t - 1) & 1) ? taps[i] : -taps[i]; }
            int accI = 0, accQ = 0;
            if (active) {
                const int8_t *r = rc + smax + off + kfirst;
#pragma unroll
                for (int j = 0; j < SPV; j++) {
                    int k = kfirst + j;
                    int c = (k >= 0 && k < n) ? (int)r[j] : 0;
                    accI += c * I[j];
                    accQ += c * Q[j];
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                accI += __shfl_xor_sync(0xffffffffu, accI, o);
                accQ += __shfl_xor_sync(0xffffffffu, accQ, o);
            }
            if (lane == 0) {
                atomicAdd(&sm->sumI[t], (unsigned long long)(long long)accI);
                atomicAdd(&sm->sumQ[t], (unsigned long long)(long long)accQ);
            }
        }
    }
    __syncthreads();
}

#endif
