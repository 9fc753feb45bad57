/* trk_device.cuh - one code period of the tracking correlator, executed by one
 * CTA.  Device-side counterpart of correlator() (src/sdrcmn.c:687-722):
 *   mixcarr  (src/sdrcmn.c:633-669)  carrier wipe-off with the 32-entry table
 *   rescode  (src/sdrcmn.c:608-621)  nearest-chip replica, n + 2*smax samples
 *   dot_23 / dot_22 (src/sdrcmn.c:310-354)  1+2*ns tap sums
 *
 * Formulation (all integer, exact):
 *   m[k]   = carrier-mixed sample k (I and Q), P[k] = sum_{i<k} m[i]  (prefix sums)
 *   r[p]   = replica value at window position p, piecewise constant (one run per
 *            chip, ~16 samples at 16.368 MS/s); B_q = first position of run q
 *   tap(off) = sum_k m[k] r[k+off]
 *            = r_last * P[n] + sum_q (r_{q-1} - r_q) * P[clip(B_q - off, 0, n)]
 * so a tap costs one prefix lookup per chip SIGN CHANGE (~512 per period)
 * instead of one multiply-add per sample (16368), and all 1+2*ns taps share the
 * prefix arrays.  Sums are kept modulo 2^32; the true values (|sum| < 2^31) are
 * recovered exactly.  The reference accumulates the same integers in doubles,
 * so II/QQ are bit-identical; the NCO index sequences come from nco_exact.h in
 * integer form (nco_iseg_t), so no FP64 arithmetic is done per sample.
 *
 * CTA layout: 1024 threads = two NCO warps (0: carrier NCO lane + epilogue,
 * 1: code NCO lane) + 960 workers.  Worker u owns TRK_SPAN = 20 bytes
 * [20u, 20u+20) of the 4-byte-aligned span that covers the window (five
 * coalesced 32-bit loads).  Workers take a branch-free path when their span lies
 * inside one NCO segment (almost always), so the per-sample work is independent
 * integer arithmetic with full instruction-level parallelism.
 */
#ifndef SDRB200_TRK_DEVICE_CUH
#define SDRB200_TRK_DEVICE_CUH

#include <cuda_runtime.h>
#include <stdint.h>
#include "nco_exact.h"
#include "sdrb200_internal.h"

#define TRK_WORKERS   960                    /* warps 2..31                        */
#define TRK_THREADS   1024
#define TRK_NCO_A     0                      /* carrier NCO lane; warp 0 = epilogue warp */
#define TRK_NCO_B     32                     /* code NCO lane (warp 1)             */
#define TRK_SPAN      20                     /* bytes of IF owned by one worker per round */
#define TRK_MAXSEG    96                     /* NCO segments held on the device    */
/* rows of the prefix store: one per 20-byte span of the (4-byte aligned) window */
#define TRK_ROWS(nbytes) (((nbytes) + 3 + TRK_SPAN - 1) / TRK_SPAN + 1)
/* dynamic shared memory of one channel, in 32-bit words: prefix rows
 * (S+1 pairs each), row bases (pairs), run-boundary queue                       */
#define TRK_WORK_WORDS(n, dtype, qcap) \
    ((size_t)TRK_ROWS((n) * (dtype)) * ((TRK_SPAN / (dtype)) + 2) * 2 + (size_t)(qcap))

/* carrier table: cost[i]=floor(cos(2*pi*i/32)*32+0.5), sint likewise
 * (src/sdrcmn.c:641-646); trk.cu re-derives the table with libm at start-up
 * and refuses to run if it differs                 */
#define SDRB200_COST_INIT {32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6,0,6,12,18,23,27,30,31}
#define SDRB200_SINT_INIT {0,6,12,18,23,27,30,31,32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6}
static __constant__ int8_t c_cost[32] = SDRB200_COST_INIT;
static __constant__ int8_t c_sint[32] = SDRB200_SINT_INIT;

struct trk_ncohdr_t { int32_t nseg; int32_t pad; double x_end; };

struct trk_smem_t {
    trk_ncohdr_t carr, code;                /* segment counts, accumulators after the last addition */
    nco_iseg_t  segs[4][TRK_MAXSEG];        /* segment storage: {carrier, code} x {this period, walked ahead} */
    nco_iseg_t *icarr;                      /* carrier NCO segments of this period (segs[0] or segs[2]) */
    nco_iseg_t *icode;                      /* code NCO segments of this period (segs[1] or segs[3])    */
    unsigned int sum[2 * SDRB200_MAXTAPS];  /* tap sums mod 2^32: I[0..], Q[MAXTAPS..] */
    unsigned int wtot[2][32];               /* per-warp totals of the block scan   */
    unsigned int carry[2];                  /* prefix carried between rounds       */
    int          lut[32];                   /* cos | sin << 16                     */
    int          qcount;                    /* run boundaries queued this period   */
    int          nslow[2];                  /* rows / spans left to the cooperative path */
    int          slow[2][128];
    int          vlast;                     /* replica value at the last position  */
    double   remc, remp;
    int32_t  err;
    int32_t  pad;
    int8_t   codetab[1024];                 /* +-1 chips                           */
};

/* the dynamic work area of one channel */
struct trk_work_t {
    uint2        *P2;                       /* [rows][S+1] running (I,Q) sums inside a row */
    uint2        *base;                     /* [rows] sums of all earlier rows             */
    unsigned int *queue;                    /* [qcap] p | sign<<31 of every replica sign change */
    int           qcap;
};
template <int DTYPE>
__device__ __forceinline__ trk_work_t trk_work_layout(unsigned int *w, int maxn, int qcap)
{
    trk_work_t k;
    const int rows = TRK_ROWS(maxn * DTYPE);
    k.P2 = (uint2 *)w;
    k.base = k.P2 + (size_t)rows * (TRK_SPAN / DTYPE + 1);
    k.queue = (unsigned int *)(k.base + rows);
    k.qcap = qcap;
    return k;
}

/* optional cycle stamps inside the worker phases (one thread of one team; pfp == NULL: off) */
struct trk_wprof_t { long long *p; long long t0; };
__device__ __forceinline__ void trk_wstamp(trk_wprof_t *w, int i)
{
    if (w && w->p) {
        long long t1; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t1)::"memory");
        atomicAdd((unsigned long long *)&w->p[i], (unsigned long long)(t1 - w->t0)); w->t0 = t1;   /* RED: no stall */
    }
}

template <int NW>
__device__ __forceinline__ void trk_bar_workers()
{
    asm volatile("bar.sync 1, %0;" ::"n"(NW) : "memory");
}

/* last segment with k0 <= k */
__device__ __forceinline__ int trk_iseg_find(const nco_iseg_t *seg, int nseg, int k)
{
    int lo = 0, hi = nseg - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (seg[mid].k0 <= k) lo = mid; else hi = mid - 1;
    }
    return lo;
}

/* cursor over the (int) values of an NCO at consecutive indices */
struct trk_cursor_t {
    const nco_iseg_t *seg;
    int nseg, s, next_k0, shift, negm;
    unsigned long long m;
    long long D;
    __device__ __forceinline__ void load(int k)
    {
        const nco_iseg_t *p = &seg[s];
        shift = p->shift; negm = -p->neg; D = p->D;
        m = p->X0 + (unsigned long long)((long long)(k - p->k0) * p->D);
        if (shift >= 64) { shift = 0; m = 0ull; D = 0; }       /* |value| < 1: (int) gives 0 */
        next_k0 = (s + 1 < nseg) ? seg[s + 1].k0 : 0x7fffffff;
    }
    __device__ __forceinline__ void start(const nco_iseg_t *sg, int n, int k)
    {
        seg = sg; nseg = n; s = trk_iseg_find(sg, n, k); load(k);
    }
    /* low 32 bits of (int) value at the current index k, then step to index k+1 */
    __device__ __forceinline__ int value_and_advance(int k)
    {
        const int t = (int)(unsigned int)(m >> shift);
        const int v = (t ^ negm) - negm;
        if (k + 1 >= next_k0) { s++; load(k + 1); } else m += (unsigned long long)D;
        return v;
    }
};

/* (int) value of an NCO at an arbitrary index k (own segment search) */
__device__ __forceinline__ int trk_value_at(const nco_iseg_t *seg, int nseg, int k)
{
    const nco_iseg_t *p = &seg[trk_iseg_find(seg, nseg, k)];
    const int sh = p->shift;
    if (sh >= 64) return 0;
    const unsigned long long m = p->X0 + (unsigned long long)((long long)(k - p->k0) * p->D);
    const int t = (int)(unsigned int)(m >> sh);
    return p->neg ? -t : t;
}

/* ---- phase 0: constants of the CTA (once per launch) ------------------------ */
__device__ __forceinline__ void trk_init_smem(trk_smem_t *sm)
{
    if (threadIdx.x < 32) sm->lut[threadIdx.x] = ((int)c_cost[threadIdx.x] & 0xffff) | ((int)c_sint[threadIdx.x] << 16);
    if (threadIdx.x == 0) { sm->err = 0; sm->icarr = sm->segs[0]; sm->icode = sm->segs[1]; }
}

/* ---- phase 1: NCO segment tables (two lanes, one per NCO warp) -------------- *
 * trk_zero_period clears the accumulators of the period; the two walks run on   *
 * lane TRK_NCO_A (carrier) and TRK_NCO_B (code).  The carrier remainder         *
 * (mixcarr's subtract loop, src/sdrcmn.c:665-668) is NOT computed here:         *
 * trk_carrier_rem() does it on the NCO lane while the workers are busy.         */
/* optional per-warp time stamps (build with -DTRK_DEBUG): slot s, warp w -> g_dbg[s*32+w] */
#ifdef TRK_DEBUG
__device__ long long g_dbg[4096];
__device__ int g_dbg_on;
__device__ __forceinline__ long long dbg_clock() { long long t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory"); return t; }
#define DBG(slot) do { if (g_dbg_on && blockIdx.x == 0 && (threadIdx.x & 31) == 0) g_dbg[(slot) * 32 + (threadIdx.x >> 5)] = dbg_clock(); } while (0)
#else
#define DBG(slot)
#endif

/* out of line on purpose: the walk is a long single-lane dependency chain on 64-bit
 * values; keeping it out of the 1024-thread kernel body (64-register cap) avoids spills.
 * (Measured and dropped: a warp-cooperative variant with the per-binade constants
 * precomputed one per lane and fetched by shuffle; the same constants in a shared-memory
 * table built once per frequency; the fast chain as a separate tight function (125
 * instead of ~200 instructions per binade).  None changed the ~21 k cycles of a code
 * walk: the time is the dependent 64-bit integer/convert latency of one lane sharing its
 * scheduler with seven busy worker warps, not the instruction count.  Handing the rows
 * to the worker warps of schedulers 2 and 3 first - so that the warps a period leaves
 * idle sit beside the NCO lanes - shortened the lane chains by ~1 k cycles but not the
 * period, and cost 7 % with 2048-sample periods (four busy warps on two schedulers).)  */
__device__ __noinline__ int trk_walk_call(double x, double step, int count, double wrap, nco_iseg_t *seg, double *x_end)
{
    return nco_walk_i(x, step, count, wrap, seg, TRK_MAXSEG, x_end);
}
__device__ __noinline__ double trk_rem_call(double x_end) { return nco_carrier_rem(x_end); }

__device__ __forceinline__ void trk_zero_period(trk_smem_t *sm)
{
    if (threadIdx.x < 2 * SDRB200_MAXTAPS) sm->sum[threadIdx.x] = 0u;
    if (threadIdx.x < 2) sm->carry[threadIdx.x] = 0u;
    if (threadIdx.x == 2) { sm->qcount = 0; sm->nslow[0] = 0; sm->nslow[1] = 0; }
}
__device__ __forceinline__ void trk_walk_carrier(trk_smem_t *sm, double ti, int n, double freq, double phi0)
{
    if (threadIdx.x == TRK_NCO_A) {
        double phi, ps, xe;
        nco_carrier_init(freq, phi0, ti, &phi, &ps);
        const int ns = trk_walk_call(phi, ps, n, 0.0, sm->icarr, &xe);
        if (ns < 0) sm->err = 1;
        sm->carr.nseg = ns < 0 ? 0 : ns; sm->carr.x_end = xe;
    }
    __syncwarp();
}
__device__ __forceinline__ void trk_walk_code(trk_smem_t *sm, double ti, int n, double crate, double coff,
                                              int smax, int coden)
{
    if (threadIdx.x == TRK_NCO_B) {
        double ci = __dmul_rn(ti, crate);                   /* ti*crate, sdrcmn.c:709 */
        double c0 = nco_code_init(coff, smax, ci, coden), xe;
        const int ns = trk_walk_call(c0, ci, n + 2 * smax, (double)coden, sm->icode, &xe);
        if (ns < 0) sm->err = 1;
        sm->code.nseg = ns < 0 ? 0 : ns; sm->code.x_end = xe;
        sm->remc = nco_code_rem(xe, smax, ci);
    }
    __syncwarp();
}
/* all of phase 1, for callers that follow it with one __syncthreads() */
__device__ __forceinline__ void trk_build_tables(trk_smem_t *sm, double ti, int n, double freq, double phi0,
                                                 double crate, double coff, int smax, int coden)
{
    trk_zero_period(sm);
    if ((threadIdx.x >> 5) == (TRK_NCO_A >> 5)) trk_walk_carrier(sm, ti, n, freq, phi0);
    if ((threadIdx.x >> 5) == (TRK_NCO_B >> 5)) trk_walk_code(sm, ti, n, crate, coff, smax, coden);
}
/* split barriers of the period loop: the workers start mixing as soon as the carrier
 * table exists (barrier 2: carrier warp + workers) and meet the code table before the
 * tap sums (barrier 3: code warp + workers, also the mix -> taps barrier)            */
__device__ __forceinline__ void trk_bar_carrier() { asm volatile("bar.sync 2, %0;" ::"n"(TRK_WORKERS + 32) : "memory"); }
__device__ __forceinline__ void trk_bar_code()    { asm volatile("bar.sync 3, %0;" ::"n"(TRK_WORKERS + 32) : "memory"); }
__device__ __forceinline__ void trk_carrier_rem(trk_smem_t *sm)
{
    if (threadIdx.x == TRK_NCO_A) sm->remp = trk_rem_call(sm->carr.x_end);
    __syncwarp();
}



/* byte b of w, sign-extended: one PRMT with sign replication */
__device__ __forceinline__ int trk_sbyte(unsigned int w, int b)
{
    const unsigned int n8 = 8u | (unsigned int)b;
    int d;                                  /* prmt: selector bit 3 = replicate the sign of the byte */
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(0u), "r"((n8 << 12) | (n8 << 8) | (n8 << 4) | (unsigned int)b));
    return d;
}

/* ---- phase 2 (workers): wipe-off + prefix sums ------------------------------- *
 * data: sample 0 of the window in global memory (any alignment; the aligned     *
 * words touching the window must be readable - the IF stream's guard and the    *
 * primitive's arena guarantee it).  Row u of wk.P2 receives the running (I,Q)   *
 * sums of worker u's samples, wk.base[u] the sums of all earlier rows.           */
/* NW worker threads, the first of them thread T0 of the CTA (defaults: the single-CTA layout) */
template <int DTYPE, int NW = TRK_WORKERS, int T0 = 64>
__device__ __forceinline__ void trk_mix_prefix(trk_smem_t *sm, const int8_t *__restrict__ data, int n,
                                               const trk_work_t &wk, trk_wprof_t *wp = nullptr)
{
    constexpr int S = TRK_SPAN / DTYPE, RS = S + 1;     /* samples per row, row stride (pairs) */
    constexpr int ROUND_BYTES = NW * TRK_SPAN, NWARP = NW / 32;
    const int u = threadIdx.x - T0, lane = threadIdx.x & 31, wwarp = u >> 5;
    const uintptr_t addr = (uintptr_t)data;
    const int a = (int)(addr & 3);
    const unsigned int *wbase = (const unsigned int *)(addr - a);
    const int nbytes = n * DTYPE + a;
    const int rounds = (nbytes + ROUND_BYTES - 1) / ROUND_BYTES;
    DBG(2);
    for (int r = 0; r < rounds; r++) {
        const int row = r * NW + u;
        const int byte0 = row * TRK_SPAN;
        const int kfirst = (byte0 - a) / DTYPE;          /* exact: a is even for I/Q streams */
        unsigned int runI = 0u, runQ = 0u;
        bool slow = false;
        if (byte0 < nbytes) {
            unsigned int w[TRK_SPAN / 4];
#pragma unroll
            for (int i = 0; i < TRK_SPAN / 4; i++) w[i] = (byte0 + 4 * i < nbytes) ? __ldg(wbase + (byte0 >> 2) + i) : 0u;
#define TRK_BYTE(j) trk_sbyte(w[(j) >> 2], (j) & 3)
            trk_cursor_t cur;
            cur.start(sm->icarr, sm->carr.nseg, max(kfirst, 0));
            trk_wstamp(wp, 26);
            uint2 *rowp = wk.P2 + (size_t)row * RS;
#define TRK_MIX(j, idx)                                                                             \
            {                                                                                       \
                const int cs = sm->lut[(idx) & SDRB200_CMASK];                                      \
                const int c = (int)(short)cs, sn = cs >> 16;                                        \
                if (DTYPE == 1) {                                                                   \
                    const int x = TRK_BYTE(j);                                                      \
                    runI += (unsigned int)(c * x); runQ += (unsigned int)(sn * x);   /* sdrcmn.c:660-661 */ \
                } else {                                                                            \
                    const int x0 = TRK_BYTE(2 * (j)), x1 = TRK_BYTE(2 * (j) + 1);                   \
                    runI += (unsigned int)(c * x0 - sn * x1);                        /* sdrcmn.c:654-655 */ \
                    runQ += (unsigned int)(sn * x0 + c * x1);                                       \
                }                                                                                   \
            }
            if (kfirst >= 0 && kfirst + S <= n && kfirst + S <= cur.next_k0) {
                /* the whole row lies in one NCO segment: branch-free, incremental phase */
                unsigned long long m = cur.m;
                const unsigned long long D = (unsigned long long)cur.D;
                const int sh = cur.shift, ng = cur.negm;
#pragma unroll
                for (int j = 0; j < S; j++) {
                    const int tt = (int)(unsigned int)(m >> sh);
                    TRK_MIX(j, (tt ^ ng) - ng)
                    rowp[j] = make_uint2(runI, runQ);
                    m += D;
                }
            } else slow = true;
#undef TRK_MIX
#undef TRK_BYTE
        }
        /* rows that touch the window edges or more than one NCO segment go to a shared
         * list; every warp then takes rows from it round-robin, one LANE per sample (own
         * segment search, samples re-read from global memory) and a shuffle scan forms
         * the running sums.  Such rows cluster at the window start (low binades), so
         * sharing them keeps one warp from becoming the tail of the phase.            */
        if (slow) {
            const int i = atomicAdd(&sm->nslow[0], 1);
            if (i < 128) sm->slow[0][i] = row; else sm->err = 3;
        }
        trk_wstamp(wp, 16);
        trk_bar_workers<NW>();
        trk_wstamp(wp, 17);
        {
            const int ns_ = min(sm->nslow[0], 128);
            for (int it = wwarp; it < ns_; it += NWARP) {
                const int srow = sm->slow[0][it];
                const int kk = (srow * TRK_SPAN - a) / DTYPE + lane;
                unsigned int I = 0u, Q = 0u;
                if (lane < S && (unsigned int)kk < (unsigned int)n) {
                    const int cs = sm->lut[trk_value_at(sm->icarr, sm->carr.nseg, kk) & SDRB200_CMASK];
                    const int c = (int)(short)cs, sn = cs >> 16;
                    if (DTYPE == 1) {
                        const int x = data[kk];
                        I = (unsigned int)(c * x); Q = (unsigned int)(sn * x);
                    } else {
                        const int x0 = data[2 * kk], x1 = data[2 * kk + 1];
                        I = (unsigned int)(c * x0 - sn * x1); Q = (unsigned int)(sn * x0 + c * x1);
                    }
                }
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned int uI = __shfl_up_sync(0xffffffffu, I, o), uQ = __shfl_up_sync(0xffffffffu, Q, o);
                    if (lane >= o) { I += uI; Q += uQ; }
                }
                if (lane < S) wk.P2[(size_t)srow * RS + lane] = make_uint2(I, Q);
                if (lane == S - 1) wk.base[srow] = make_uint2(I, Q);      /* row total, picked up by the owner */
            }
        }
        trk_wstamp(wp, 18);
        trk_bar_workers<NW>();
        trk_wstamp(wp, 19);
        if (slow) { const uint2 tt = wk.base[row]; runI = tt.x; runQ = tt.y; }
        if (u == 0) sm->nslow[0] = 0;                     /* next round */
        DBG(3);
        /* block-wide exclusive scan of the row totals over the worker warps */
        unsigned int incI = runI, incQ = runQ;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int uI = __shfl_up_sync(0xffffffffu, incI, o), uQ = __shfl_up_sync(0xffffffffu, incQ, o);
            if (lane >= o) { incI += uI; incQ += uQ; }
        }
        if (lane == 31) { sm->wtot[0][wwarp] = incI; sm->wtot[1][wwarp] = incQ; }
        trk_bar_workers<NW>();
        /* lane l holds the total of worker warp l; inclusive scan over lanes */
        unsigned int wI = (lane < NWARP) ? sm->wtot[0][lane] : 0u, wQ = (lane < NWARP) ? sm->wtot[1][lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int uI = __shfl_up_sync(0xffffffffu, wI, o), uQ = __shfl_up_sync(0xffffffffu, wQ, o);
            if (lane >= o) { wI += uI; wQ += uQ; }
        }
        const unsigned int allI = __shfl_sync(0xffffffffu, wI, 31), allQ = __shfl_sync(0xffffffffu, wQ, 31);
        unsigned int prevI = __shfl_sync(0xffffffffu, wI, (wwarp + 31) & 31), prevQ = __shfl_sync(0xffffffffu, wQ, (wwarp + 31) & 31);
        if (wwarp == 0) { prevI = 0u; prevQ = 0u; }
        if (byte0 < nbytes + TRK_SPAN)
            wk.base[row] = make_uint2(sm->carry[0] + prevI + incI - runI, sm->carry[1] + prevQ + incQ - runQ);
        if (rounds > 1) {
            trk_bar_workers<NW>();                       /* everyone has read wtot / carry */
            if (u == 0) { sm->carry[0] += allI; sm->carry[1] += allQ; }
            trk_bar_workers<NW>();
        }
        trk_wstamp(wp, 20);
    }
}

/* both words of the same prefix: (sum I, sum Q) over samples [0, k) */
template <int DTYPE>
__device__ __forceinline__ uint2 trk_prefix2_at(const trk_work_t &wk, int a, int k)
{
    constexpr int S = TRK_SPAN / DTYPE, RS = S + 1;
    if (k <= 0) return make_uint2(0u, 0u);
    const int byte = (k - 1) * DTYPE + a;
    const int row = byte / TRK_SPAN;
    const int j = (byte - row * TRK_SPAN) / DTYPE;
    const uint2 b = wk.base[row], p = wk.P2[(size_t)row * RS + j];
    return make_uint2(b.x + p.x, b.y + p.y);
}

/* ---- phase 3 (workers): replica sign changes -> tap sums --------------------- *
 * taps[i] (i < ns): correlator offsets in samples (trk.corrp).  Output order as  *
 * the reference: {P, E1, L1, E2, L2, ...} (sdrcmn.c:682-683); E_i uses the       *
 * replica shifted by -taps[i], L_i by +taps[i] (sdrcmn.c:712-715).               *
 *  (a) every worker walks its share of the n + 2*smax replica positions with the *
 *      integer code NCO and queues the positions where the chip VALUE changes;   *
 *  (b) the queue is consumed with one LANE PER TAP SUM (I and Q of every tap):   *
 *      a lane looks its prefix value up and accumulates privately, so there is   *
 *      no cross-lane reduction and no divergence.                                *
 * Preceded by a CTA barrier (prefix rows complete), followed by one before       *
 * sm->sum is read.                                                               */
template <int DTYPE, int NW = TRK_WORKERS, int T0 = 64>
__device__ __forceinline__ void trk_tap_sums(trk_smem_t *sm, const trk_work_t &wk, const int8_t *data, int n,
                                             const int *taps, int ns, int smax, trk_wprof_t *wp = nullptr)
{
    const int t = threadIdx.x - T0, lane = threadIdx.x & 31, wwarp = t >> 5;
    const int a = (int)((uintptr_t)data & 3);
    const int N = n + 2 * smax, ntap = 1 + 2 * ns;
    const int PP = (N + NW - 1) / NW;
    const int p0 = t * PP, p1 = min(p0 + PP, N);
    /* (a) */
    DBG(5);
    {
        bool slow = false;
        if (p0 < N) {
            trk_cursor_t cur;
            slow = true;
            if (p0 > 0) {
                cur.start(sm->icode, sm->code.nseg, p0 - 1);
                trk_wstamp(wp, 27);
                if (p1 <= cur.next_k0 && cur.negm == 0 && cur.D > 0) {
                    /* the span and its left neighbour lie in one segment of a rising code
                     * phase: chips prev+1 .. clast start inside the span; the start of chip c
                     * is the smallest j with (m + j*D) >> sh >= c, found from a floating-point
                     * estimate and corrected with exact integer tests                        */
                    slow = false;
                    const unsigned long long D = (unsigned long long)cur.D;
                    const int prev = (int)(unsigned int)(cur.m >> cur.shift);
                    const int clast = (int)(unsigned int)((cur.m + (unsigned long long)(p1 - p0) * D) >> cur.shift);
                    const float invD = 1.0f / (float)cur.D;
                    int vprev = (int)sm->codetab[prev];
                    for (int c = prev + 1; c <= clast; c++) {
                        const unsigned long long target = (unsigned long long)c << cur.shift;
                        long long j = (long long)((float)(target - cur.m) * invD);
                        j = max(1ll, min(j, (long long)(p1 - p0)));
                        while (j > 1 && cur.m + (unsigned long long)(j - 1) * D >= target) j--;
                        while (cur.m + (unsigned long long)j * D < target) j++;
                        const int v = (int)sm->codetab[c];
                        if (v != vprev) {
                            const int slot = atomicAdd(&sm->qcount, 1);
                            if (slot < wk.qcap) wk.queue[slot] = (unsigned int)(p0 - 1 + (int)j) | (vprev < v ? 0x80000000u : 0u);
                            else sm->err = 2;
                        }
                        vprev = v;
                    }
                    if (p1 == N) sm->vlast = vprev;
                }
            }
        }
        /* spans at the window start or across NCO segments: shared list, then every warp
         * takes spans round-robin with one LANE per position                            */
        if (slow) {
            const int i = atomicAdd(&sm->nslow[1], 1);
            if (i < 128) sm->slow[1][i] = t; else sm->err = 3;
        }
        trk_wstamp(wp, 21);
        trk_bar_workers<NW>();
        trk_wstamp(wp, 22);
        const int ns_ = min(sm->nslow[1], 128);
        for (int it = wwarp; it < ns_; it += NW / 32) {
            const int q0 = sm->slow[1][it] * PP, q1 = min(q0 + PP, N);
            for (int pb = q0; pb < q1; pb += 32) {
                const int p = pb + lane;
                if (p < q1) {
                    const int v = (int)sm->codetab[trk_value_at(sm->icode, sm->code.nseg, p)];
                    const int vl = (p > 0) ? (int)sm->codetab[trk_value_at(sm->icode, sm->code.nseg, p - 1)] : v;
                    if (v != vl) {
                        const int slot = atomicAdd(&sm->qcount, 1);
                        if (slot < wk.qcap) wk.queue[slot] = (unsigned int)p | (vl < v ? 0x80000000u : 0u);
                        else sm->err = 2;
                    }
                    if (p == N - 1) sm->vlast = v;
                }
            }
        }
    }
    DBG(6);
    trk_wstamp(wp, 23);
    trk_bar_workers<NW>();
    trk_wstamp(wp, 24);
    DBG(7);
    /* (b) one lane per (tap, queue entry): a lane reads the (I,Q) prefix pair of its position
     * once and accumulates both sums, so a tap costs one lookup per entry; the 32 / ntap lane
     * groups of a warp take consecutive entries                                            */
    const int qn = min(sm->qcount, wk.qcap);
    const int tpw = min(ntap, 32), G = 32 / tpw;
    const int g = lane / tpw, tl = lane - g * tpw;
    for (int tap0 = 0; tap0 < ntap; tap0 += tpw) {
        const int tap = tap0 + tl;
        if (g >= G || tap >= ntap) continue;
        int off = 0;
        if (tap > 0) { const int s = taps[(tap - 1) >> 1]; off = ((tap - 1) & 1) ? s : -s; }
        const int bias = smax + off;
        unsigned int accI = 0u, accQ = 0u;
#pragma unroll 4
        for (int e = wwarp * G + g; e < qn; e += (NW / 32) * G) {
            const unsigned int entry = wk.queue[e];
            const unsigned int delta = (entry >> 31) ? 0xfffffffeu : 2u;        /* r_{q-1} - r_q */
            const uint2 pf = trk_prefix2_at<DTYPE>(wk, a, min(max((int)(entry & 0x7fffffffu) - bias, 0), n));
            accI += delta * pf.x; accQ += delta * pf.y;
        }
        if (wwarp == 0 && g == 0) {                                             /* r_last * P[n] */
            const uint2 pf = trk_prefix2_at<DTYPE>(wk, a, n);
            accI += (unsigned int)sm->vlast * pf.x; accQ += (unsigned int)sm->vlast * pf.y;
        }
        if (accI) atomicAdd(&sm->sum[tap], accI);
        if (accQ) atomicAdd(&sm->sum[SDRB200_MAXTAPS + tap], accQ);
    }
    trk_wstamp(wp, 25);
    DBG(8);
}

#endif
