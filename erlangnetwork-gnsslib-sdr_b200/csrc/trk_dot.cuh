/* trk_dot.cuh - correlator() of one code period (src/sdrcmn.c:687-722) as packed int8 dot
 * products; the worker side of the batched tracking kernel k_trk_cl.
 *
 * The reference forms I[k] = cost[idx_k]*x[k], Q[k] = sint[idx_k]*x[k] (mixcarr,
 * src/sdrcmn.c:647-663), the replica r[p] = code[(int)coff_p] (rescode, src/sdrcmn.c:608-621)
 * and 1 + 2*ns sums  sum_k I[k]*r[k + smax + off]  (dot_23 / dot_22, src/sdrcmn.c:310-354).
 * All factors are small integers, so the sums are exact in int32 in any order.  Here:
 *
 *   x      stays packed as loaded: one 32-bit word = 4 real samples (2 I/Q pairs);
 *   C4/S4  the carrier table values of the word's samples packed the same way (int8 each;
 *          for I/Q data {c,-s,c,-s} and {s,c,s,c}, src/sdrcmn.c:654-655), built once per word
 *          from the integer NCO segments (nco_exact.h) and shared by all taps;
 *   r      one BIT per window position (1: replica = +1) in a shared-memory array that the
 *          team fills once per period from the code NCO segments; a tap reads its 20 bits with
 *          one funnel shift and expands them to byte masks through a small table;
 *   tap    = 2 * sum_{r=+1} x*c  -  sum_all x*c      (r = +-1)
 *          = 2 * dp4a(x & mask, C4) - dp4a(x, C4)    per word, I and Q alike.
 *
 * No prefix sums, no queues, no atomics: a thread's 20 bytes are independent of every other
 * thread's, the only exchanges are the code-bit array (one barrier) and the per-warp partial
 * sums (one barrier).  Rows that touch more than one NCO segment take a per-sample cursor path.
 */
#ifndef SDRB200_TRK_DOT_CUH
#define SDRB200_TRK_DOT_CUH

#include "trk_device.cuh"

/* cycle stamps inside the worker phases cost registers in the hot loops: compiled in with -DTRKD_PROFILE only */
#ifdef TRKD_PROFILE
#define TRKD_STAMP(i) trk_wstamp(wp, (i))
#else
#define TRKD_STAMP(i) do { } while (0)
#endif
#define TRKD_PAD   64                        /* guard positions in front of the code-bit array */
#define TRKD_NPART (2 * SDRB200_MAXTAPS + 2) /* per-warp partial sums: I[tap].., Q[tap].., all I, all Q */

/* dynamic work area of one team, in 32-bit words (the mask table is aligned to its 2 KB) */
#define TRKD_CBIT_WORDS(n, smax2) ((((n) + (smax2) + TRKD_PAD + 31) >> 5) + 4)
#define TRKD_WORK_WORDS(n, smax2, nwarp) \
    ((size_t)512 + 512 + 64 + 256 + 64 + 2 * (size_t)(nwarp) * TRKD_NPART + TRKD_CBIT_WORDS(n, smax2))

struct trkd_work_t {
    uint2        *mlut;                      /* [256] byte (or nibble) of code bits -> byte masks of two words */
    int          *clut;                      /* [2][32] carrier words {c, s, -s, c}; [1]: entry (i+1)&31 for negative phases */
    uint8_t      *cpos;                      /* [1024] 1: chip value is +1                 */
    int          *tbias;                     /* [MAXTAPS] smax + offset of tap t (order P, E1, L1, E2, ..., src/sdrcmn.c:712-715) */
    unsigned int *part;                      /* [2][nwarp][TRKD_NPART] partial sums, by period parity */
    unsigned int *cbits;                     /* code bits: bit p + TRKD_PAD <-> window position p */
};

__device__ __forceinline__ trkd_work_t trkd_work_layout(unsigned int *w, int nwarp)
{
    trkd_work_t k;
    const unsigned int sa = (unsigned int)__cvta_generic_to_shared(w);
    w += ((2048u - (sa & 2047u)) & 2047u) >> 2;      /* < 512 words of slack */
    k.mlut = (uint2 *)w;                     /* 512 words */
    k.clut = (int *)(w + 512);               /* 64        */
    k.cpos = (uint8_t *)(w + 512 + 64);      /* 256       */
    k.tbias = (int *)(w + 512 + 64 + 256);   /* 64        */
    k.part = w + 512 + 64 + 256 + 64;
    k.cbits = k.part + 2 * (size_t)nwarp * TRKD_NPART;
    return k;
}

/* once per launch (all threads of the CTA, followed by a barrier of the caller) */
template <int DTYPE>
__device__ __forceinline__ void trkd_init(const trkd_work_t &wk, const trk_smem_t *sm, int clen, const int *taps, int ns,
                                          int nthreads)
{
    for (int b = threadIdx.x; b < 256; b += nthreads) {
        unsigned int m[2] = {0u, 0u};
        if (DTYPE == 1) {
#pragma unroll
            for (int i = 0; i < 8; i++) if ((b >> i) & 1) m[i >> 2] |= 0xFFu << (8 * (i & 3));
        } else {
#pragma unroll
            for (int i = 0; i < 4; i++) if ((b >> i) & 1) m[i >> 1] |= 0xFFFFu << (16 * (i & 1));
        }
        wk.mlut[b] = make_uint2(m[0], m[1]);
    }
    for (int i = threadIdx.x; i < 64; i += nthreads) {
        const int j = (i < 32) ? i : ((i + 1) & 31);
        const int c = (int)c_cost[j], s = (int)c_sint[j];
        wk.clut[i] = (c & 0xFF) | ((s & 0xFF) << 8) | (((-s) & 0xFF) << 16) | ((c & 0xFF) << 24);
    }
    for (int i = threadIdx.x; i < 1024; i += nthreads) wk.cpos[i] = (i < clen && sm->codetab[i] > 0) ? 1 : 0;
    if (threadIdx.x < 1 + 2 * ns) {
        const int tap = threadIdx.x, smax = taps[ns - 1];
        int off = 0;
        if (tap > 0) { const int sft = taps[(tap - 1) >> 1]; off = ((tap - 1) & 1) ? sft : -sft; }
        wk.tbias[tap] = smax + off;
    }
}

__device__ __forceinline__ unsigned int trkd_prmt(unsigned int a, unsigned int b, unsigned int sel)
{
    unsigned int d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}
__device__ __forceinline__ int trkd_dp4a(unsigned int a, unsigned int b, int c)
{
    int d;
    asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

/* index of the last segment with k0 <= k, found by the whole warp for ONE k (lane 0's) */
__device__ __forceinline__ int trkd_warp_find(const nco_iseg_t *seg, int nseg, int k, int lane)
{
    int cnt = 0;
    for (int b = 0; b < nseg; b += 32) {
        const int i = b + lane;
        const bool le = (i < nseg) && (seg[i].k0 <= k);
        cnt += __popc(__ballot_sync(0xffffffffu, le));
    }
    return cnt > 0 ? cnt - 1 : 0;
}

/* a segment in "top bits" form: the wanted field of (X0 + (k - k0) * D) >> shift sits in the top
 * BITS bits of u(k) = u0 + (k - k0) * du (mod 2^64); higher bits are dropped, which is exact for
 * a field that is taken modulo 2^BITS (carrier: & 31) or known to fit (chip index < 1024)       */
struct trkd_lin_t { unsigned long long u, du; };
template <int BITS>
__device__ __forceinline__ trkd_lin_t trkd_lin_at(const nco_iseg_t *p, int k)
{
    trkd_lin_t o;
    const int sh = p->shift;
    if (sh > 64 - BITS) { o.u = 0ull; o.du = 0ull; return o; }    /* |value| < 1: the field is 0 */
    const int ls = 64 - BITS - sh;
    const unsigned long long D = (unsigned long long)p->D;
    const unsigned long long m = p->X0 + (unsigned long long)((long long)(k - p->k0)) * D;
    o.u = m << ls; o.du = D << ls;
    return o;
}

/* the BITS-bit field of the NCO value at an arbitrary index k (one lane; s: a segment at or before k's) */
template <int BITS>
__device__ __forceinline__ unsigned int trkd_field_at(const nco_iseg_t *seg, int nseg, int s, int k, int *neg)
{
    while (s + 1 < nseg && seg[s + 1].k0 <= k) s++;
    const trkd_lin_t l = trkd_lin_at<BITS>(&seg[s], k);
    *neg = seg[s].neg;
    return (unsigned int)(l.u >> (64 - BITS));
}

/* ---- carrier words of one row ------------------------------------------------------------
 * cw[i], sw[i]: the two int8 weight words of data word i of the row (see the file header).
 * Called by whole warps.  A row inside one NCO segment is a linear phase: five index bits from
 * the top of a 64-bit accumulator per sample.  Rows that touch a segment border (a handful per
 * period, most of them at the window start where the binades are short) are done by the warp
 * together, one lane per sample, and handed to their owner by shuffles.                     */
template <int DTYPE, int NWORD>
__device__ __forceinline__ void trkd_carrier_row(const trkd_work_t &wk, const nco_iseg_t *seg, int nseg, int s, int kfirst,
                                                 bool live, int lane, unsigned int (&cw)[NWORD], unsigned int (&sw)[NWORD])
{
    constexpr int S = NWORD * 4 / DTYPE;
    const int next_k0 = (s + 1 < nseg) ? seg[s + 1].k0 : 0x7fffffff;
    const bool fast = (kfirst + S <= next_k0);
    unsigned int e[S];
#pragma unroll
    for (int j = 0; j < S; j++) e[j] = 0u;
    if (fast && live) {
        const nco_iseg_t *p = &seg[s];
        trkd_lin_t l = trkd_lin_at<5>(p, kfirst);
        const int *lut = wk.clut;
        if (p->neg) { l.u = ~l.u; l.du = 0ull - l.du; lut += 32; }   /* (-t) & 31 = (top5(~u) + 1) & 31 */
#pragma unroll
        for (int j = 0; j < S; j++) {
            e[j] = (unsigned int)lut[(unsigned int)(l.u >> 59)];
            l.u += l.du;
        }
    }
    unsigned int slow = __ballot_sync(0xffffffffu, live && !fast);
    while (slow) {
        const int src = __ffs(slow) - 1;
        slow &= slow - 1;
        const int kf = __shfl_sync(0xffffffffu, kfirst, src);
        const int ss = __shfl_sync(0xffffffffu, s, src);
        unsigned int ev = 0u;
        {
            const int k = kf + lane;
            unsigned int idx = 0u;
            if (lane < S && k >= 0) {
                int neg;
                const unsigned int t = trkd_field_at<5>(seg, nseg, ss, k, &neg);
                idx = neg ? ((32u - t) & 31u) : t;
            }
            ev = (unsigned int)wk.clut[idx];
        }
#pragma unroll
        for (int j = 0; j < S; j++) {
            const unsigned int v = __shfl_sync(0xffffffffu, ev, j);
            if (lane == src) e[j] = v;
        }
    }
#pragma unroll
    for (int i = 0; i < NWORD; i++) {
        if (DTYPE == 1) {
            const unsigned int t01 = trkd_prmt(e[4 * i], e[4 * i + 1], 0x5140u);
            const unsigned int t23 = trkd_prmt(e[4 * i + 2], e[4 * i + 3], 0x5140u);
            cw[i] = trkd_prmt(t01, t23, 0x5410u);
            sw[i] = trkd_prmt(t01, t23, 0x7632u);
        } else {
            cw[i] = trkd_prmt(e[2 * i], e[2 * i + 1], 0x6420u);      /* { c, -s, c, -s } */
            sw[i] = trkd_prmt(e[2 * i], e[2 * i + 1], 0x7531u);      /* { s,  c, s,  c } */
        }
    }
}

/* ---- code bits of the period: positions [0, N) -> wk.cbits (16 positions per item) --------
 * Same split: an item inside one code NCO segment walks a 10-bit chip index down the top of a
 * 64-bit accumulator; items on a segment border or at the window end are done by the warp with
 * one lane per position and collected with a ballot.                                        */
template <int NW, int T0>
__device__ __forceinline__ void trkd_code_bits(const trkd_work_t &wk, const nco_iseg_t *seg, int nseg, int N, int p_lo, int p_hi)
{
    constexpr int NWARP = NW / 32;
    const int t = threadIdx.x - T0, lane = threadIdx.x & 31, wwarp = t >> 5;
    const int it_lo = (p_lo + TRKD_PAD) >> 4, it_hi = (p_hi + TRKD_PAD + 15) >> 4;
    unsigned short *out = (unsigned short *)wk.cbits;
    /* consecutive items go to different warps: the borders of the short binades at the window
     * start would otherwise all land in one warp                                            */
    for (int base = it_lo; base < it_hi; base += NW) {
        /* a partial last round stays contiguous so that only the warps it needs run it */
        if (base + NW > it_hi && base + (t - lane) >= it_hi) break;
        const int it = base + ((base + NW <= it_hi) ? lane * NWARP + wwarp : t);
        const int p0 = (it << 4) - TRKD_PAD;                     /* >= 0, multiple of 16 */
        const bool valid = it < it_hi;
        const int s = trk_iseg_find(seg, nseg, valid ? p0 : 0);
        const int next_k0 = (s + 1 < nseg) ? seg[s + 1].k0 : 0x7fffffff;
        const bool fast = valid && p0 + 16 <= next_k0 && p0 + 16 <= N;
        unsigned int bits = 0u;
        if (fast) {
            trkd_lin_t l = trkd_lin_at<10>(&seg[s], p0);
#pragma unroll
            for (int j = 0; j < 16; j++) {
                bits = (bits << 1) | (unsigned int)wk.cpos[(unsigned int)(l.u >> 54)];
                l.u += l.du;
            }
            bits = __brev(bits) >> 16;
        }
        unsigned int slow = __ballot_sync(0xffffffffu, valid && !fast);
        while (slow) {
            const int src = __ffs(slow) - 1;
            slow &= slow - 1;
            const int ps = __shfl_sync(0xffffffffu, p0, src) + lane;
            const int ss = __shfl_sync(0xffffffffu, s, src);
            int bit = 0;
            if (lane < 16 && ps < N) {
                int neg;
                bit = (int)wk.cpos[trkd_field_at<10>(seg, nseg, ss, ps, &neg)];
            }
            const unsigned int b = __ballot_sync(0xffffffffu, bit) & 0xFFFFu;
            if (lane == src) bits = b;
        }
        if (valid) out[it] = (unsigned short)bits;
    }
}

__device__ __forceinline__ uint2 trkd_lds64(unsigned int addr)
{
    uint2 v;
    asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));   /* the mask table is constant */
    return v;
}

/* ---- one period: tap sums mod 2^32 into sm->sum (I at [tap], Q at [MAXTAPS + tap]) --------
 * Called by the NW worker threads (first one: thread T0).  Contains two worker barriers; the
 * caller places one more before sm->sum is consumed by other threads.  `par` alternates per
 * call (partial-sum buffers).  wait_carr() / wait_code(): the caller's waits for the carrier and
 * code segment tables cseg / kseg; they return the segment counts.  The samples are requested
 * before the first wait, the carrier words are formed before the second.
 * Lane l of a warp keeps the warp's sums of tap l (tap 32, if there is one, sits on lane 0).
 * Team ti of nteams takes an equal share of the window's rows; the sums are then partial.    */
template <int DTYPE, int SPAN, int NW, int T0, typename WaitCarr, typename WaitCode>
__device__ __forceinline__ void trkd_period(trk_smem_t *sm, const trkd_work_t &wk, const int8_t *__restrict__ data, int n,
                                            int ns, int smax, int par, int ti, int nteams,
                                            const nco_iseg_t *cseg, const nco_iseg_t *kseg,
                                            WaitCarr wait_carr, WaitCode wait_code, trk_wprof_t *wp = nullptr)
{
    constexpr int NWORD = SPAN / 4, NWARP = NW / 32;
    constexpr int BPW = 4 / DTYPE;                               /* code bits per data word */
    constexpr unsigned int PMASK = ((1u << (2 * BPW)) - 1u) << 3;
    const int u = threadIdx.x - T0, lane = threadIdx.x & 31, wwarp = u >> 5;
    const uintptr_t addr = (uintptr_t)data;
    const int a = (int)(addr & 3);
    const unsigned int *wbase = (const unsigned int *)(addr - a);
    const int nbytes = n * DTYPE + a;
    /* this team's share of the rows (SPAN bytes each) of the aligned span that covers the window */
    const int total_rows = (nbytes + SPAN - 1) / SPAN;
    const int row_lo = (int)((long long)total_rows * ti / nteams), row_hi = (int)((long long)total_rows * (ti + 1) / nteams);
    const int nrows = row_hi - row_lo;
    const int rounds = (nrows + NW - 1) / NW;
    const int ntap = 1 + 2 * ns;
    const unsigned int mbase = (unsigned int)__cvta_generic_to_shared(wk.mlut);     /* 2 KB aligned */
    int ncseg = 0;
    unsigned int x[NWORD], cw[NWORD], sw[NWORD];
    int allI = 0, allQ = 0, accI = 0, accQ = 0, acc32I = 0, acc32Q = 0;
    for (int r = 0; r < rounds; r++) {
        /* consecutive rows go to different warps (see trkd_code_bits) */
        const int ri = r * NW + lane * NWARP + wwarp;
        const bool live = ri < nrows;
        const int row = row_lo + (live ? ri : 0);
        const int byte0 = row * SPAN;
        const int kfirst = (byte0 - a) / DTYPE;                  /* exact: a is even for I/Q streams */
#pragma unroll
        for (int i = 0; i < NWORD; i++) x[i] = (live && byte0 + 4 * i < nbytes) ? __ldg(wbase + (byte0 >> 2) + i) : 0u;
        if (r == 0) ncseg = wait_carr();
        TRKD_STAMP(26);
        trkd_carrier_row<DTYPE, NWORD>(wk, cseg, ncseg, trk_iseg_find(cseg, ncseg, max(kfirst, 0)), kfirst, live, lane, cw, sw);
        /* bytes outside the window contribute nothing */
        if (live && (byte0 < a || byte0 + SPAN > nbytes)) {
#pragma unroll
            for (int i = 0; i < NWORD; i++) {
                unsigned int m = 0u;
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const int bb = byte0 + 4 * i + b;
                    if (bb >= a && bb < nbytes) m |= 0xFFu << (8 * b);
                }
                x[i] &= m;
            }
        }
        TRKD_STAMP(16);
        if (r == 0) {
            const int nkseg = wait_code();
            TRKD_STAMP(27);
            {
                const int N = n + 2 * smax;
                const int k_lo = (row_lo * SPAN - a) / DTYPE, k_hi = (row_hi * SPAN - a) / DTYPE;
                trkd_code_bits<NW, T0>(wk, kseg, nkseg, N, max(k_lo, 0), min(k_hi + 2 * smax, N));
            }
            TRKD_STAMP(21);
            trk_bar_workers<NW>();
            TRKD_STAMP(22);
        }
#pragma unroll
        for (int i = 0; i < NWORD; i++) { allI = trkd_dp4a(x[i], cw[i], allI); allQ = trkd_dp4a(x[i], sw[i], allQ); }
        /* taps: bit kfirst + (smax + off) + j of the code-bit array, j < S; rows past the window
         * (all-zero data) read position 0                                                     */
        const int kbp = kfirst + TRKD_PAD;                       /* >= 0: kfirst >= -3, bias >= 0 */
        const unsigned int *cbw = wk.cbits;
        const int *tb = wk.tbias;
        auto tap_sums = [&](int tap, int &pI, int &pQ) {
            const int bi = kbp + tb[tap];
            const unsigned int *cb = cbw + (bi >> 5);
            const unsigned int f = __funnelshift_r(cb[0], cb[1], bi);
            pI = 0; pQ = 0;
#pragma unroll
            for (int i = 0; i < NWORD; i += 2) {
                const int sh = BPW * i - 3;
                const unsigned int fi = (sh < 0) ? (f << (-sh)) : (f >> sh);
                const uint2 m = trkd_lds64((fi & PMASK) | mbase);
                const unsigned int x0 = x[i] & m.x;
                pI = trkd_dp4a(x0, cw[i], pI); pQ = trkd_dp4a(x0, sw[i], pQ);
                if (i + 1 < NWORD) {
                    const unsigned int x1 = x[i + 1] & m.y;
                    pI = trkd_dp4a(x1, cw[i + 1], pI); pQ = trkd_dp4a(x1, sw[i + 1], pQ);
                }
            }
            pI = __reduce_add_sync(0xffffffffu, pI);
            pQ = __reduce_add_sync(0xffffffffu, pQ);
        };
        const int nt32 = min(ntap, 32);
#pragma unroll 1
        for (int tap = 0; tap < nt32; tap++) {
            int pI, pQ;
            tap_sums(tap, pI, pQ);
            if (lane == tap) { accI += pI; accQ += pQ; }
        }
        if (ntap > 32) {
            int pI, pQ;
            tap_sums(32, pI, pQ);
            acc32I += pI; acc32Q += pQ;
        }
    }
    allI = __reduce_add_sync(0xffffffffu, allI);
    allQ = __reduce_add_sync(0xffffffffu, allQ);
    {
        unsigned int *part = wk.part + ((size_t)par * NWARP + wwarp) * TRKD_NPART;
        if (lane < ntap) { part[lane] = (unsigned int)accI; part[SDRB200_MAXTAPS + lane] = (unsigned int)accQ; }
        if (lane == 0) {
            part[2 * SDRB200_MAXTAPS] = (unsigned int)allI; part[2 * SDRB200_MAXTAPS + 1] = (unsigned int)allQ;
            if (ntap > 32) { part[32] = (unsigned int)acc32I; part[SDRB200_MAXTAPS + 32] = (unsigned int)acc32Q; }
        }
    }
    TRKD_STAMP(25);
    trk_bar_workers<NW>();
    TRKD_STAMP(24);
    /* tap = 2 * sum_{r=+1} - sum_all, over the warps' partial sums */
    if (u < 2 * ntap) {
        const int tap = (u < ntap) ? u : u - ntap;
        const int col = (u < ntap) ? tap : SDRB200_MAXTAPS + tap;
        const int acol = (u < ntap) ? 2 * SDRB200_MAXTAPS : 2 * SDRB200_MAXTAPS + 1;
        const unsigned int *pp = wk.part + (size_t)par * NWARP * TRKD_NPART;
        unsigned int sp = 0u, sa = 0u;
#pragma unroll 4
        for (int w = 0; w < NWARP; w++) { sp += pp[w * TRKD_NPART + col]; sa += pp[w * TRKD_NPART + acol]; }
        sm->sum[col] = 2u * sp - sa;
    }
}

#endif
