/* nco_exact.h - bit-exact, parallel-friendly replay of the reference's NCOs.
 *
 * The reference advances its carrier and code NCOs by REPEATED double addition
 * and truncates each value to an integer table index:
 *     mixcarr : phi += ps;  index = ((int)phi) & 31      src/sdrcmn.c:650-662
 *     rescode : if (coff>=len) coff-=len; chip=(int)coff; coff+=ci   :614-617
 *     mixcarr : while (prem>DPI) prem-=DPI;              src/sdrcmn.c:666-668
 * A closed form phi0+k*ps differs in the last bits and can flip an index, so
 * to stay bit-identical on a GPU the sequential sums are reproduced EXACTLY:
 *
 *   All doubles inside one binade [2^e,2^(e+1)) are multiples of u=2^(e-52).
 *   fl(x+s) for x in that binade (and the result still in it) equals x plus s
 *   rounded to a multiple of u; the rounding is the same at every step except
 *   that an exact tie (s mod u == u/2) is resolved by the parity of x/u.  After
 *   ONE in-binade step the parity is fixed (the result of a tie is even and the
 *   increment keeps it even), so from the second in-binade step on the
 *   increment d is constant and x_(k+j) = x_k + j*d holds EXACTLY (all terms are
 *   multiples of u below 2^(e+1), hence representable; j*d and the sum are
 *   exact in double arithmetic).
 *
 * nco_walk therefore performs real additions only around binade changes /
 * wraps and jumps over the interior of each binade with an EXACT run length
 * (nco_jump), emitting a short table of linear segments (about two per binade
 * crossed: ~30 entries for 16368 samples).  Any thread can
 * then evaluate x_k for its own k from the table.  Verified against the plain
 * sequential loops over millions of random cases in tests/test_nco_exact.py.
 */
#ifndef SDRB200_NCO_EXACT_H
#define SDRB200_NCO_EXACT_H

#include <stdint.h>
#include <string.h>
#include <math.h>

#ifdef __CUDACC__
#define NCO_HD __host__ __device__ __forceinline__
#define NCO_ADD(a, b) nco_add_rn((a), (b))
#define NCO_MUL(a, b) nco_mul_rn((a), (b))
#define NCO_DIV(a, b) nco_div_rn((a), (b))
#else
#define NCO_HD static inline
#define NCO_ADD(a, b) ((a) + (b))
#define NCO_MUL(a, b) ((a) * (b))
#define NCO_DIV(a, b) ((a) / (b))
#endif

#ifdef __CUDACC__
/* explicit round-to-nearest ops: never contracted into FMA by nvcc */
__host__ __device__ __forceinline__ double nco_add_rn(double a, double b)
{
#ifdef __CUDA_ARCH__
    return __dadd_rn(a, b);
#else
    volatile double r = a + b; return r;
#endif
}
__host__ __device__ __forceinline__ double nco_mul_rn(double a, double b)
{
#ifdef __CUDA_ARCH__
    return __dmul_rn(a, b);
#else
    volatile double r = a * b; return r;
#endif
}
__host__ __device__ __forceinline__ double nco_div_rn(double a, double b)
{
#ifdef __CUDA_ARCH__
    return __ddiv_rn(a, b);
#else
    volatile double r = a / b; return r;
#endif
}
#endif

#define NCO_MAXSEG 192
#define NCO_TWO_PI 6.2831853071795864      /* DPI = 2.0*3.1415926535897932, src/sdr.h:103-104 */

/* values x_k for k in [k0, next.k0): x_k = x0 + (k-k0)*d, exactly */
typedef struct {
    int32_t k0;
    int32_t pad;
    double  x0;
    double  d;
} nco_seg_t;

typedef struct {
    int32_t   nseg;
    int32_t   count;        /* number of values described                     */
    double    x_end;        /* accumulator after the last addition            */
    nco_seg_t seg[NCO_MAXSEG];
} nco_table_t;

NCO_HD uint64_t nco_bits(double x)
{
#ifdef __CUDA_ARCH__
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t b; memcpy(&b, &x, sizeof(b)); return b;
#endif
}
NCO_HD double nco_from_bits(uint64_t b)
{
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    double x; memcpy(&x, &b, sizeof(x)); return x;
#endif
}
/* same sign and same exponent (and not zero) */
NCO_HD int nco_same_binade(double a, double b)
{
    uint64_t ba = nco_bits(a), bb = nco_bits(b);
    return ((ba >> 52) == (bb >> 52)) && ((ba << 1) != 0) && ((bb << 1) != 0);
}
/* |lower bound| of the binade holding x (x != 0): mantissa cleared */
NCO_HD double nco_binade_lo(double x)
{
    return nco_from_bits(nco_bits(x) & 0x7FF0000000000000ULL);
}

/* all values of one run share the binade of its first value and stay on the
 * permitted side of `limit` (limit_on 0: none, 1: value < limit, 2: value > limit) */
NCO_HD int nco_in_run(double x1, double y, int limit_on, double limit)
{
    if (!nco_same_binade(x1, y)) return 0;
    if (limit_on == 1 && y >= limit) return 0;
    if (limit_on == 2 && !(y > limit)) return 0;
    return 1;
}

/* EXACT run length: the largest J in [0, maxj] such that x1 + j*d stays in the
 * run for every j <= J (d != 0 keeps the sequence monotone).  An estimate
 * room * inv_d (inv_d: approximate reciprocal of |d|, computed once per walk) is
 * corrected with exact double arithmetic: inside one
 * binade j*d and x1 + j*d are exact, and a sum that leaves the binade compares
 * on the correct side of its border after rounding.  *y = x1 + J*d.             */
NCO_HD int32_t nco_jump(double x1, double d, double inv_d, int limit_on, double limit, int32_t maxj, double *y)
{
    double lo = nco_binade_lo(x1), hi = lo + lo, m = fabs(x1), room, yy, y2;
    int growing = ((x1 > 0.0) == (d > 0.0));
    int32_t J;
    double est;
    if (d == 0.0 || maxj <= 0) { *y = x1; return maxj > 0 ? maxj : 0; }
    room = growing ? (hi - m) : (m - lo);
    if (limit_on == 1 && d > 0.0) { double r2 = limit - x1; if (r2 < room) room = r2; }
    if (limit_on == 2 && d < 0.0) { double r2 = x1 - limit; if (r2 < room) room = r2; }
    est = NCO_MUL(room, inv_d);         /* inv_d ~ 1/|d| (any accuracy: J is corrected below) */
    if (!(est > 0.0)) J = 0;
    else if (est >= (double)maxj) J = maxj;
    else J = (int32_t)est;
    yy = NCO_ADD(x1, NCO_MUL((double)J, d));
    while (J > 0 && !nco_in_run(x1, yy, limit_on, limit)) { J--; yy = NCO_ADD(x1, NCO_MUL((double)J, d)); }
    while (J < maxj) {
        y2 = NCO_ADD(yy, d);
        if (!nco_in_run(x1, y2, limit_on, limit)) break;
        yy = y2; J++;
    }
    *y = yy;
    return J;
}

/* approximate 1/|x| (single-precision accuracy is plenty for nco_jump) */
NCO_HD double nco_rcp_est(double x)
{
    return (double)(1.0f / (float)fabs(x));
}

#define NCO_EMIT(K0, X0, DD) do { if (ns < NCO_MAXSEG) { nco_seg_t *sg_ = &t->seg[ns]; sg_->k0 = (K0); sg_->pad = 0; \
                                   sg_->x0 = (X0); sg_->d = (DD); } ns++; } while (0)

/* Sequence   for k in 0..count-1: { if (wrap>0 && x>=wrap) x-=wrap; use x; x+=step; }
 * (wrap = 0: no wrap test, the carrier NCO; wrap = code length: the code NCO).
 * Fills t with segments describing every USED value; t->x_end = x after the
 * last addition.  Returns 0, or -1 if the table overflowed.                    */
NCO_HD int nco_walk(double x, double step, int32_t count, double wrap, nco_table_t *t)
{
    int32_t k = 0, ns = 0;
    const int lim = (wrap > 0.0) ? 1 : 0;
    const double inv = nco_rcp_est(step);
    while (k < count) {
        double x1, x2, d;
        int32_t J;
        if (lim && x >= wrap) x = NCO_ADD(x, -wrap);
        x1 = NCO_ADD(x, step);
        if (x1 == x) {                      /* step is zero or below half an ulp: */
            NCO_EMIT(k, x, 0.0);            /* the accumulator never moves again  */
            k = count;
            break;
        }
        if (k + 1 >= count || (lim && x1 >= wrap) || !nco_same_binade(x, x1)) {
            NCO_EMIT(k, x, 0.0); k += 1; x = x1; continue;
        }
        x2 = NCO_ADD(x1, step);
        if (k + 2 >= count || (lim && x2 >= wrap) || !nco_same_binade(x1, x2)) {
            NCO_EMIT(k, x, NCO_ADD(x1, -x)); k += 2; x = x2; continue;
        }
        d = NCO_ADD(x2, -x1);
        /* steady state from x1: values x1 + j*d, j = 0..J, J >= 1 (x2 is j = 1),
         * J exact: the addition after the run leaves the binade (or wraps)      */
        J = nco_jump(x1, d, inv, lim, wrap, count - 2 - k, &x2);
        NCO_EMIT(k, x, 0.0);
        NCO_EMIT(k + 1, x1, d);
        x = NCO_ADD(x2, step);
        k += 2 + J;
    }
    t->x_end = x;
    t->count = count;
    t->nseg = ns;
    return (ns > NCO_MAXSEG) ? -1 : 0;
}

/* value used at index k (0 <= k < count); seg hint speeds up monotone scans */
NCO_HD double nco_value(const nco_table_t *t, int32_t k, int32_t *hint)
{
    int32_t s = hint ? *hint : 0;
    if (s >= t->nseg || t->seg[s].k0 > k) s = 0;
    while (s + 1 < t->nseg && t->seg[s + 1].k0 <= k) s++;
    if (hint) *hint = s;
    return NCO_ADD(t->seg[s].x0, NCO_MUL((double)(k - t->seg[s].k0), t->seg[s].d));
}


/* ---- integer form of a segment ------------------------------------------------
 * All values of a segment lie in ONE binade [2^e, 2^(e+1)) and are multiples of
 * u = 2^(e-52): value_j = (X0 + j*D) * u with X0 < 2^53 and X0 + j*D < 2^53.
 * The C cast (int)value, which the reference applies to every NCO value
 * (src/sdrcmn.c:615,652,659), is therefore sign * ((X0 + j*D) >> (52 - e)):
 * two integer multiply-adds and a shift per sample instead of FP64 arithmetic. */
typedef struct {
    int32_t  k0;
    int32_t  shift;      /* 52 - e, clamped to 64 (|value| < 1 -> 0)            */
    uint64_t X0;
    int64_t  D;          /* signed change of the MAGNITUDE per step             */
    int32_t  neg;        /* values are negative                                 */
    int32_t  pad;
} nco_iseg_t;

NCO_HD void nco_iseg_make(const nco_seg_t *s, nco_iseg_t *o)
{
    const uint64_t bx = nco_bits(s->x0), bd = nco_bits(s->d);
    const int ex = (int)((bx >> 52) & 0x7FF), ed = (int)((bd >> 52) & 0x7FF);
    o->k0 = s->k0; o->pad = 0;
    o->neg = (int32_t)(bx >> 63);
    if (ex == 0) { o->X0 = 0; o->D = 0; o->shift = 64; return; }      /* zero / subnormal: (int) gives 0 */
    o->X0 = (bx & 0x000FFFFFFFFFFFFFULL) | 0x0010000000000000ULL;
    {
        const int sh = 52 - (ex - 1023);
        o->shift = sh > 64 ? 64 : (sh < 0 ? -1 : sh);                   /* -1: |value| >= 2^53, unsupported */
    }
    if (ed == 0 || (bd << 1) == 0) { o->D = 0; return; }
    {
        const uint64_t md = (bd & 0x000FFFFFFFFFFFFFULL) | 0x0010000000000000ULL;
        const int dsh = ex - ed;                                        /* >= 0: |d| < 2^(e+1) */
        const uint64_t mag = (dsh < 0) ? (md << 1) : (dsh > 63 ? 0 : (md >> dsh));
        const int grows = ((bx >> 63) == (bd >> 63));
        o->D = grows ? (int64_t)mag : -(int64_t)mag;
    }
}

/* (int)value at index k of the segment (C truncation toward zero) */
NCO_HD int64_t nco_iseg_trunc(const nco_iseg_t *s, int32_t k)
{
    const uint64_t m = s->X0 + (uint64_t)((int64_t)(k - s->k0) * s->D);
    const int64_t t = (s->shift >= 64) ? 0 : (int64_t)(m >> s->shift);
    return s->neg ? -t : t;
}

/* ---- integer walker ----------------------------------------------------------
 * Same sequence as nco_walk, but the interior of a binade is handled on the 53-bit
 * mantissas: with M(x) = |x| / ulp(binade), a steady run is M1 + j*D with the integer
 * D = M(x2) - M(x1) taken from two REAL additions (so the rounding of the increment,
 * including an exact tie, is whatever the hardware did), the run length is one
 * integer division (float estimate + exact correction) and the step that leaves
 * the binade is again a real addition.  Segments come out directly in integer form. */
#define NCO_MANT(b) (((b) & 0x000FFFFFFFFFFFFFULL) | 0x0010000000000000ULL)

NCO_HD uint32_t nco_udiv_est(uint64_t num, uint64_t den, uint32_t cap)
{
    /* floor(num / den) clamped to cap; den > 0 */
#ifdef __CUDA_ARCH__
    float q = __fdividef(__ull2float_rz(num), __ull2float_ru(den));
#else
    float q = (float)num / (float)den;
#endif
    uint64_t J;
    int64_t r;
    if (q >= (float)cap) {
        /* cap * den may not fit: compare through the quotient side */
        if (num / den >= cap) return cap;               /* rare: exact division */
        J = num / den;
        return (uint32_t)J;
    }
    J = (uint64_t)q;
    r = (int64_t)(num - J * den);
    while (r < 0) { J--; r += (int64_t)den; }
    while (r >= (int64_t)den) { J++; r -= (int64_t)den; }
    return J > cap ? cap : (uint32_t)J;
}

/* further steps a steady run can take from mantissa M1 (exponent field ex, sign of
 * the values irrelevant) with signed mantissa increment D before it leaves its binade
 * or violates the limit (limit_on 1: value < limit for growing positive values,
 * 2: value > limit for shrinking positive values), clamped to maxj                */
NCO_HD int32_t nco_ijump(uint64_t M1, int64_t D, int ex, int positive, int limit_on, double limit, int32_t maxj)
{
    uint64_t num;
    if (maxj <= 0) return 0;
    if (D == 0) return maxj;
    if (D > 0) {
        uint64_t Xmax = 0x001FFFFFFFFFFFFFULL;
        if (limit_on == 1 && positive) {
            const uint64_t bl = nco_bits(limit);
            if ((int)((bl >> 52) & 0x7FF) == ex) { const uint64_t ml = NCO_MANT(bl) - 1; if (ml < Xmax) Xmax = ml; }
        }
        num = Xmax - M1;
        return (int32_t)nco_udiv_est(num, (uint64_t)D, (uint32_t)maxj);
    } else {
        uint64_t Xmin = 0x0010000000000000ULL;
        if (limit_on == 2 && positive) {
            const uint64_t bl = nco_bits(limit);
            if ((int)((bl >> 52) & 0x7FF) == ex) { const uint64_t ml = NCO_MANT(bl) + 1; if (ml > Xmin) Xmin = ml; }
        }
        num = M1 - Xmin;
        return (int32_t)nco_udiv_est(num, (uint64_t)(-D), (uint32_t)maxj);
    }
}

NCO_HD void nco_iseg_set(nco_iseg_t *o, int32_t k0, uint64_t bits, int64_t D)
{
    const int ex = (int)((bits >> 52) & 0x7FF);
    o->k0 = k0; o->pad = 0;
    o->neg = (int32_t)(bits >> 63);
    if (ex == 0) { o->X0 = 0; o->D = 0; o->shift = 64; return; }       /* zero / subnormal: (int) gives 0 */
    o->X0 = NCO_MANT(bits);
    o->D = D;
    {
        const int sh = 52 - (ex - 1023);
        o->shift = sh > 64 ? 64 : (sh < 0 ? -1 : sh);
    }
}

/* integer-segment version of nco_walk: returns the number of segments written to
 * seg[0..maxseg), -1 on overflow, -2 if a value is too large to truncate;
 * *x_end = accumulator after the last addition                                    */
NCO_HD int nco_walk_i(double x, double step, int32_t count, double wrap, nco_iseg_t *seg, int maxseg, double *x_end)
{
    int32_t k = 0, ns = 0, range_bad = 0;
    const int lim = (wrap > 0.0) ? 1 : 0;
#define NCO_IEMIT(K0, BITS, DD) do { if (ns < maxseg) { nco_iseg_set(&seg[ns], (K0), (BITS), (DD)); if (seg[ns].shift < 0) range_bad = 1; } ns++; } while (0)
    /* constants of the fast chain (positive x growing by a positive normal step) */
    const uint64_t bstep = nco_bits(step), bwrap = nco_bits(wrap);
    const int ed = (int)((bstep >> 52) & 0x7FF), ewrap = (int)((bwrap >> 52) & 0x7FF);
    const uint64_t Dm = NCO_MANT(bstep);
    const int chain_ok = (step > 0.0) && ed > 0 && ed < 0x7FF;
    while (k < count) {
        double x1, x2;
        uint64_t b0, b1, b2;
        int64_t D0, D;
        int32_t J;
        if (lim && x >= wrap) x = NCO_ADD(x, -wrap);
        /* ---- fast chain: whole binades on the mantissa, no floating point.  For x in
         * binade ex = ed + s (s >= 1) every addition rounds step to the same multiple
         * D of ulp(x) unless step's low s bits are an exact tie; the addition that
         * leaves the binade rounds (X + step/ulp) / 2 to nearest even.                */
        while (chain_ok) {
            const uint64_t bx = nco_bits(x);
            const int ex = (int)(bx >> 52);                      /* sign bit set -> large -> s out of range */
            const int sft = ex - ed;
            uint64_t X, q, r, half, Dn, Xmax, num, rr;
            uint32_t Jn;
            float est;
            if (sft < 1 || sft > 52) break;
            X = NCO_MANT(bx);
            q = Dm >> sft; r = Dm & ((1ULL << sft) - 1); half = 1ULL << (sft - 1);
            if (r == half) break;                               /* tie: the generic path sorts out the parity */
            Dn = q + (r > half ? 1 : 0);
            Xmax = 0x001FFFFFFFFFFFFFULL;
            if (lim && ex == ewrap) Xmax = NCO_MANT(bwrap) - 1;
            if (X > Xmax) break;
            num = Xmax - X;
#ifdef __CUDA_ARCH__
            est = __fdividef(__ull2float_rz(num), __ull2float_ru(Dn));
#else
            est = (float)num / (float)Dn;
#endif
            if (!(est < (float)(count - k - 3))) break;         /* the run reaches the end: generic tail */
            Jn = (uint32_t)est;
            rr = num - (uint64_t)Jn * Dn;
            while ((int64_t)rr < 0) { Jn--; rr += Dn; }
            while (rr >= Dn) { Jn++; rr -= Dn; }
            if ((int32_t)Jn >= count - k - 3) break;
            NCO_IEMIT(k, bx, (int64_t)Dn);
            k += (int32_t)Jn + 1;
            {
                const uint64_t XJ = Xmax - rr;                   /* X + Jn*Dn */
                if (Xmax != 0x001FFFFFFFFFFFFFULL) {
                    /* stopped by the wrap limit: one real addition, then the wrap test above */
                    x = NCO_ADD(nco_from_bits((bx & 0xFFF0000000000000ULL) | (XJ & 0x000FFFFFFFFFFFFFULL)), step);
                    if (lim && x >= wrap) x = NCO_ADD(x, -wrap);
                } else {
                    const uint64_t T = XJ + q, Th = T >> 1;
                    const uint64_t Xn = Th + ((T & 1) & ((r != 0 ? 1 : 0) | (Th & 1)));
                    x = nco_from_bits(((uint64_t)(ex + 1) << 52) | (Xn & 0x000FFFFFFFFFFFFFULL));
                }
            }
        }
        if (k >= count) break;
        x1 = NCO_ADD(x, step);
        b0 = nco_bits(x);
        if (x1 == x) { NCO_IEMIT(k, b0, 0); k = count; break; }
        if (k + 1 >= count || (lim && x1 >= wrap) || !nco_same_binade(x, x1)) {
            NCO_IEMIT(k, b0, 0); k += 1; x = x1; continue;
        }
        b1 = nco_bits(x1);
        D0 = (int64_t)NCO_MANT(b1) - (int64_t)NCO_MANT(b0);
        x2 = NCO_ADD(x1, step);
        if (k + 2 >= count || (lim && x2 >= wrap) || !nco_same_binade(x1, x2)) {
            NCO_IEMIT(k, b0, D0); k += 2; x = x2; continue;
        }
        b2 = nco_bits(x2);
        D = (int64_t)NCO_MANT(b2) - (int64_t)NCO_MANT(b1);
        /* steady state from x1: mantissas M1 + j*D, j = 0..J, J >= 1 (x2 is j = 1) */
        J = nco_ijump(NCO_MANT(b1), D, (int)((b1 >> 52) & 0x7FF), x1 > 0.0, lim, wrap, count - 2 - k);
        if (D0 == D) NCO_IEMIT(k, b0, D);
        else { NCO_IEMIT(k, b0, 0); NCO_IEMIT(k + 1, b1, D); }
        {
            const uint64_t MJ = NCO_MANT(b1) + (uint64_t)((int64_t)J * D);
            const double xJ = nco_from_bits((b1 & 0xFFF0000000000000ULL) | (MJ & 0x000FFFFFFFFFFFFFULL));
            x = NCO_ADD(xJ, step);
        }
        k += 2 + J;
    }
#undef NCO_IEMIT
    *x_end = x;
    if (range_bad) return -2;                  /* |value| >= 2^53: no integer index */
    return (ns > maxseg) ? -1 : ns;
}

/* while (x > 2pi) x -= 2pi;   (src/sdrcmn.c:666-668), exactly.
 *
 * Fast path: P = 2pi is a constant, so for x in binade 2^(2+s) (s >= 2, i.e. x >= 16) the
 * subtraction is known in units of ulp(x) without probing: P / ulp = MP / 2^s = q + r/2^s
 * (MP: 53-bit mantissa of P).  While the exact difference stays at or above the binade's
 * lower border it is rounded on the same grid, so every step subtracts D = q (r < half) or
 * q + 1 (r > half) from the mantissa X.  J = floor((X - 2^52 - 1) / D) such steps keep
 * X - J*D >= 2^52 + 1, for which the exact difference (> X_j - 1/2) is provably inside
 * the binade; one integer division replaces them.  The step that may leave the binade is a
 * REAL subtraction.  An exact tie (r == half, binade [64,128) for this P) takes one real
 * step first: its result is even, after which D is the even one of {q, q+1}.
 * Everything else (x < 16, huge x) runs the generic probing loop.                        */
/* per-binade constants of the fast path for P = 2*pi (mantissa 0x1921FB54442D18): entry s is
 * D = P / ulp rounded for x in [2^(2+s), 2^(3+s)) - q + (r > half), or the even one of
 * {q, q+1} for the tie binade - and a float reciprocal of D / 2^21 (biased low by 2^-22 so
 * that the run-length estimate rarely overshoots).  tests/test_nco_exact.py recomputes them. */
#define NCO_R2PI_TIES 0x10U   /* binades s whose rounding is an exact tie */
static const uint64_t nco_r2pi_D_host[32] = {
    0x001921FB54442D18ULL,
    0x000C90FDAA22168CULL,
    0x0006487ED5110B46ULL,
    0x0003243F6A8885A3ULL,
    0x0001921FB54442D2ULL,
    0x0000C90FDAA22169ULL,
    0x00006487ED5110B4ULL,
    0x00003243F6A8885AULL,
    0x00001921FB54442DULL,
    0x00000C90FDAA2217ULL,
    0x000006487ED5110BULL,
    0x000003243F6A8886ULL,
    0x000001921FB54443ULL,
    0x000000C90FDAA221ULL,
    0x0000006487ED5111ULL,
    0x0000003243F6A888ULL,
    0x0000001921FB5444ULL,
    0x0000000C90FDAA22ULL,
    0x00000006487ED511ULL,
    0x00000003243F6A89ULL,
    0x00000001921FB544ULL,
    0x00000000C90FDAA2ULL,
    0x000000006487ED51ULL,
    0x000000003243F6A9ULL,
    0x000000001921FB54ULL,
    0x000000000C90FDAAULL,
    0x0000000006487ED5ULL,
    0x0000000003243F6BULL,
    0x0000000001921FB5ULL,
    0x0000000000C90FDBULL,
    0x00000000006487EDULL,
    0x00000000003243F7ULL
};
static const float nco_r2pi_rcp_host[32] = {
    2.964491708e-10f,
    5.928982305e-10f,
    1.185796461e-09f,
    2.371592922e-09f,
    4.743185844e-09f,
    9.486371688e-09f,
    1.897274338e-08f,
    3.794548675e-08f,
    7.589097351e-08f,
    1.517819470e-07f,
    3.035638940e-07f,
    6.071277880e-07f,
    1.214255576e-06f,
    2.428511152e-06f,
    4.857022304e-06f,
    9.714044609e-06f,
    1.942808922e-05f,
    3.885617843e-05f,
    7.771235687e-05f,
    1.554247137e-04f,
    3.108494275e-04f,
    6.216988550e-04f,
    1.243397710e-03f,
    2.486795420e-03f,
    4.973590840e-03f,
    9.947181679e-03f,
    1.989436336e-02f,
    3.978872672e-02f,
    7.957745343e-02f,
    1.591549069e-01f,
    3.183098137e-01f,
    6.366195679e-01f
};
#ifdef __CUDACC__
static __device__ __constant__ uint64_t nco_r2pi_D_dev[32] = {
    0x001921FB54442D18ULL,
    0x000C90FDAA22168CULL,
    0x0006487ED5110B46ULL,
    0x0003243F6A8885A3ULL,
    0x0001921FB54442D2ULL,
    0x0000C90FDAA22169ULL,
    0x00006487ED5110B4ULL,
    0x00003243F6A8885AULL,
    0x00001921FB54442DULL,
    0x00000C90FDAA2217ULL,
    0x000006487ED5110BULL,
    0x000003243F6A8886ULL,
    0x000001921FB54443ULL,
    0x000000C90FDAA221ULL,
    0x0000006487ED5111ULL,
    0x0000003243F6A888ULL,
    0x0000001921FB5444ULL,
    0x0000000C90FDAA22ULL,
    0x00000006487ED511ULL,
    0x00000003243F6A89ULL,
    0x00000001921FB544ULL,
    0x00000000C90FDAA2ULL,
    0x000000006487ED51ULL,
    0x000000003243F6A9ULL,
    0x000000001921FB54ULL,
    0x000000000C90FDAAULL,
    0x0000000006487ED5ULL,
    0x0000000003243F6BULL,
    0x0000000001921FB5ULL,
    0x0000000000C90FDBULL,
    0x00000000006487EDULL,
    0x00000000003243F7ULL
};
static __device__ __constant__ float nco_r2pi_rcp_dev[32] = {
    2.964491708e-10f,
    5.928982305e-10f,
    1.185796461e-09f,
    2.371592922e-09f,
    4.743185844e-09f,
    9.486371688e-09f,
    1.897274338e-08f,
    3.794548675e-08f,
    7.589097351e-08f,
    1.517819470e-07f,
    3.035638940e-07f,
    6.071277880e-07f,
    1.214255576e-06f,
    2.428511152e-06f,
    4.857022304e-06f,
    9.714044609e-06f,
    1.942808922e-05f,
    3.885617843e-05f,
    7.771235687e-05f,
    1.554247137e-04f,
    3.108494275e-04f,
    6.216988550e-04f,
    1.243397710e-03f,
    2.486795420e-03f,
    4.973590840e-03f,
    9.947181679e-03f,
    1.989436336e-02f,
    3.978872672e-02f,
    7.957745343e-02f,
    1.591549069e-01f,
    3.183098137e-01f,
    6.366195679e-01f
};
#endif

NCO_HD double nco_reduce_2pi(double x)
{
    const double P = NCO_TWO_PI;
    while (x > P) {
        double x1, x2;
        uint64_t b1, b2;
        int64_t D;
        int32_t J;
        {
            const uint64_t bx = nco_bits(x);
            const int s = (int)(bx >> 52) - (1023 + 2);             /* x > P > 0: sign bit clear */
            if ((unsigned int)(s - 2) <= 28u) {
#ifdef __CUDA_ARCH__
                const uint64_t Dq = nco_r2pi_D_dev[s]; const float rcp = nco_r2pi_rcp_dev[s];
#else
                const uint64_t Dq = nco_r2pi_D_host[s]; const float rcp = nco_r2pi_rcp_host[s];
#endif
                uint64_t X = NCO_MANT(bx);
                if ((NCO_R2PI_TIES >> s) & 1u) {
                    x1 = NCO_ADD(x, -P);
                    if (!nco_same_binade(x, x1)) { x = x1; continue; }
                    X = NCO_MANT(nco_bits(x1));                     /* even after the tie */
                }
                if (X > 0x0010000000000001ULL) {
                    /* J = floor(num / Dq): float estimate from the top 31 bits, exact fix-up */
                    const uint64_t num = X - 0x0010000000000001ULL;
                    const uint32_t Jn = (uint32_t)((float)(uint32_t)(num >> 21) * rcp);
                    int64_t rr = (int64_t)(num - (uint64_t)Jn * Dq);
                    /* bounded corrections (loops here are turned into a 64-bit division by the
                     * compiler); an estimate further off takes the exact remainder          */
                    if (rr < 0) rr += (int64_t)Dq;
                    if (rr >= (int64_t)Dq) rr -= (int64_t)Dq;
                    if (rr >= (int64_t)Dq) rr -= (int64_t)Dq;
                    if (rr < 0 || rr >= (int64_t)Dq) rr = (int64_t)(num % Dq);
                    X = 0x0010000000000001ULL + (uint64_t)rr;
                }
                x = NCO_ADD(nco_from_bits((bx & 0xFFF0000000000000ULL) | (X & 0x000FFFFFFFFFFFFFULL)), -P);
                continue;
            }
        }
        x1 = NCO_ADD(x, -P);
        if (!(x1 > P) || !nco_same_binade(x, x1)) { x = x1; continue; }
        x2 = NCO_ADD(x1, -P);
        if (!(x2 > P) || !nco_same_binade(x1, x2)) { x = x2; continue; }
        b1 = nco_bits(x1); b2 = nco_bits(x2);
        D = (int64_t)NCO_MANT(b2) - (int64_t)NCO_MANT(b1);            /* < 0 */
        J = nco_ijump(NCO_MANT(b1), D, (int)((b1 >> 52) & 0x7FF), 1, 2, P, 0x7FFFFFF0);
        {
            const uint64_t MJ = NCO_MANT(b1) + (uint64_t)((int64_t)J * D);
            x = nco_from_bits((b1 & 0xFFF0000000000000ULL) | (MJ & 0x000FFFFFFFFFFFFFULL));
        }
    }
    return x;
}

/* ==== lean double-domain walker (the tracking kernel's NCO lanes) ================
 * Same sequences as nco_walk / nco_reduce_2pi (nco_walk_l, nco_reduce_2pi_d below), arranged so
 * that the DEPENDENT chain per binade is ~8 floating-point operations and nothing else:
 *
 *   inside binade [2^e, 2^(e+1)) every addition of `step` adds d_e = step rounded to the
 *   binade's grid (see the head of this file); d_e depends on step and e only, so a table
 *   of d_e (and of 1/d_e, biased low) is built ONCE per step value - in parallel, one lane
 *   per binade, off the chain (nco_steptab_entry).  An exact tie (step's dropped bits are
 *   exactly half a grid unit) is resolved by the parity of x: one REAL addition makes the
 *   mantissa even, after which the increment is the even one of {q, q+1} grid units.
 *   The run length is floor(room * inv) with one exact upward correction; the addition that
 *   leaves the binade (or reaches the wrap limit) is a REAL addition, so its rounding is
 *   the hardware's.  Values outside the table's domain (x <= 0, step <= 0, x below step's
 *   binade, subnormals) take the probing iteration of nco_walk.
 * Segments come out in the integer form the correlator threads use (nco_iseg_t).          */
#define NCO_STEPTAB_N 54
typedef struct {
    double   di[NCO_STEPTAB_N][2];  /* [s] = {d, inv}: increment in binade ed + s (s = 1..53; 53: |step| below
                                     * one ulp) and (1 / d) * (1 - 2^-30); d = 0: entry unusable             */
    uint64_t Dn[NCO_STEPTAB_N];     /* [s]: d in units of the binade's grid                                  */
    uint64_t ties;                  /* bit s: binade ed + s rounds |step| on an exact tie                    */
    int32_t  ed;                    /* biased exponent of step; 0: table unusable                            */
    int32_t  pad;
} nco_steptab_t;

/* entry s (1..53) of the table for `step` (any sign, normal): returns 0 if binade ed + s has no
 * normal grid representation of the increment (never for receiver frequencies)              */
NCO_HD int nco_steptab_entry(double step, int s, double *d, double *inv, int *tie, uint64_t *Dn_out)
{
    const uint64_t bstep = nco_bits(step);
    const int ed = (int)((bstep >> 52) & 0x7FF);
    const uint64_t Dm = NCO_MANT(bstep);
    const uint64_t q = Dm >> s, r = Dm & ((1ULL << s) - 1), half = 1ULL << (s - 1);
    uint64_t Dn;
    const int eg = ed + s - 52;                         /* biased exponent of one grid unit */
    *tie = (r == half);
    if (r == half) Dn = q + (q & 1);                    /* the even one of {q, q+1}         */
    else Dn = q + (r > half ? 1 : 0);
    *Dn_out = Dn;
    if (eg < 1 || eg > 2046 || ed + s > 2046 || Dn == 0) { *d = 0.0; *inv = 0.0; return 0; }
    *d = NCO_MUL((double)(int64_t)Dn, nco_from_bits((uint64_t)eg << 52));      /* exact   */
    *inv = NCO_MUL(NCO_DIV(1.0, *d), 0x1.fffffffcp-1);                          /* 1 - 2^-30 */
    return 1;
}
NCO_HD void nco_steptab_build(double step, nco_steptab_t *t)
{
    const uint64_t bstep = nco_bits(step);
    const int ed = (int)((bstep >> 52) & 0x7FF);
    int s;
    t->ties = 0; t->pad = 0;
    t->ed = (step != 0.0 && ed > 0 && ed < 0x7FF) ? ed : 0;       /* entries depend on |step| only */
    t->di[0][0] = 0.0; t->di[0][1] = 0.0; t->Dn[0] = 0;
    for (s = 1; s <= 53; s++) {
        int tie = 0;
        t->Dn[s] = 0;
        if (!t->ed || !nco_steptab_entry(step, s, &t->di[s][0], &t->di[s][1], &tie, &t->Dn[s])) { t->di[s][0] = 0.0; t->di[s][1] = 0.0; }
        if (tie) t->ties |= 1ULL << s;
    }
}

/* floor(v) for 0 <= v < 2^51 and v*w + c, on the fixed-latency FP64 pipe of the device */
#ifdef __CUDA_ARCH__
#define NCO_FLOOR51(v) (__dadd_rn(__dadd_rd((v), 4503599627370496.0), -4503599627370496.0))
#define NCO_FMA(a, b, c) __fma_rn((a), (b), (c))
#else
#define NCO_FLOOR51(v) floor(v)
#define NCO_FMA(a, b, c) ((a) * (b) + (c))          /* product and sum are exact where it is used */
#endif

/* The walk in two stages.
 * Stage 1, nco_walk_raw - the serial chain, kept to the bare recurrence: one RAW record per run
 *   {bits of the run's first value, index of that value (as a double), kind}, kind 0 = a single
 *   value, 1 = |x| growing by the table increment of its binade, 2 = shrinking by it.  It works on
 *   the magnitude m = |x| so that one table (built for |step|) serves both signs:
 *     growing   (x and step of one sign): values m + j*d stay on the binade's grid while < top
 *               (top = next binade border, or the wrap limit inside its binade);
 *     shrinking (opposite signs): values m - j*d stay on the grid while >= lo + ulp (then the
 *               exact difference is still inside the binade);
 *     J = floor(room * inv) with inv biased low by 2^-30: never too large, at most one too
 *     small for runs shorter than 2^29 -> ONE upward correction, no loop;
 *     the addition that leaves the run is a real one.  An exact tie takes one real step while
 *     the mantissa is odd.  Everything else (|x| below the table, zero, subnormal, huge) is
 *     stepped one real addition at a time.  The run that reaches `count` is cut afterwards, so
 *     the loop carries no clamp.
 * Stage 2, nco_raw_to_iseg - one record -> the integer segment the correlator threads use;
 *   independent per record (the tracking kernel converts with one lane per record).          */
typedef struct {
    uint64_t bx;            /* bits of the first value of the run */
    double   kf;            /* its index                          */
    int32_t  kind, pad;
} nco_raw_t;

#define NCO_RAW_EMIT(BX, KF, KIND) do { if (ns < maxseg) { raw[ns].bx = (BX); raw[ns].kf = (KF); raw[ns].kind = (KIND); } ns++; } while (0)

/* The common case - x > 0 growing by step > 0 (the carrier at IF, every code NCO) - as a tight
 * loop: straight-line arithmetic, three branches per binade.  Leaves at the first value that
 * needs anything else (|x| below the table, a tie with an odd mantissa, ...), or when the
 * walk is complete (returns 1; *px is then the accumulator after the last addition).  `lim`
 * is a literal at both call sites.                                                         */
NCO_HD int nco_walk_tight(double *px, double *pKf, int32_t *pns, const double step, const double Cf, const double wrap,
                          const int lim, const int ed, const int ewrap, const uint64_t ties,
                          const nco_steptab_t *tab, nco_raw_t *raw, const int maxseg)
{
    double x = *px, Kf = *pKf;
    int32_t ns = *pns;
    int ended = 0;
    for (;;) {
        const uint64_t bx = nco_bits(x);
        const int ex = (int)(bx >> 52);                      /* x > 0 here */
        const int s = ex - ed;
        double d, inv, top, Jf, y, y2, Kn;
        if ((unsigned int)(s - 1) >= 53u || (lim && ex > ewrap)) break;
        d = tab->di[s][0]; inv = tab->di[s][1];
        if (d == 0.0 || (((ties >> s) & bx) & 1ULL)) break;
        top = (lim && ex == ewrap) ? wrap : nco_from_bits((uint64_t)(ex + 1) << 52);
        Jf = NCO_FLOOR51(NCO_MUL(NCO_ADD(top, -x), inv));
        y = NCO_FMA(Jf, d, x);
        y2 = NCO_ADD(y, d);
        Kn = NCO_ADD(Kf, NCO_ADD(Jf, 1.0));
        if (y2 < top) { y = y2; Kn = NCO_ADD(Kn, 1.0); }
        if (ns < maxseg) { raw[ns].bx = bx; raw[ns].kf = Kf; raw[ns].kind = 1; }
        ns++;
        if (Kn >= Cf) {                                      /* the walk ends inside (or with) this run */
            y = NCO_FMA(NCO_ADD(NCO_ADD(Cf, -Kf), -1.0), d, x);
            x = NCO_ADD(y, step);
            Kf = Cf; ended = 1;
            break;
        }
        Kf = Kn;
        x = NCO_ADD(y, step);                                /* leaves the run: a real addition */
        if (lim && x >= wrap) x = NCO_ADD(x, -wrap);
        if (!(x > 0.0)) break;
    }
    *px = x; *pKf = Kf; *pns = ns;
    return ended;
}

NCO_HD int nco_walk_raw(double x, double step, int32_t count, double wrap, const nco_steptab_t *tab,
                        nco_raw_t *raw, int maxseg, double *x_end)
{
    int32_t ns = 0;
    const int lim = (wrap > 0.0) ? 1 : 0;
    const int ed = tab->ed;                                  /* table of |step| */
    const int ewrap = (int)((nco_bits(wrap) >> 52) & 0x7FF);
    const uint64_t ties = tab->ties;
    const uint64_t sstep = nco_bits(step) >> 63;
    const double Cf = (double)count;
    double Kf = 0.0;
    while (Kf < Cf) {
        if (lim && x >= wrap) x = NCO_ADD(x, -wrap);
        if (ed && step > 0.0 && x > 0.0) {
            const int ended = lim ? nco_walk_tight(&x, &Kf, &ns, step, Cf, wrap, 1, ed, ewrap, ties, tab, raw, maxseg)
                                  : nco_walk_tight(&x, &Kf, &ns, step, Cf, wrap, 0, ed, ewrap, ties, tab, raw, maxseg);
            if (ended) break;
            /* x is the next value to use (wrap already applied): one general iteration handles it */
        }
        {
            const uint64_t bx = nco_bits(x);
            const uint64_t bm = bx & 0x7FFFFFFFFFFFFFFFULL;
            const int ex = (int)(bm >> 52);
            const int s = ex - ed;
            const uint64_t neg = bx >> 63;
            if (ed && (unsigned int)(s - 1) < 53u && !(lim && ex > ewrap)) {
                const double d = tab->di[s][0], inv = tab->di[s][1];
                const double m = nco_from_bits(bm);
                const int growing = (neg == sstep);
                if (d != 0.0 && !(((ties >> s) & 1ULL) && (bm & 1ULL))) {
                    double Jf, y, y2, Kn;
                    int up;
                    if (growing) {
                        const double top = (lim && ex == ewrap) ? wrap : nco_from_bits((uint64_t)(ex + 1) << 52);
                        Jf = NCO_FLOOR51(NCO_MUL(NCO_ADD(top, -m), inv));
                        y = NCO_FMA(Jf, d, m);
                        y2 = NCO_ADD(y, d);
                        up = (y2 < top);
                    } else {
                        const double lo1 = nco_from_bits(((uint64_t)ex << 52) | 1ULL);
                        if (!(m >= lo1)) goto single;        /* on the border: the real subtraction decides */
                        Jf = NCO_FLOOR51(NCO_MUL(NCO_ADD(m, -lo1), inv));
                        y = NCO_FMA(Jf, -d, m);
                        y2 = NCO_ADD(y, -d);
                        up = (y2 >= lo1);
                    }
                    if (up) { y = y2; Jf = NCO_ADD(Jf, 1.0); }
                    NCO_RAW_EMIT(bx, Kf, growing ? 1 : 2);
                    Kn = NCO_ADD(Kf, NCO_ADD(Jf, 1.0));
                    if (Kn > Cf) {                           /* the walk ends inside this run */
                        const double Jc = NCO_ADD(NCO_ADD(Cf, -Kf), -1.0);
                        y = NCO_FMA(Jc, growing ? d : -d, m);
                        Kn = Cf;
                    }
                    Kf = Kn;
                    x = NCO_ADD(neg ? -y : y, step);         /* leaves the run (or ends the walk): real */
                    continue;
                }
            }
        }
single:
        {
            const double x1 = NCO_ADD(x, step);
            NCO_RAW_EMIT(nco_bits(x), Kf, 0);
            if (x1 == x) { Kf = Cf; break; }                /* zero or sub-ulp step: never moves again */
            Kf = NCO_ADD(Kf, 1.0);
            x = x1;
        }
    }
    *x_end = x;
    return (ns > maxseg) ? -1 : ns;
}
#undef NCO_RAW_EMIT

/* returns 0, or 1 if a value is too large to truncate (|value| >= 2^53) */
NCO_HD int nco_raw_to_iseg(const nco_raw_t *r, const nco_steptab_t *tab, nco_iseg_t *o)
{
    const uint64_t bm = r->bx & 0x7FFFFFFFFFFFFFFFULL;
    const int ex = (int)(bm >> 52);
    if (r->kind == 0) { nco_iseg_set(o, (int32_t)r->kf, r->bx, 0); return o->shift < 0; }
    {
        const int sh = 1075 - ex;                            /* 52 - (ex - 1023) */
        const int64_t Dn = (int64_t)tab->Dn[ex - tab->ed];
        o->k0 = (int32_t)r->kf; o->pad = 0; o->neg = (int32_t)(r->bx >> 63); o->X0 = NCO_MANT(bm);
        o->D = (r->kind == 1) ? Dn : -Dn;
        o->shift = sh > 64 ? 64 : sh;
        if (sh < 0) { o->shift = -1; return 1; }
    }
    return 0;
}

/* both stages in sequence (contract of nco_walk_i): host tests, device self-test */
NCO_HD int nco_walk_l(double x, double step, int32_t count, double wrap, const nco_steptab_t *tab,
                      nco_iseg_t *seg, int maxseg, double *x_end)
{
    nco_raw_t raw[NCO_MAXSEG];
    int i, bad = 0;
    const int ns = nco_walk_raw(x, step, count, wrap, tab, raw, maxseg < NCO_MAXSEG ? maxseg : NCO_MAXSEG, x_end);
    if (ns < 0) return -1;
    for (i = 0; i < ns; i++) bad |= nco_raw_to_iseg(&raw[i], tab, &seg[i]);
    return bad ? -2 : ns;
}

/* 2*pi per binade [2^(2+s), 2^(3+s)): rounded to the binade's grid (the even neighbour for the
 * tie binade s = 4) and its reciprocal biased low by 2^-30; tests/test_nco_exact.py recomputes */
#define NCO_R2PI_TABLES \
    { 0x1.921fb54442d18p+2, 0x1.921fb54442d18p+2, 0x1.921fb54442d18p+2, 0x1.921fb54442d18p+2, \
      0x1.921fb54442d20p+2, 0x1.921fb54442d20p+2, 0x1.921fb54442d00p+2, 0x1.921fb54442d00p+2, \
      0x1.921fb54442d00p+2, 0x1.921fb54442e00p+2, 0x1.921fb54442c00p+2, 0x1.921fb54443000p+2, \
      0x1.921fb54443000p+2, 0x1.921fb54442000p+2, 0x1.921fb54444000p+2, 0x1.921fb54440000p+2, \
      0x1.921fb54440000p+2, 0x1.921fb54440000p+2, 0x1.921fb54440000p+2, 0x1.921fb54480000p+2, \
      0x1.921fb54400000p+2, 0x1.921fb54400000p+2, 0x1.921fb54400000p+2, 0x1.921fb54800000p+2, \
      0x1.921fb54000000p+2, 0x1.921fb54000000p+2, 0x1.921fb54000000p+2, 0x1.921fb58000000p+2, \
      0x1.921fb50000000p+2, 0x1.921fb60000000p+2, 0x1.921fb40000000p+2, 0x1.921fb80000000p+2 }
#define NCO_R2PI_INVS \
    { 0x1.45f306d784bc1p-3, 0x1.45f306d784bc1p-3, 0x1.45f306d784bc1p-3, 0x1.45f306d784bc1p-3, \
      0x1.45f306d784bbap-3, 0x1.45f306d784bbap-3, 0x1.45f306d784bd4p-3, 0x1.45f306d784bd4p-3, \
      0x1.45f306d784bd4p-3, 0x1.45f306d784b05p-3, 0x1.45f306d784ca4p-3, 0x1.45f306d784966p-3, \
      0x1.45f306d784966p-3, 0x1.45f306d78565ep-3, 0x1.45f306d783c6ep-3, 0x1.45f306d78704ep-3, \
      0x1.45f306d78704ep-3, 0x1.45f306d78704ep-3, 0x1.45f306d78704ep-3, 0x1.45f306d753248p-3, \
      0x1.45f306d7bae54p-3, 0x1.45f306d7bae54p-3, 0x1.45f306d7bae54p-3, 0x1.45f306d47cdf5p-3, \
      0x1.45f306daf8eb3p-3, 0x1.45f306daf8eb3p-3, 0x1.45f306daf8eb3p-3, 0x1.45f306a7188c7p-3, \
      0x1.45f3070ed94a0p-3, 0x1.45f3063f57cf2p-3, 0x1.45f307de5ac5dp-3, 0x1.45f304a054dcbp-3 }
static const double nco_r2pi_Pd_host[32] = NCO_R2PI_TABLES;
static const double nco_r2pi_inv_host[32] = NCO_R2PI_INVS;
#ifdef __CUDACC__
static __device__ __constant__ double nco_r2pi_Pd_dev[32] = NCO_R2PI_TABLES;
static __device__ __constant__ double nco_r2pi_inv_dev[32] = NCO_R2PI_INVS;
#endif

/* while (x > 2pi) x -= 2pi, double-domain chain (same result as nco_reduce_2pi).  In binade
 * 2^(2+s), s >= 1, a subtraction whose exact result stays at or above the binade's lower
 * border lo is x - Pd[s]; J = floor((x - lo - ulp) / Pd) such steps keep x - J*Pd >= lo + ulp,
 * for which the exact difference x - 2pi > x - Pd - ulp/2 is inside the binade.           */
NCO_HD double nco_reduce_2pi_d(double x)
{
    const double P = NCO_TWO_PI;
    while (x > P) {
        const uint64_t bx = nco_bits(x);
        const int s = (int)(bx >> 52) - (1023 + 2);
        if ((unsigned int)(s - 1) <= 26u) {                 /* x < 2^30: runs shorter than 2^29 */
#ifdef __CUDA_ARCH__
            const double Pd = nco_r2pi_Pd_dev[s], inv = nco_r2pi_inv_dev[s];
#else
            const double Pd = nco_r2pi_Pd_host[s], inv = nco_r2pi_inv_host[s];
#endif
            const double lo = nco_from_bits(bx & 0xFFF0000000000000ULL);
            const double lo1 = nco_from_bits((bx & 0xFFF0000000000000ULL) | 1ULL);    /* lo + ulp */
            double y, y2, J;
            if ((NCO_R2PI_TIES >> s) & 1u) {
                const double x1 = NCO_ADD(x, -P);
                if (!(x1 >= lo)) { x = x1; continue; }
                x = x1;                                     /* even mantissa from here on */
            }
            if (x >= lo1) {
                /* inv is biased low by 2^-30: J is never too large and (runs < 2^29) at most one
                 * too small - as in nco_walk_l                                                */
                J = NCO_FLOOR51(NCO_MUL(NCO_ADD(x, -lo1), inv));
                y = NCO_FMA(J, -Pd, x);                     /* exact */
                y2 = NCO_ADD(y, -Pd);
                x = (y2 >= lo1) ? y2 : y;
            }
            x = NCO_ADD(x, -P);                             /* may leave the binade: real */
            continue;
        }
        x = NCO_ADD(x, -P);
    }
    return x;
}

/* ---- the reference's scalar prologue / epilogue arithmetic, step for step ---- */

/* mixcarr: phi = phi0*CDIV/DPI ; ps = freq*CDIV*ti          src/sdrcmn.c:647-648 */
NCO_HD void nco_carrier_init(double freq, double phi0, double ti, double *phi, double *ps)
{
    *phi = NCO_DIV(NCO_MUL(phi0, 32.0), NCO_TWO_PI);
    *ps  = NCO_MUL(NCO_MUL(freq, 32.0), ti);
}
/* mixcarr: prem = phi*DPI/CDIV ; while(prem>DPI) prem-=DPI  src/sdrcmn.c:665-668 */
NCO_HD double nco_carrier_rem(double phi_end)
{
    return nco_reduce_2pi_d(NCO_MUL(NCO_MUL(phi_end, NCO_TWO_PI), 0.03125));   /* /32: an exact scaling */
}
/* rescode: coff-=smax*ci ; coff-=floor(coff/len)*len        src/sdrcmn.c:611-612 */
NCO_HD double nco_code_init(double coff, int smax, double ci, int len)
{
    coff = NCO_ADD(coff, -NCO_MUL((double)smax, ci));
    coff = NCO_ADD(coff, -NCO_MUL(floor(NCO_DIV(coff, (double)len)), (double)len));
    return coff;
}
/* rescode: return coff - smax*ci                            src/sdrcmn.c:619    */
NCO_HD double nco_code_rem(double coff_end, int smax, double ci)
{
    return NCO_ADD(coff_end, -NCO_MUL((double)smax, ci));
}
/* sdrtracking: (int)((clen-remcode)/(codefreq/f_sf))        src/sdrtrk.c:31-32  */
NCO_HD int32_t nco_currnsamp(int clen, double remcode, double codefreq, double f_sf)
{
    return (int32_t)NCO_DIV(NCO_ADD((double)clen, -remcode), NCO_DIV(codefreq, f_sf));
}

#endif /* SDRB200_NCO_EXACT_H */
