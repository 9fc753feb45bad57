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
 * nco_walk_* therefore performs real additions only around binade changes /
 * wraps and jumps over the interior of each binade, emitting a short table of
 * linear segments (a few dozen entries for 16368 samples).  Any thread can
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

/* How many further constant-increment steps may be taken from x1 (increment d,
 * same binade assumed for x1 and x1+d) such that every value stays inside the
 * binade of x1 and strictly below `limit` in the direction of travel
 * (limit_on: 0 none, 1: values must stay < limit, 2: values must stay > limit).
 * Conservative by two steps; the caller finishes with real additions.          */
NCO_HD int32_t nco_room(double x1, double d, int limit_on, double limit, int32_t maxj)
{
    double lo = nco_binade_lo(x1);          /* magnitude bounds [lo, 2lo)      */
    double hi = lo + lo;
    double m = fabs(x1);
    double room, j;
    int growing = ((x1 > 0.0) == (d > 0.0));
    if (d == 0.0) return maxj;
    room = growing ? (hi - m) : (m - lo);
    if (limit_on == 1 && d > 0.0) { double r2 = limit - x1; if (r2 < room) room = r2; }
    if (limit_on == 2 && d < 0.0) { double r2 = x1 - limit; if (r2 < room) room = r2; }
    if (!(room > 0.0)) return 0;
    j = floor(room / fabs(d)) - 2.0;
    if (j < 0.0) return 0;
    if (j > (double)maxj) return maxj;
    return (int32_t)j;
}

NCO_HD void nco_emit(nco_table_t *t, int32_t k0, double x0, double d)
{
    if (t->nseg < NCO_MAXSEG) {
        t->seg[t->nseg].k0 = k0; t->seg[t->nseg].pad = 0;
        t->seg[t->nseg].x0 = x0; t->seg[t->nseg].d = d;
        t->nseg++;
    } else {
        t->nseg = NCO_MAXSEG + 1;           /* overflow marker (caller checks) */
    }
}

/* Sequence   for k in 0..count-1: { if (wrap>0 && x>=wrap) x-=wrap; use x; x+=step; }
 * (wrap = 0: no wrap test, the carrier NCO; wrap = code length: the code NCO).
 * Fills t with segments describing every USED value; t->x_end = x after the
 * last addition.  Returns 0, or -1 if the table overflowed.                    */
NCO_HD int nco_walk(double x, double step, int32_t count, double wrap, nco_table_t *t)
{
    int32_t k = 0;
    const int lim = (wrap > 0.0) ? 1 : 0;
    t->nseg = 0; t->count = count;
    while (k < count) {
        double x1, x2, d;
        int32_t J;
        if (lim && x >= wrap) x = NCO_ADD(x, -wrap);
        x1 = NCO_ADD(x, step);
        if (x1 == x) {                      /* step is zero or below half an ulp: */
            nco_emit(t, k, x, 0.0);         /* the accumulator never moves again  */
            k = count;
            break;
        }
        if (k + 1 >= count || (lim && x1 >= wrap) || !nco_same_binade(x, x1)) {
            nco_emit(t, k, x, 0.0); k += 1; x = x1; continue;
        }
        x2 = NCO_ADD(x1, step);
        if (k + 2 >= count || (lim && x2 >= wrap) || !nco_same_binade(x1, x2)) {
            nco_emit(t, k, x, NCO_ADD(x1, -x)); k += 2; x = x2; continue;
        }
        d = NCO_ADD(x2, -x1);
        /* steady state from x1: values x1 + j*d, j = 0..J, J >= 1 (x2 is j = 1) */
        J = nco_room(x1, d, lim, wrap, count - 2 - k);
        if (J < 1) J = 1;
        nco_emit(t, k, x, 0.0);
        nco_emit(t, k + 1, x1, d);
        x = NCO_ADD(NCO_ADD(x1, NCO_MUL((double)J, d)), step);
        k += 2 + J;
    }
    t->x_end = x;
    return (t->nseg > NCO_MAXSEG) ? -1 : 0;
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

/* while (x > 2pi) x -= 2pi;   (src/sdrcmn.c:666-668), exactly */
NCO_HD double nco_reduce_2pi(double x)
{
    const double P = NCO_TWO_PI;
    while (x > P) {
        double x1 = NCO_ADD(x, -P), x2, d;
        int32_t J;
        if (!(x1 > P) || !nco_same_binade(x, x1)) { x = x1; continue; }
        x2 = NCO_ADD(x1, -P);
        if (!(x2 > P) || !nco_same_binade(x1, x2)) { x = x2; continue; }
        d = NCO_ADD(x2, -x1);
        J = nco_room(x1, d, 2, P, 0x7FFFFFF0);
        if (J < 1) J = 1;
        x = NCO_ADD(x1, NCO_MUL((double)J, d));
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
    return nco_reduce_2pi(NCO_DIV(NCO_MUL(phi_end, NCO_TWO_PI), 32.0));
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
