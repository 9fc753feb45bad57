/* offt.c - ORACLE-ONLY arbitrary-length complex DFT (see offt.h).
 *
 * Mixed-radix decimation-in-frequency Stockham autosort transform.
 * Radices 4, 2, 3, 5 have dedicated butterflies; any other prime factor p is
 * handled by a direct O(p^2) butterfly that exploits conjugate symmetry of
 * the p-th roots of unity (so 32736 = 2^5*3*11*31, 40000 = 2^6*5^4 and
 * 654720 = 2^7*3*5*11*31 all work).  All arithmetic is double precision.
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "offt.h"

#define OFFT_MAXFAC 64

struct offt_plan_s {
    int n, sign;
    int nfac;
    int fac[OFFT_MAXFAC];
    double *tw;          /* 2*n doubles: cos, sign*sin of 2*pi*k/n          */
};

static int factorise(int n, int *fac)
{
    int nf = 0, p;
    while (n % 4 == 0) { fac[nf++] = 4; n /= 4; }
    while (n % 2 == 0) { fac[nf++] = 2; n /= 2; }
    for (p = 3; (long long)p * p <= n; p += 2)
        while (n % p == 0) { fac[nf++] = p; n /= p; }
    if (n > 1) fac[nf++] = n;
    return nf;
}

offt_plan *offt_create(int n, int sign)
{
    offt_plan *p;
    int k;
    if (n < 1 || (sign != 1 && sign != -1)) return NULL;
    p = (offt_plan *)calloc(1, sizeof(*p));
    if (!p) return NULL;
    p->n = n; p->sign = sign;
    p->nfac = (n > 1) ? factorise(n, p->fac) : 0;
    p->tw = (double *)malloc(sizeof(double) * 2 * (size_t)n);
    if (!p->tw) { free(p); return NULL; }
    for (k = 0; k < n; k++) {
        long double a = 6.283185307179586476925286766559005768L * (long double)k / (long double)n;
        p->tw[2 * k]     = (double)cosl(a);
        p->tw[2 * k + 1] = (double)(sign * sinl(a));
    }
    return p;
}

void offt_destroy(offt_plan *p)
{
    if (!p) return;
    free(p->tw);
    free(p);
}

int offt_length(const offt_plan *p) { return p ? p->n : 0; }

/* one Stockham stage: radix r, current sub-length ncur = r*m, stride s */
static void stage(const offt_plan *pl, int r, int m, int s,
                  const double *src, double *dst)
{
    const int n = pl->n;
    const double *tw = pl->tw;
    const int rstep = n / r;          /* w_r = w_n^(n/r) */
    int p, q, t, u;

    if (r == 2) {
        for (p = 0; p < m; p++) {
            const double wr = tw[2 * (size_t)((long long)p * s % n)];
            const double wi = tw[2 * (size_t)((long long)p * s % n) + 1];
            const double *a = src + 2 * (size_t)s * p;
            const double *b = src + 2 * (size_t)s * (p + m);
            double *y0 = dst + 2 * (size_t)s * (2 * p);
            double *y1 = dst + 2 * (size_t)s * (2 * p + 1);
            for (q = 0; q < s; q++) {
                double ar = a[2*q], ai = a[2*q+1], br = b[2*q], bi = b[2*q+1];
                double dr = ar - br, di = ai - bi;
                y0[2*q] = ar + br; y0[2*q+1] = ai + bi;
                y1[2*q] = dr * wr - di * wi; y1[2*q+1] = dr * wi + di * wr;
            }
        }
        return;
    }
    if (r == 4) {
        const double sg = (double)pl->sign;    /* multiply by sign*i */
        for (p = 0; p < m; p++) {
            size_t i1 = (size_t)((long long)p * s % n);
            size_t i2 = (size_t)((long long)2 * p * s % n);
            size_t i3 = (size_t)((long long)3 * p * s % n);
            const double w1r = tw[2*i1], w1i = tw[2*i1+1];
            const double w2r = tw[2*i2], w2i = tw[2*i2+1];
            const double w3r = tw[2*i3], w3i = tw[2*i3+1];
            const double *a0 = src + 2 * (size_t)s * p;
            const double *a1 = src + 2 * (size_t)s * (p + m);
            const double *a2 = src + 2 * (size_t)s * (p + 2 * m);
            const double *a3 = src + 2 * (size_t)s * (p + 3 * m);
            double *y0 = dst + 2 * (size_t)s * (4 * p);
            double *y1 = y0 + 2 * (size_t)s;
            double *y2 = y1 + 2 * (size_t)s;
            double *y3 = y2 + 2 * (size_t)s;
            for (q = 0; q < s; q++) {
                double x0r = a0[2*q], x0i = a0[2*q+1], x1r = a1[2*q], x1i = a1[2*q+1];
                double x2r = a2[2*q], x2i = a2[2*q+1], x3r = a3[2*q], x3i = a3[2*q+1];
                double s02r = x0r + x2r, s02i = x0i + x2i, d02r = x0r - x2r, d02i = x0i - x2i;
                double s13r = x1r + x3r, s13i = x1i + x3i, d13r = x1r - x3r, d13i = x1i - x3i;
                /* j*d13 where j = sign*i : (re,im) -> sg*(-im, re) */
                double jr = -sg * d13i, ji = sg * d13r;
                double b1r = d02r + jr, b1i = d02i + ji;
                double b2r = s02r - s13r, b2i = s02i - s13i;
                double b3r = d02r - jr, b3i = d02i - ji;
                y0[2*q] = s02r + s13r; y0[2*q+1] = s02i + s13i;
                y1[2*q] = b1r * w1r - b1i * w1i; y1[2*q+1] = b1r * w1i + b1i * w1r;
                y2[2*q] = b2r * w2r - b2i * w2i; y2[2*q+1] = b2r * w2i + b2i * w2r;
                y3[2*q] = b3r * w3r - b3i * w3i; y3[2*q+1] = b3r * w3i + b3i * w3r;
            }
        }
        return;
    }
    /* generic odd radix r (3, 5 and every other prime) */
    {
        const int h = (r - 1) / 2;
        double *sp = (double *)malloc(sizeof(double) * 4 * (size_t)(h + 1));
        double *dm = sp + 2 * (size_t)(h + 1);
        for (p = 0; p < m; p++) {
            for (q = 0; q < s; q++) {
                const double *a0 = src + 2 * ((size_t)s * p + q);
                double *y = dst + 2 * ((size_t)s * ((size_t)r * p) + q);
                double sumr = a0[0], sumi = a0[1];
                for (t = 1; t <= h; t++) {
                    const double *at = src + 2 * ((size_t)s * (p + (size_t)t * m) + q);
                    const double *ar = src + 2 * ((size_t)s * (p + (size_t)(r - t) * m) + q);
                    sp[2*t] = at[0] + ar[0]; sp[2*t+1] = at[1] + ar[1];
                    dm[2*t] = at[0] - ar[0]; dm[2*t+1] = at[1] - ar[1];
                    sumr += sp[2*t]; sumi += sp[2*t+1];
                }
                y[0] = sumr; y[1] = sumi;                 /* u = 0, twiddle 1 */
                for (u = 1; u <= h; u++) {
                    double Ar = a0[0], Ai = a0[1], Br = 0.0, Bi = 0.0;
                    for (t = 1; t <= h; t++) {
                        size_t k = (size_t)((long long)t * u % r) * (size_t)rstep;
                        double c = tw[2*k], sn = tw[2*k+1];   /* sn carries the sign */
                        Ar += sp[2*t] * c;  Ai += sp[2*t+1] * c;
                        Br += dm[2*t] * sn; Bi += dm[2*t+1] * sn;
                    }
                    /* b_u = A + i*B ; b_{r-u} = A - i*B   (i*(x+iy) = -y + ix) */
                    {
                        double b1r = Ar - Bi, b1i = Ai + Br;
                        double b2r = Ar + Bi, b2i = Ai - Br;
                        size_t k1 = (size_t)((long long)p * u % n * s % n);
                        size_t k2 = (size_t)((long long)p * (r - u) % n * s % n);
                        double w1r = tw[2*k1], w1i = tw[2*k1+1];
                        double w2r = tw[2*k2], w2i = tw[2*k2+1];
                        double *y1 = y + 2 * (size_t)s * u;
                        double *y2 = y + 2 * (size_t)s * (r - u);
                        y1[0] = b1r * w1r - b1i * w1i; y1[1] = b1r * w1i + b1i * w1r;
                        y2[0] = b2r * w2r - b2i * w2i; y2[1] = b2r * w2i + b2i * w2r;
                    }
                }
            }
        }
        free(sp);
    }
}

void offt_exec_f64(const offt_plan *pl, double *x)
{
    int n, ncur, s, i;
    double *y, *src, *dst, *tmp;
    if (!pl || pl->n <= 1) return;
    n = pl->n;
    y = (double *)malloc(sizeof(double) * 2 * (size_t)n);
    if (!y) return;
    src = x; dst = y; ncur = n; s = 1;
    for (i = 0; i < pl->nfac; i++) {
        int r = pl->fac[i], m = ncur / r;
        stage(pl, r, m, s, src, dst);
        tmp = src; src = dst; dst = tmp;
        ncur = m; s *= r;
    }
    if (src != x) memcpy(x, src, sizeof(double) * 2 * (size_t)n);
    free(y);
}

void offt_exec_f32(const offt_plan *pl, float *x)
{
    int n, i;
    double *w;
    if (!pl || pl->n <= 1) return;
    n = pl->n;
    w = (double *)malloc(sizeof(double) * 2 * (size_t)n);
    if (!w) return;
    for (i = 0; i < 2 * n; i++) w[i] = (double)x[i];
    offt_exec_f64(pl, w);
    for (i = 0; i < 2 * n; i++) x[i] = (float)w[i];
    free(w);
}
