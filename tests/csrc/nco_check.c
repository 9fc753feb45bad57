/* nco_check.c - CPU check of csrc/nco_exact.h against the plain sequential
 * loops it must reproduce (the loops restate src/sdrcmn.c:608-621, 647-668). */
#include <stdio.h>
#include <stdlib.h>
#include "nco_exact.h"

static uint64_t rs;
static double urand(void)
{
    rs = rs * 6364136223846793005ULL + 1442695040888963407ULL;
    return (double)(rs >> 11) * (1.0 / 9007199254740992.0);
}

int max_isegments = 0;

/* integer walker against the literal loop: (int) cast of every value and the end value */
int check_walk_i(double x0, double step, int n, double wrap)
{
    static nco_iseg_t seg[NCO_MAXSEG];
    double p = x0, xe;
    int k, bad = 0, s = 0;
    const int ns = nco_walk_i(x0, step, n, wrap, seg, NCO_MAXSEG, &xe);
    if (ns < 0) return 1000000;
    if (ns > max_isegments) max_isegments = ns;
    for (k = 0; k < n; k++, p += step) {
        if (wrap > 0.0 && p >= wrap) p -= wrap;
        while (s + 1 < ns && seg[s + 1].k0 <= k) s++;
        if (fabs(p) < 9.0e15) {
            if (seg[s].shift < 0 || nco_iseg_trunc(&seg[s], k) != (int64_t)p) bad++;
        }
        if (seg[s].shift >= 0 && seg[s].shift <= 52 && fabs(p) >= 1.0) {
            /* full value, not only its integer part */
            const uint64_t m = seg[s].X0 + (uint64_t)((int64_t)(k - seg[s].k0) * seg[s].D);
            if (ldexp((double)m, -seg[s].shift) != fabs(p)) bad++;
        }
    }
    if (p != xe) bad++;
    return bad;
}

int max_tsegments = 0;

/* the tracking kernel's walker (step table + integer segments out) against the literal loop */
int check_walk_t(double x0, double step, int n, double wrap)
{
    static nco_iseg_t seg[NCO_MAXSEG];
    static nco_steptab_t tab;
    double p = x0, xe;
    int k, bad = 0, s = 0, ns;
    nco_steptab_build(step, &tab);
    ns = nco_walk_l(x0, step, n, wrap, &tab, seg, NCO_MAXSEG, &xe);
    if (ns < 0) return 1000000;
    if (ns > max_tsegments) max_tsegments = ns;
    for (k = 0; k < n; k++, p += step) {
        if (wrap > 0.0 && p >= wrap) p -= wrap;
        while (s + 1 < ns && seg[s + 1].k0 <= k) s++;
        if (fabs(p) < 9.0e15) {
            if (seg[s].shift < 0 || nco_iseg_trunc(&seg[s], k) != (int64_t)p) bad++;
        }
        if (seg[s].shift >= 0 && seg[s].shift <= 52 && fabs(p) >= 1.0) {
            const uint64_t m = seg[s].X0 + (uint64_t)((int64_t)(k - seg[s].k0) * seg[s].D);
            if (ldexp((double)m, -seg[s].shift) != fabs(p)) bad++;
        }
    }
    if (p != xe) bad++;
    return bad;
}

/* returns mismatches; carrier NCO */
int check_carrier(double freq, double phi0, double ti, int n)
{
    static nco_table_t t;
    double phi, ps, p, prem;
    int k, bad = 0, hint = 0;
    nco_carrier_init(freq, phi0, ti, &phi, &ps);
    if (nco_walk(phi, ps, n, 0.0, &t)) return -1;
    p = phi;
    for (k = 0; k < n; k++, p += ps) {
        double v = nco_value(&t, k, &hint);
        nco_iseg_t is;
        if (v != p) bad++;
        nco_iseg_make(&t.seg[hint], &is);                 /* integer form: same (int) cast */
        if (is.shift < 0 || (int)nco_iseg_trunc(&is, k) != (int)p) bad++;
    }
    if (p != t.x_end) bad++;
    bad += check_walk_i(phi, ps, n, 0.0);
    bad += check_walk_t(phi, ps, n, 0.0);
    prem = p * NCO_TWO_PI / 32.0;
    while (prem > NCO_TWO_PI) prem -= NCO_TWO_PI;
    if (prem != nco_carrier_rem(t.x_end)) bad++;
    return bad;
}

int check_code(double coff, int smax, double ci, int len, int n)
{
    static nco_table_t t;
    double c, c0;
    int k, bad = 0, hint = 0, N = n + 2 * smax;
    c0 = nco_code_init(coff, smax, ci, len);
    if (nco_walk(c0, ci, N, (double)len, &t)) return -1;
    c = coff;
    c -= smax * ci;
    c -= floor(c / len) * len;
    if (c != c0) bad++;
    for (k = 0; k < N; k++, c += ci) {
        double v;
        if (c >= len) c -= len;
        v = nco_value(&t, k, &hint);
        if (v != c) bad++;
        {
            nco_iseg_t is;
            nco_iseg_make(&t.seg[hint], &is);
            if (is.shift < 0 || (int)nco_iseg_trunc(&is, k) != (int)c) bad++;
        }
    }
    if (c != t.x_end) bad++;
    bad += check_walk_i(c0, ci, N, (double)len);
    bad += check_walk_t(c0, ci, N, (double)len);
    if (c - smax * ci != nco_code_rem(t.x_end, smax, ci)) bad++;
    return bad;
}

int max_segments = 0;

/* mode 0: IF-style carrier (fs/4 +- doppler), 1: baseband carrier (+-10 kHz),
 * 2: code 16.368 MS/s, 3: code 2.048 MS/s, 5: zero / sub-ulp steps,
 * other: adversarial steps with short mantissas */
int fuzz(unsigned long long seed, int iters, int mode)
{
    int i, bad = 0;
    rs = seed;
    for (i = 0; i < iters; i++) {
        int r;
        if (mode == 0) {
            double fs = 16.368e6, f = 4.092e6 + (urand() - 0.5) * 2e4;
            r = check_carrier(f, urand() * NCO_TWO_PI, 1.0 / fs, 16360 + (int)(urand() * 20));
        } else if (mode == 1) {
            double fs = 2.048e6, f = (urand() - 0.5) * 2e4;
            r = check_carrier(f, urand() * NCO_TWO_PI, 1.0 / fs, 2040 + (int)(urand() * 20));
        } else if (mode == 2) {
            double fs = 16.368e6, cr = 1.023e6 + (urand() - 0.5) * 20.0;
            r = check_code(urand() * 0.07, 18, (1.0 / fs) * cr, 1023, 16360 + (int)(urand() * 20));
        } else if (mode == 3) {
            double fs = 2.048e6, cr = 1.023e6 + (urand() - 0.5) * 20.0;
            r = check_code(urand() * 1023.0, 4, (1.0 / fs) * cr, 1023, 2040 + (int)(urand() * 20));
        } else if (mode == 5) {
            /* zero and sub-ulp steps (an acquisition bin at exactly 0 Hz) */
            double f = (urand() < 0.5) ? 0.0 : (urand() - 0.5) * 1e-6;
            r = check_carrier(f, (urand() < 0.3) ? 0.0 : urand() * NCO_TWO_PI, 1.0 / 2.048e6, 2048);
        } else {
            /* steps with few mantissa bits force exact ties in many binades */
            double step = ldexp(floor(urand() * 4096.0) + 0.5, -(int)(urand() * 12));
            double x0 = ldexp(floor(urand() * 65536.0), -(int)(urand() * 20));
            if (urand() < 0.3) step = -step;
            {
                static nco_table_t t;
                double p = x0; int k, n = 3000 + (int)(urand() * 3000), hint = 0;
                r = 0;
                if (nco_walk(x0, step, n, 0.0, &t)) r = -1;
                else {
                    for (k = 0; k < n; k++, p += step) if (nco_value(&t, k, &hint) != p) r++;
                    if (p != t.x_end) r++;
                    if (t.nseg > max_segments) max_segments = t.nseg;
                    r += check_walk_i(x0, step, n, 0.0);
                    r += check_walk_t(x0, step, n, 0.0);
                }
            }
        }
        if (r) { bad++; if (bad < 5) fprintf(stderr, "mismatch mode %d iter %d r=%d\n", mode, i, r); }
    }
    return bad;
}

/* 2*pi reduction (src/sdrcmn.c:666-668) against the literal loop: random and structured x */
int fuzz_reduce(int n)
{
    const int iters = n / 5;
        int i, b2 = 0;
        rs = 99;
        for (i = 0; i < iters * 5; i++) {
            double x = urand() * ldexp(1.0, (int)(urand() * 18)), y = x;
            while (y > NCO_TWO_PI) y -= NCO_TWO_PI;
            if (y != nco_reduce_2pi(x) || y != nco_reduce_2pi_d(x)) b2++;
        }
        /* structured: mantissas just above a binade border, near multiples of 2*pi, large x */
        for (i = 0; i < iters * 5; i++) {
            const int e = 3 + (int)(urand() * 18);
            double x, y;
            switch (i % 4) {
            case 0: x = ldexp(1.0, e) + ldexp(floor(urand() * 4096.0), e - 52); break;
            case 1: x = floor(urand() * ldexp(1.0, e - 2)) * NCO_TWO_PI + (urand() - 0.5) * 1e-9; break;
            case 2: x = ldexp(2.0, e) - ldexp(floor(urand() * 4096.0), e - 52); break;
            default: x = ldexp(1.0 + urand(), e); break;
            }
            y = x;
            while (y > NCO_TWO_PI) y -= NCO_TWO_PI;
            if (y != nco_reduce_2pi(x) || y != nco_reduce_2pi_d(x)) b2++;
        }
    return b2;
}

#ifdef NCO_MAIN
int main(int argc, char **argv)
{
    int iters = argc > 1 ? atoi(argv[1]) : 2000, m, bad = 0;
    for (m = 0; m < 6; m++) { int b = fuzz(1234 + m, iters, m); printf("mode %d: %d bad of %d\n", m, b, iters); bad += b; }
    printf("max segments (mode 4) %d, integer walker %d, lean table walker %d\n", max_segments, max_isegments, max_tsegments);
    { int b2 = fuzz_reduce(iters * 5); printf("reduce_2pi: %d bad of %d\n", b2, iters * 10); bad += b2; }
    return bad != 0;
}
#endif
