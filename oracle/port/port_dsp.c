/* port_dsp.c - CPU restatement of the reference's DSP primitives
 * (TEST INFRASTRUCTURE; see port.h).  The arithmetic order is the reference's:
 * NCO phases advance by repeated double addition and are truncated with a C
 * cast; tap sums add integer products into doubles; FFT-domain work is float. */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "port.h"
#include "../offt.h"

/* carrier table of src/sdrcmn.c:639-646, built once */
static short g_cos[P_CDIV], g_sin[P_CDIV];
static int g_lut_ready = 0;
static void lut_build(void)
{
    int k;
    for (k = 0; k < P_CDIV; k++) {
        g_cos[k] = (short)floor(cos(P_DPI / P_CDIV * k) / P_CSCALE + 0.5);
        g_sin[k] = (short)floor(sin(P_DPI / P_CDIV * k) / P_CSCALE + 0.5);
    }
    g_lut_ready = 1;
}

/* src/sdrcmn.c:633-669 */
double p_mixcarr(const signed char *data, int dtype, double ti, int n, double freq,
                 double phi0, short *I, short *Q)
{
    double phi = phi0 * P_CDIV / P_DPI;           /* :647 */
    const double ps = freq * P_CDIV * ti;          /* :648 */
    double rem;
    int k, idx;
    if (!g_lut_ready) lut_build();
    if (dtype == 2) {                              /* :651-657 */
        for (k = 0; k < n; k++, phi += ps) {
            const int x0 = data[2 * k], x1 = data[2 * k + 1];
            idx = ((int)phi) & P_CMASK;
            I[k] = (short)(g_cos[idx] * x0 - g_sin[idx] * x1);
            Q[k] = (short)(g_sin[idx] * x0 + g_cos[idx] * x1);
        }
    } else if (dtype == 1) {                       /* :658-664 */
        for (k = 0; k < n; k++, phi += ps) {
            idx = ((int)phi) & P_CMASK;
            I[k] = (short)(g_cos[idx] * data[k]);
            Q[k] = (short)(g_sin[idx] * data[k]);
        }
    }
    rem = phi * P_DPI / P_CDIV;                    /* :665 */
    while (rem > P_DPI) rem -= P_DPI;              /* :666 */
    return rem;
}

/* src/sdrcmn.c:608-621 */
double p_rescode(const short *code, int len, double coff, int smax, double ci, int n, short *rcode)
{
    int k;
    const int total = n + 2 * smax;
    coff -= smax * ci;
    coff -= floor(coff / len) * len;
    for (k = 0; k < total; k++, coff += ci) {
        if (coff >= len) coff -= len;
        rcode[k] = code[(int)coff];
    }
    return coff - smax * ci;
}

/* dot_21 / dot_22 / dot_23 (src/sdrcmn.c:287-354) for nb = 1, 2, 3 replicas:
 * d1[j] = sum a1[i]*b[j][i], d2[j] = sum a2[i]*b[j][i]; int products, double sums */
void p_dots(const short *a1, const short *a2, const short *const *b, int nb, int n,
            double *d1, double *d2)
{
    int i, j;
    for (j = 0; j < nb; j++) d1[j] = d2[j] = 0.0;
    for (i = 0; i < n; i++)
        for (j = 0; j < nb; j++) {
            d1[j] += a1[i] * b[j][i];
            d2[j] += a2[i] * b[j][i];
        }
}

/* src/sdrcmn.c:687-722 */
void p_correlator(const signed char *data, int dtype, double ti, int n, double freq,
                  double phi0, double crate, double coff, const int *s, int ns,
                  double *II, double *QQ, double *remc, double *remp,
                  const short *code, int coden)
{
    const int smax = s[ns - 1];                    /* :694 */
    short *mI = (short *)malloc(sizeof(short) * (size_t)(n + 64));
    short *mQ = (short *)malloc(sizeof(short) * (size_t)(n + 64));
    short *rep = (short *)malloc(sizeof(short) * (size_t)(n + 2 * smax));
    const short *prompt, *b[3];
    int i;
    if (!mI || !mQ || !rep) { free(mI); free(mQ); free(rep); return; }
    prompt = rep + smax;
    *remp = p_mixcarr(data, dtype, ti, n, freq, phi0, mI, mQ);            /* :706 */
    *remc = p_rescode(code, coden, coff, smax, ti * crate, n, rep);       /* :709 */
    b[0] = prompt; b[1] = prompt - s[0]; b[2] = prompt + s[0];            /* P, E1, L1  :712 */
    p_dots(mI, mQ, b, 3, n, II, QQ);
    for (i = 1; i < ns; i++) {                                            /* E_i, L_i   :713-715 */
        b[0] = prompt - s[i]; b[1] = prompt + s[i];
        p_dots(mI, mQ, b, 2, n, II + 1 + 2 * i, QQ + 1 + 2 * i);
    }
    for (i = 0; i < 1 + 2 * ns; i++) { II[i] *= P_CSCALE; QQ[i] *= P_CSCALE; }   /* :716-719 */
    free(mI); free(mQ); free(rep);
}

/* src/sdrcmn.c:185-195 */
void p_cpxcpx(const short *I, const short *Q, double scale, int n, float *cpx)
{
    int i;
    const float sc = (float)scale;
    for (i = 0; i < n; i++) {
        cpx[2 * i] = I[i] * sc;
        cpx[2 * i + 1] = Q ? Q[i] * sc : 0.0f;
    }
}

/* cpxfft / cpxifft (src/sdrcmn.c:134-175): in-place unnormalised DFT.  FFTW is
 * replaced by oracle/offt.c exactly as in the reference build of the oracle.   */
void p_fft(float *cpx, int n, int inverse)
{
    static struct { int n, sign; offt_plan *p; } cache[8];
    static int used = 0;
    const int sign = inverse ? +1 : -1;
    offt_plan *p = NULL;
    int i;
    for (i = 0; i < used; i++) if (cache[i].n == n && cache[i].sign == sign) p = cache[i].p;
    if (!p) {
        p = offt_create(n, sign);
        if (!p) return;
        if (used < 8) { cache[used].n = n; cache[used].sign = sign; cache[used].p = p; used++; }
    }
    offt_exec_f32(p, cpx);
}

/* src/sdrcmn.c:228-251 */
void p_cpxconv(float *a, const float *b, int m, int n, int flagsum, double *conv)
{
    const float m2 = (float)m * m;
    int i;
    p_fft(a, m, 0);
    for (i = 0; i < m; i++) {                      /* a = -(a * conj(b))  :236-240 */
        const float ar = a[2 * i], ai = a[2 * i + 1], br = b[2 * i], bi = b[2 * i + 1];
        const float re = -ar * br - ai * bi;
        a[2 * i + 1] = ar * bi - ai * br;
        a[2 * i] = re;
    }
    p_fft(a, m, 1);
    for (i = 0; i < n; i++) {
        const float pw = (a[2 * i] * a[2 * i] + a[2 * i + 1] * a[2 * i + 1]) / m2;
        if (flagsum) conv[i] += pw; else conv[i] = pw;
    }
}

/* src/sdrcmn.c:261-276 */
void p_cpxpspec(float *cpx, int n, int flagsum, double *pspec)
{
    int i;
    p_fft(cpx, n, 0);
    for (i = 0; i < n; i++) {
        const float pw = cpx[2 * i] * cpx[2 * i] + cpx[2 * i + 1] * cpx[2 * i + 1];
        if (flagsum) pspec[i] += pw; else pspec[i] = pw;
    }
}

/* src/sdrcmn.c:738-773 */
void p_pcorrelator(const signed char *data, int dtype, double ti, int n, const double *freq,
                   int nfreq, double crate, int m, const float *codex, double *P)
{
    signed char *win = (signed char *)calloc((size_t)m * dtype, 1);        /* zero padding :756 */
    short *mI = (short *)malloc(sizeof(short) * (size_t)(m + 64));
    short *mQ = (short *)malloc(sizeof(short) * (size_t)(m + 64));
    float *x = (float *)malloc(sizeof(float) * 2 * (size_t)m);
    int f;
    (void)crate;
    if (win && mI && mQ && x) {
        memcpy(win, data, (size_t)2 * n * dtype);                          /* :757 */
        for (f = 0; f < nfreq; f++) {
            p_mixcarr(win, dtype, ti, m, freq[f], 0.0, mI, mQ);            /* :761 */
            p_cpxcpx(mI, mQ, P_CSCALE / m, m, x);                          /* :764 */
            p_cpxconv(x, codex, m, n, 1, P + (size_t)f * n);               /* :767 */
        }
    }
    free(win); free(mI); free(mQ); free(x);
}

static int outside(int i, int exs, int exe)
{
    return (exs <= exe) ? (i < exs || i > exe) : (i < exs && i > exe);
}

/* src/sdrcmn.c:461-476: seeded with d[0] even when index 0 is excluded;
 * strict '<' so the first maximum wins */
double p_maxvd(const double *d, int n, int exs, int exe, int *ind)
{
    double best = d[0];
    int i;
    *ind = 0;
    for (i = 1; i < n; i++)
        if (outside(i, exs, exe) && best < d[i]) { best = d[i]; *ind = i; }
    return best;
}

/* src/sdrcmn.c:487-497 */
double p_meanvd(const double *d, int n, int exs, int exe)
{
    double sum = 0.0;
    int i, skipped = 0;
    for (i = 0; i < n; i++) {
        if (outside(i, exs, exe)) sum += d[i]; else skipped++;
    }
    return sum / (n - skipped);
}
