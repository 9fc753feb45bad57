/* port_acq.c - channel set-up and acquisition, restated for the CPU
 * (TEST INFRASTRUCTURE; see port.h). */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "port.h"

static void loop_fill(p_loop_t *l, double dllb, double pllb, double fllb)
{
    /* inittrkprmstruct, src/sdrinit.c:400-423 */
    l->dllb = dllb; l->pllb = pllb; l->fllb = fllb;
    l->dllw2 = (dllb / 0.53) * (dllb / 0.53);
    l->dllaw = 1.414 * (dllb / 0.53);
    l->pllw2 = (pllb / 0.53) * (pllb / 0.53);
    l->pllaw = 1.414 * (pllb / 0.53);
    l->fllw = fllb / 0.25;
}

/* initsdrch (src/sdrinit.c:583-657) with initacqstruct (:385-394),
 * inittrkstruct (:432-480) and the bit-sync part of initnavstruct (:489-568);
 * cfg->acq_* override the compile-time grid exactly as oracle/replay.c does. */
int p_chan_init(p_chan_t *c, const p_cfg_t *cfg, int prn, int ctype)
{
    int i;
    short *rc;
    memset(c, 0, sizeof(*c));
    c->prn = prn; c->ctype = ctype; c->dtype = cfg->dtype;
    c->f_sf = cfg->f_sf; c->f_if = cfg->f_if;
    c->ti = 1 / cfg->f_sf;                                   /* :598 */
    c->clen = p_gencode(prn, ctype, c->code, &c->crate);     /* :601 */
    if (c->clen <= 0) return -1;
    c->ci = c->ti * c->crate;                                /* :606 */
    c->ctime = c->clen / c->crate;
    c->nsamp = (int)(cfg->f_sf * c->ctime);
    c->nsampchip = (int)(c->nsamp / c->clen);
    if (ctype == 20) {                                       /* FDMA, :612-615 */
        c->f_cf = 1.60200E9 + 0.56250E6 * prn;
        c->foffset = 0.56250E6 * prn;
    } else {
        c->f_cf = cfg->f_cf; c->foffset = 0.0;               /* :619-620 */
    }
    c->intg = 10;                                            /* ACQINTG_*, src/sdr.h:141-145 */
    c->nfreq = 2 * (7000 / 200) + 1;                         /* ACQHBAND/ACQSTEP, :393 */
    {
        double step = 200.0;
        if (cfg->acq_intg > 0) c->intg = cfg->acq_intg;
        if (cfg->acq_step > 0.0 && cfg->acq_hband > 0.0) {
            step = cfg->acq_step;
            c->nfreq = 2 * (int)(cfg->acq_hband / cfg->acq_step) + 1;
        }
        c->nfft = 2 * c->nsamp;                              /* :625 */
        c->freq = (double *)malloc(sizeof(double) * c->nfreq);
        if (!c->freq) return -1;
        for (i = 0; i < c->nfreq; i++)                       /* :633-635 */
            c->freq[i] = c->f_if + ((i - (c->nfreq - 1) / 2) * step) + c->foffset;
    }
    /* tracking taps and loop schedule, :441-474 */
    loop_fill(&c->prm1, cfg->dllb[0], cfg->pllb[0], cfg->fllb[0]);
    loop_fill(&c->prm2, cfg->dllb[1], cfg->pllb[1], cfg->fllb[1]);
    c->corrn = cfg->corrn;
    if (c->corrn < 1 || c->corrn > 16) return -1;
    for (i = 0; i < c->corrn; i++) {
        c->corrp[i] = cfg->corrd * (i + 1);
        if (c->corrp[i] == cfg->corrp) { c->ne = 2 * (i + 1) - 1; c->nl = 2 * (i + 1); }
    }
    {
        const int loop = (ctype == 27) ? 2 : 10;             /* LOOP_*, src/sdr.h:152-154 */
        c->loopms = loop * (int)(c->ctime * 1000);
    }
    c->rate = (ctype == 1) ? 20 : (ctype == 20) ? 10 : 2;    /* NAVRATE_*, src/sdr.h:159-171 */
    /* frequency-domain replica, :645-653 */
    rc = (short *)calloc((size_t)c->nfft, sizeof(short));
    c->xcode = (float *)malloc(sizeof(float) * 2 * (size_t)c->nfft);
    if (!rc || !c->xcode) { free(rc); return -1; }
    p_rescode(c->code, c->clen, 0, 0, c->ci, c->nsamp, rc);
    p_cpxcpx(rc, NULL, 1.0, c->nfft, c->xcode);
    p_fft(c->xcode, c->nfft, 0);
    free(rc);
    return 0;
}

void p_chan_free(p_chan_t *c)
{
    free(c->freq); free(c->xcode);
    c->freq = NULL; c->xcode = NULL;
}

/* src/sdracq.c:71-95 */
int p_checkacquisition(const double *P, p_chan_t *c)
{
    int flat, codei, freqi, exs, exe, dummy;
    const int n = c->nsamp;
    const double maxP = p_maxvd(P, n * c->nfreq, -1, -1, &flat);
    double meanP, maxP2;
    codei = flat % n;                                        /* ind2sub, src/sdrcmn.c:574-578 */
    freqi = c->nfreq * flat / (n * c->nfreq);
    exs = codei - 2 * c->nsampchip; if (exs < 0) exs += n;
    exe = codei + 2 * c->nsampchip; if (exe >= n) exe -= n;
    meanP = p_meanvd(P + (size_t)freqi * n, n, exs, exe);
    c->cn0 = 10 * log10(maxP / meanP / c->ctime);
    maxP2 = p_maxvd(P + (size_t)freqi * n, n, exs, exe, &dummy);
    c->peakr = maxP / maxP2;
    c->acqcodei = codei;
    c->freqi = freqi;
    c->acqfreq = c->freq[freqi];
    return c->peakr > P_ACQTH;
}

/* src/sdracq.c:14-62; `stream` is the IF addressed by absolute sample index,
 * `avail` = fendbuffsize*buffcnt, the samples received so far */
uint64_t p_acquisition(p_chan_t *c, const signed char *stream, uint64_t avail, double *power)
{
    uint64_t buffloc = avail - (uint64_t)(c->intg + 1) * c->nsamp;      /* :24 */
    int i;
    c->nintg_used = 0;
    for (i = 0; i < c->intg; i++) {
        const signed char *win = stream + buffloc * c->dtype;            /* 2*nsamp samples, :31 */
        buffloc += c->nsamp;
        p_pcorrelator(win, c->dtype, c->ti, c->nsamp, c->freq, c->nfreq, c->crate, c->nfft,
                      c->xcode, power);
        c->nintg_used++;
        if (p_checkacquisition(power, c)) { c->flagacq = 1; break; }
    }
    if (c->flagacq) {                                                    /* :51-56 */
        buffloc += -(uint64_t)(i + 1) * c->nsamp + c->acqcodei;
        c->carrfreq = c->acqfreq;
        c->codefreq = c->crate;
    }
    return buffloc;
}
