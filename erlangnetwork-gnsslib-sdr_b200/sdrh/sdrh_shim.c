/* sdrh_shim.c - the reference's own sdr.h names on top of libsdrb200.
 *
 * This is the drop-in: compile it against the reference's public header
 * (-I<reference>/src ... see the Makefile) and link the result in place of
 * sdracq.o, sdrtrk.o and the hot functions of sdrcmn.o.  sdrmain.c, sdrinit.c,
 * sdrspec.c, sdrnav.c ... keep calling the same symbols with the same arguments
 * (prototypes: src/sdr.h:610-619, 650-668, 680-688); the arithmetic now runs on
 * the GPU.  Error behaviour follows the reference: these functions return void
 * (or a buffer location), so a failure is printed with SDRPRINTF and the
 * outputs are left untouched / flagacq, flagtrk stay OFF (src/sdrcmn.c:697-702).
 *
 * Symbols the host program keeps providing: sdrini, sdrstat, hreadmtx,
 * rcvgetbuff(), sdrnavigation(), sleepms().
 *
 * pll(), dll(), cumsumcorr(), clearcumsumcorr() are a few scalar double
 * operations per call on fields of the caller's sdrch_t; as per-call symbols
 * they stay host C (a kernel launch per 20 flops would only add latency).  The
 * batched engine (sdrb200_trk_*) runs the same updates on the device.
 */
#include "sdr.h"
#include "sdrb200.h"

static void fail(const char *fn, int rc)
{
    if (rc) SDRPRINTF("error: %s: libsdrb200 rc=%d %s\n", fn, rc, sdrb200_last_error());
}

/* ------------------------------------------------------------ FFT family */
void cpxfft(fftwf_plan plan, cpx_t *cpx, int n)          /* src/sdrcmn.c:134-150 */
{
    (void)plan;                                          /* always NULL in the reference */
    fail("cpxfft", sdrb200_cpxfft((float *)cpx, n, 0));
}
void cpxifft(fftwf_plan plan, cpx_t *cpx, int n)         /* src/sdrcmn.c:158-175 */
{
    (void)plan;
    fail("cpxifft", sdrb200_cpxfft((float *)cpx, n, 1));
}
void cpxcpx(const short *II, const short *QQ, double scale, int n, cpx_t *cpx)   /* :185-195 */
{
    fail("cpxcpx", sdrb200_cpxcpx(II, QQ, scale, n, (float *)cpx));
}
void cpxconv(fftwf_plan plan, fftwf_plan iplan, cpx_t *cpxa, cpx_t *cpxb, int m, int n,
             int flagsum, double *conv)                  /* src/sdrcmn.c:228-251 */
{
    (void)plan; (void)iplan;
    fail("cpxconv", sdrb200_cpxconv((float *)cpxa, (const float *)cpxb, m, n, flagsum, conv));
}
void cpxpspec(fftwf_plan plan, cpx_t *cpx, int n, int flagsum, double *pspec)    /* :261-276 */
{
    (void)plan;
    fail("cpxpspec", sdrb200_cpxpspec((float *)cpx, n, flagsum, pspec));
}

/* ------------------------------------------------------ correlator family */
void dot_21(const short *a1, const short *a2, const short *b, int n, double *d1, double *d2)
{
    const short *r[1] = {b};
    fail("dot_21", sdrb200_dot2x(a1, a2, r, 1, n, d1, d2));
}
void dot_22(const short *a1, const short *a2, const short *b1, const short *b2, int n,
            double *d1, double *d2)
{
    const short *r[2] = {b1, b2};
    fail("dot_22", sdrb200_dot2x(a1, a2, r, 2, n, d1, d2));
}
void dot_23(const short *a1, const short *a2, const short *b1, const short *b2, const short *b3,
            int n, double *d1, double *d2)
{
    const short *r[3] = {b1, b2, b3};
    fail("dot_23", sdrb200_dot2x(a1, a2, r, 3, n, d1, d2));
}
double mixcarr(const char *data, int dtype, double ti, int n, double freq, double phi0,
               short *II, short *QQ)                     /* src/sdrcmn.c:633-669 */
{
    double rem = 0.0;
    fail("mixcarr", sdrb200_mixcarr((const int8_t *)data, dtype, ti, n, freq, phi0, II, QQ, &rem));
    return rem;
}
double rescode(const short *code, int len, double coff, int smax, double ci, int n, short *rcode)
{
    double rem = 0.0;                                    /* src/sdrcmn.c:608-621 */
    fail("rescode", sdrb200_rescode(code, len, coff, smax, ci, n, rcode, &rem));
    return rem;
}
void correlator(const char *data, int dtype, double ti, int n, double freq, double phi0,
                double crate, double coff, int *s, int ns, double *II, double *QQ,
                double *remc, double *remp, short *codein, int coden)   /* src/sdrcmn.c:687-722 */
{
    fail("correlator", sdrb200_correlator((const int8_t *)data, dtype, ti, n, freq, phi0, crate, coff,
                                          s, ns, II, QQ, remc, remp, codein, coden));
}
void pcorrelator(const char *data, int dtype, double ti, int n, double *freq, int nfreq,
                 double crate, int m, cpx_t *codex, double *P)          /* src/sdrcmn.c:738-773 */
{
    fail("pcorrelator", sdrb200_pcorrelator((const int8_t *)data, dtype, ti, n, freq, nfreq, crate, m,
                                            (const float *)codex, P));
}

/* ------------------------------------------------------------ acquisition */
int checkacquisition(double *P, sdrch_t *sdr)            /* src/sdracq.c:71-95 */
{
    double cn0, peakr;
    int codei, freqi;
    int rc = sdrb200_checkacquisition(P, sdr->nsamp, sdr->acq.nfreq, sdr->nsampchip, sdr->ctime,
                                      &cn0, &peakr, &codei, &freqi);
    if (rc) { fail("checkacquisition", rc); return 0; }
    sdr->acq.cn0 = cn0;
    sdr->acq.peakr = peakr;
    sdr->acq.acqcodei = codei;
    sdr->acq.freqi = freqi;
    sdr->acq.acqfreq = sdr->acq.freq[freqi];
    return sdr->acq.peakr > ACQTH;
}

/* one batched-engine plan per channel struct, created on first use (the grid,
 * code and sample rate of a channel never change after initsdrch)             */
static struct { const sdrch_t *key; sdrb200_acq_t *plan; } g_plan[MAXSAT];
static pthread_mutex_t g_plan_lock = PTHREAD_MUTEX_INITIALIZER;
static __thread int g_last_nintg = 0;
int sdrb200_sdrh_fill_power = 1;        /* copy the power matrix back (the caller plots it) */
int sdrb200_sdrh_last_nintg(void) { return g_last_nintg; }

static sdrb200_acq_t *plan_of(sdrch_t *sdr)
{
    sdrb200_acq_t *plan = NULL;
    int i, slot = -1;
    pthread_mutex_lock(&g_plan_lock);
    for (i = 0; i < MAXSAT; i++) {
        if (g_plan[i].key == sdr) { plan = g_plan[i].plan; break; }
        if (!g_plan[i].key && slot < 0) slot = i;
    }
    if (!plan && slot >= 0) {
        sdrb200_acq_cfg_t c;
        int8_t *code = (int8_t *)malloc((size_t)sdr->clen);
        memset(&c, 0, sizeof(c));
        if (code) {
            for (i = 0; i < sdr->clen; i++) code[i] = (int8_t)sdr->code[i];
            c.f_sf = sdr->f_sf; c.dtype = sdr->dtype; c.nsamp = sdr->nsamp; c.nsampchip = sdr->nsampchip;
            c.ctime = sdr->ctime; c.nfreq = sdr->acq.nfreq; c.freq = sdr->acq.freq; c.nprn = 1;
            c.clen = sdr->clen; c.crate = sdr->crate; c.codes = code; c.intg = sdr->acq.intg;
            c.ncoh = 1; c.threshold = ACQTH;
            if (sdrb200_acq_create(&plan, &c) == SDRB200_OK) {
                g_plan[slot].key = sdr; g_plan[slot].plan = plan;
                if (sdrb200_sdrh_fill_power) sdrb200_acq_keep_power(plan, 1);
            }
            else plan = NULL;
            free(code);
        }
    }
    pthread_mutex_unlock(&g_plan_lock);
    return plan;
}

/* forget (and free) the plan of a channel; call before freesdrch()/re-init */
void sdrb200_sdrh_release(const sdrch_t *sdr)
{
    int i;
    pthread_mutex_lock(&g_plan_lock);
    for (i = 0; i < MAXSAT; i++)
        if (g_plan[i].key == sdr || !sdr) {
            if (g_plan[i].plan) sdrb200_acq_destroy(g_plan[i].plan);
            g_plan[i].key = NULL; g_plan[i].plan = NULL;
        }
    pthread_mutex_unlock(&g_plan_lock);
}

uint64_t sdracquisition(sdrch_t *sdr, double *power)     /* src/sdracq.c:14-62 */
{
    const int nsamp = sdr->nsamp, intg = sdr->acq.intg, dt = sdr->dtype;
    sdrb200_acq_result_t r;
    sdrb200_acq_t *plan;
    uint64_t buffloc, b0;
    char *data;
    int i, rc;

    data = (char *)sdrmalloc(sizeof(char) * (size_t)(intg + 1) * nsamp * dt);
    mlock(hreadmtx);
    buffloc = (sdrstat.fendbuffsize * sdrstat.buffcnt) - (intg + 1) * nsamp;
    unmlock(hreadmtx);
    b0 = buffloc;
    plan = plan_of(sdr);
    if (!data || !plan) {
        SDRPRINTF("error: sdracquisition: %s\n", plan ? "memory allocation" : sdrb200_last_error());
        sdrfree(data);
        return buffloc;
    }
    /* the same windows the reference reads one by one: 2*nsamp samples, hop nsamp */
    for (i = 0; i < intg; i++)
        rcvgetbuff(&sdrini, buffloc + (uint64_t)i * nsamp, 2 * nsamp, sdr->ftype, dt,
                   data + (size_t)i * nsamp * dt);
    if (power && sdrb200_sdrh_fill_power) sdrb200_acq_keep_power(plan, 1);
    rc = sdrb200_acq_run_host(plan, (const int8_t *)data, (uint64_t)(intg + 1) * nsamp, &r);
    if (rc) {
        fail("sdracquisition", rc);
        sdrfree(data);
        return buffloc;
    }
    if (power && sdrb200_sdrh_fill_power) {
        /* the caller's matrix is zero on entry (src/sdrmain.c:252): power = accumulated P */
        fail("sdracquisition(power)", sdrb200_acq_power(plan, 0, power));
    }
    g_last_nintg = r.nintg;
    sdr->acq.cn0 = r.cn0; sdr->acq.peakr = r.peakr;
    sdr->acq.acqcodei = r.codei; sdr->acq.freqi = r.freqi; sdr->acq.acqfreq = r.acqfreq;
    if (r.acquired) sdr->flagacq = ON;

    SDRPRINTF("%s, C/N0=%4.1f, peak=%3.1f, codei=%5d, freq=%8.1f\n", sdr->satstr, sdr->acq.cn0,
              sdr->acq.peakr, sdr->acq.acqcodei, sdr->acq.acqfreq - sdr->f_if - sdr->foffset);
    if (sdr->flagacq) {
        buffloc = b0 + (uint64_t)sdr->acq.acqcodei;      /* top of the code, src/sdracq.c:53 */
        sdr->trk.carrfreq = sdr->acq.acqfreq;
        sdr->trk.codefreq = sdr->crate;
    } else {
        buffloc = b0 + (uint64_t)intg * nsamp;
        sleepms(ACQSLEEP);
    }
    sdrfree(data);
    return buffloc;
}

/* --------------------------------------------------------------- tracking */
uint64_t sdrtracking(sdrch_t *sdr, uint64_t buffloc, uint64_t cnt)   /* src/sdrtrk.c:15-54 */
{
    char *data;
    uint64_t bufflocnow;

    sdr->flagtrk = OFF;
    data = (char *)sdrmalloc(sizeof(char) * (sdr->nsamp + 100) * sdr->dtype);
    mlock(hreadmtx);
    bufflocnow = sdrstat.fendbuffsize * sdrstat.buffcnt - sdr->nsamp;
    unmlock(hreadmtx);

    if (data && bufflocnow > buffloc) {
        double II[SDRB200_MAXTAPS], QQ[SDRB200_MAXTAPS], remc, remp;
        const int nt = 1 + 2 * sdr->trk.corrn;
        int rc;
        sdr->currnsamp = (int)((sdr->clen - sdr->trk.remcode) / (sdr->trk.codefreq / sdr->f_sf));
        rcvgetbuff(&sdrini, buffloc, sdr->currnsamp, sdr->ftype, sdr->dtype, data);
        /* same byte count as the reference's memcpy (src/sdrtrk.c:35-36) */
        memcpy(sdr->trk.oldI, sdr->trk.II, 1 + 2 * sdr->trk.corrn * sizeof(double));
        memcpy(sdr->trk.oldQ, sdr->trk.QQ, 1 + 2 * sdr->trk.corrn * sizeof(double));
        sdr->trk.oldremcode = sdr->trk.remcode;
        sdr->trk.oldremcarr = sdr->trk.remcarr;
        rc = sdrb200_correlator((const int8_t *)data, sdr->dtype, sdr->ti, sdr->currnsamp,
                                sdr->trk.carrfreq, sdr->trk.oldremcarr, sdr->trk.codefreq,
                                sdr->trk.oldremcode, sdr->trk.corrp, sdr->trk.corrn, II, QQ,
                                &remc, &remp, sdr->code, sdr->clen);
        if (rc) {
            fail("sdrtracking", rc);                     /* flagtrk stays OFF */
        } else {
            /* cosine sums go to trk.QQ, sine sums to trk.II (argument order of src/sdrtrk.c:40-43) */
            memcpy(sdr->trk.QQ, II, sizeof(double) * nt);
            memcpy(sdr->trk.II, QQ, sizeof(double) * nt);
            sdr->trk.remcode = remc;
            sdr->trk.remcarr = remp;
            sdrnavigation(sdr, buffloc, cnt);
            sdr->flagtrk = ON;
        }
    } else {
        sleepms(1);
    }
    sdrfree(data);
    return bufflocnow;
}

void cumsumcorr(sdrtrk_t *trk, int polarity)             /* src/sdrtrk.c:64-76 */
{
    int k;
    const int nt = 1 + 2 * trk->corrn;
    for (k = 0; k < nt; k++) {
        trk->II[k] *= polarity;
        trk->QQ[k] *= polarity;
        trk->oldsumI[k] += trk->oldI[k];
        trk->oldsumQ[k] += trk->oldQ[k];
        trk->sumI[k] += trk->II[k];
        trk->sumQ[k] += trk->QQ[k];
    }
}
void clearcumsumcorr(sdrtrk_t *trk)                      /* src/sdrtrk.c:77-86 */
{
    const size_t nb = sizeof(double) * (size_t)(1 + 2 * trk->corrn);
    memset(trk->oldsumI, 0, nb); memset(trk->oldsumQ, 0, nb);
    memset(trk->sumI, 0, nb);    memset(trk->sumQ, 0, nb);
}
void pll(sdrch_t *sdr, sdrtrkprm_t *prm, double dt)      /* src/sdrtrk.c:94-125 */
{
    sdrtrk_t *t = &sdr->trk;
    const double ip = t->sumI[0], qp = t->sumQ[0], ip0 = t->oldsumI[0], qp0 = t->oldsumQ[0];
    const double perr = (ip > 0 ? atan2(qp, ip) : atan2(-qp, -ip)) / PI;
    const double a1 = (ip == 0) ? PI / 2 : atan(qp / ip);
    const double a0 = (ip0 == 0) ? PI / 2 : atan(qp0 / ip0);
    double ferr = a1 - a0;
    if (ferr > PI / 2) ferr = PI - ferr;
    if (ferr < -PI / 2) ferr = -PI - ferr;
    t->carrNco += prm->pllaw * (perr - t->carrErr) + prm->pllw2 * dt * perr + prm->fllw * dt * ferr;
    t->carrfreq = sdr->acq.acqfreq + t->carrNco;
    t->carrErr = perr;
    t->freqErr = ferr;
}
void dll(sdrch_t *sdr, sdrtrkprm_t *prm, double dt)      /* src/sdrtrk.c:133-150 */
{
    sdrtrk_t *t = &sdr->trk;
    const double ie = t->sumI[t->ne], il = t->sumI[t->nl], qe = t->sumQ[t->ne], ql = t->sumQ[t->nl];
    const double cerr = (sqrt(ie * ie + qe * qe) - sqrt(il * il + ql * ql)) /
                        (sqrt(ie * ie + qe * qe) + sqrt(il * il + ql * ql));
    t->codeNco += prm->dllaw * (cerr - t->codeErr) + prm->dllw2 * dt * cerr;
    t->codefreq = sdr->crate - t->codeNco + (t->carrfreq - sdr->f_if - sdr->foffset) / (sdr->f_cf / sdr->crate);
    t->codeErr = cerr;
}
