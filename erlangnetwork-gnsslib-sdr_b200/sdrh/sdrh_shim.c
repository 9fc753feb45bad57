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

/* GPU of a channel: channel number modulo the GPUs of the box (sdrthread runs one pthread per
 * channel, src/sdrmain.c:144-149; the CUDA current device is per thread)                      */
static int g_ndev = -1;
static int device_of(const sdrch_t *sdr)
{
    int dev;
    if (g_ndev < 0) { int n = sdrb200_device_count(); g_ndev = n > 0 ? n : 0; }
    if (g_ndev <= 0) return -1;
    dev = (sdr->no > 0 ? sdr->no - 1 : 0) % g_ndev;
    sdrb200_set_device(dev);
    return dev;
}

/* what a cached plan / tracking engine was built from: the reference re-initialises sdrch[] slots in
 * place (src/sdrmain.c:27, startsdr/quitsdr) and never tells this library, so every use compares */
typedef struct { int nsamp, nfreq, intg, dtype, clen, corrn, no; double f_sf, f0, f1, crate; uint64_t codehash; } chan_print_t;
static void chan_print(const sdrch_t *sdr, chan_print_t *fp)
{
    uint64_t h = 1469598103934665603ULL;
    int i;
    memset(fp, 0, sizeof(*fp));
    fp->nsamp = sdr->nsamp; fp->nfreq = sdr->acq.nfreq; fp->intg = sdr->acq.intg; fp->dtype = sdr->dtype;
    fp->clen = sdr->clen; fp->corrn = sdr->trk.corrn; fp->no = sdr->no; fp->f_sf = sdr->f_sf; fp->crate = sdr->crate;
    fp->f0 = (sdr->acq.freq && sdr->acq.nfreq > 0) ? sdr->acq.freq[0] : 0.0;
    fp->f1 = (sdr->acq.freq && sdr->acq.nfreq > 0) ? sdr->acq.freq[sdr->acq.nfreq - 1] : 0.0;
    for (i = 0; sdr->code && i < sdr->clen; i++) { h ^= (uint64_t)(uint16_t)sdr->code[i]; h *= 1099511628211ULL; }
    fp->codehash = h;
}

/* one batched-engine plan per channel struct, created on first use and re-created when the
 * channel's parameters have changed                                                          */
static struct { const sdrch_t *key; sdrb200_acq_t *plan; chan_print_t fp; } g_plan[MAXSAT];
static pthread_mutex_t g_plan_lock = PTHREAD_MUTEX_INITIALIZER;
static __thread int g_last_nintg = 0;
int sdrb200_sdrh_fill_power = 1;        /* copy the power matrix back (the caller plots it) */
int sdrb200_sdrh_last_nintg(void) { return g_last_nintg; }

static sdrb200_acq_t *plan_of(sdrch_t *sdr, const char **why)
{
    sdrb200_acq_t *plan = NULL;
    chan_print_t fp;
    int i, slot = -1;
    chan_print(sdr, &fp);
    pthread_mutex_lock(&g_plan_lock);
    for (i = 0; i < MAXSAT; i++) {
        if (g_plan[i].key == sdr) {
            if (memcmp(&g_plan[i].fp, &fp, sizeof(fp)) == 0) plan = g_plan[i].plan;
            else {                                       /* the slot was re-initialised: forget the old plan */
                if (g_plan[i].plan) sdrb200_acq_destroy(g_plan[i].plan);
                g_plan[i].key = NULL; g_plan[i].plan = NULL;
                if (slot < 0) slot = i;
            }
            break;
        }
        if (!g_plan[i].key && slot < 0) slot = i;
    }
    if (!plan && slot < 0) for (i = 0; i < MAXSAT; i++) if (!g_plan[i].key) { slot = i; break; }
    if (!plan && slot < 0) *why = "no free plan slot (MAXSAT channels cached; call sdrb200_sdrh_release)";
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
                g_plan[slot].key = sdr; g_plan[slot].plan = plan; g_plan[slot].fp = fp;
                if (sdrb200_sdrh_fill_power) sdrb200_acq_keep_power(plan, 1);
            }
            else { plan = NULL; *why = sdrb200_last_error(); }
            free(code);
        } else *why = "memory allocation";
    }
    pthread_mutex_unlock(&g_plan_lock);
    return plan;
}

static void trk_release(const sdrch_t *sdr);

/* forget (and free) the plan and the tracking engine of a channel; call before freesdrch()/re-init
 * (not required: both are rebuilt when the channel's parameters change)                       */
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
    trk_release(sdr);
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
    device_of(sdr);
    {
        const char *why = "memory allocation";
        plan = data ? plan_of(sdr, &why) : NULL;
        if (!data || !plan) {
            SDRPRINTF("error: sdracquisition: %s\n", why);
            sdrfree(data);
            sleepms(ACQSLEEP);                           /* no progress: wait as the reference does, src/sdracq.c:58-60 */
            return buffloc;
        }
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
        sleepms(ACQSLEEP);
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

/* --------------------------------------------------------------- tracking
 * sdrthread calls sdrtracking() once per code period and then cumsumcorr/pll/dll on the result
 * (src/sdrmain.c:264-312).  One blocking H2D -> kernel -> D2H per period would make the GPU a slow
 * coprocessor, so the drop-in runs the BATCHED engine (sdrb200_trk_*: the same correlator, bit sync
 * and loop filters on the device, parity-tested against the reference) a chunk of periods AHEAD on
 * the samples that have already arrived, and serves the per-period calls from its records:
 *   sdrtracking  <- currnsamp, II, QQ, remcode, remcarr of the record whose (cnt, buffloc) match
 *   pll / dll    <- the loop-filter state the device computed for that period
 * sdrnavigation, cumsumcorr, setobsdata, writelog stay the host program's.  The IF reaches the device
 * through ONE mirror stream per GPU and front-end type, fed with rcvgetbuff() (so every front end's
 * own unpacking applies) as the receive thread advances sdrstat.buffcnt.  Anything unexpected (a
 * record that does not match the caller's counters, an engine error) drops the channel back to the
 * per-period path below for good.                                                             */
#define LA_CHUNK   250                       /* code periods per engine launch            */
#define LA_MAXDEV  16
#define LA_MIRROR  (64u << 20)               /* samples held by a mirror stream           */
#define LA_PUSH    (1u << 20)                /* samples per rcvgetbuff + push             */
int sdrb200_sdrh_lookahead = 1;              /* 0: always the per-period path             */
static unsigned long long g_la_served = 0;   /* code periods served from engine records   */
unsigned long long sdrb200_sdrh_lookahead_served(void) { return __atomic_load_n(&g_la_served, __ATOMIC_RELAXED); }

typedef struct {
    const sdrch_t *key; chan_print_t fp;
    sdrb200_trk_t *eng; sdrb200_trk_rec_t *rec;
    int nrec, pos, dead, final, dev, ftype;
    int8_t *code;
} la_t;
static la_t g_la[MAXSAT];
static pthread_mutex_t g_la_lock = PTHREAD_MUTEX_INITIALIZER;
typedef struct { sdrb200_stream_t *st; uint64_t count, start; int dtype; pthread_mutex_t lock; int init; } mirror_t;
static mirror_t g_mirror[LA_MAXDEV][2];
static __thread const sdrch_t *g_cur_sdr = NULL;      /* record the calling thread's last sdrtracking served */
static __thread const sdrb200_trk_rec_t *g_cur_rec = NULL;

static void trk_release(const sdrch_t *sdr)
{
    int i;
    pthread_mutex_lock(&g_la_lock);
    for (i = 0; i < MAXSAT; i++)
        if (g_la[i].key == sdr || !sdr) {
            if (g_la[i].eng) sdrb200_trk_destroy(g_la[i].eng);
            free(g_la[i].rec); free(g_la[i].code);
            memset(&g_la[i], 0, sizeof(g_la[i]));
        }
    if (!sdr) {                                          /* everything: the mirrors of the host ring too */
        int d, f;
        for (d = 0; d < LA_MAXDEV; d++)
            for (f = 0; f < 2; f++)
                if (g_mirror[d][f].st) {
                    sdrb200_set_device(d);
                    sdrb200_stream_destroy(g_mirror[d][f].st);
                    g_mirror[d][f].st = NULL; g_mirror[d][f].count = 0; g_mirror[d][f].start = 0;
                }
    }
    pthread_mutex_unlock(&g_la_lock);
}

/* bring the mirror of (device, front-end type) up to `upto` samples; returns the stream or NULL */
static sdrb200_stream_t *mirror_update(int dev, int ftype, int dtype, uint64_t first, uint64_t upto)
{
    mirror_t *m;
    char *tmp = NULL;
    if (dev < 0 || dev >= LA_MAXDEV || ftype < 1 || ftype > 2) return NULL;
    m = &g_mirror[dev][ftype - 1];
    pthread_mutex_lock(&g_la_lock);
    if (!m->init) { pthread_mutex_init(&m->lock, NULL); m->init = 1; }
    pthread_mutex_unlock(&g_la_lock);
    pthread_mutex_lock(&m->lock);
    if (!m->st) {
        if (sdrb200_stream_create(&m->st, dtype, LA_MIRROR) != SDRB200_OK) { m->st = NULL; pthread_mutex_unlock(&m->lock); return NULL; }
        m->dtype = dtype;
        /* the mirror starts at the block of the first sample anybody asks for (a receiver that has run
         * for hours does not re-send its past) */
        m->start = first & ~(uint64_t)65535;
        m->count = m->start;
        if (m->start) sdrb200_stream_skip(m->st, m->start);
    }
    if (m->dtype != dtype || first < m->start || upto < m->count) { pthread_mutex_unlock(&m->lock); return NULL; }
    while (m->count < upto) {
        uint64_t n = upto - m->count;
        if (n > LA_PUSH) n = LA_PUSH;
        if (!tmp) tmp = (char *)malloc((size_t)LA_PUSH * dtype);
        if (!tmp || rcvgetbuff(&sdrini, m->count, (int)n, ftype, dtype, tmp) < 0 ||
            sdrb200_stream_push_host(m->st, (const int8_t *)tmp, n) != SDRB200_OK) {
            free(tmp); pthread_mutex_unlock(&m->lock); return NULL;
        }
        m->count += n;
    }
    free(tmp);
    pthread_mutex_unlock(&m->lock);
    return m->st;
}

static la_t *la_of(sdrch_t *sdr, uint64_t buffloc, uint64_t cnt, int dev)
{
    la_t *la = NULL;
    chan_print_t fp;
    int i, slot = -1;
    chan_print(sdr, &fp);
    pthread_mutex_lock(&g_la_lock);
    for (i = 0; i < MAXSAT; i++) {
        if (g_la[i].key == sdr) {
            if (memcmp(&g_la[i].fp, &fp, sizeof(fp)) == 0 && cnt != 0) la = &g_la[i];
            else {                                       /* re-initialised channel, or a (re)start from cnt = 0 */
                if (g_la[i].eng) sdrb200_trk_destroy(g_la[i].eng);
                free(g_la[i].rec); free(g_la[i].code);
                memset(&g_la[i], 0, sizeof(g_la[i]));
                slot = i;
            }
            break;
        }
        if (!g_la[i].key && slot < 0) slot = i;
    }
    if (!la && slot >= 0) {
        la = &g_la[slot];
        memset(la, 0, sizeof(*la));
        la->key = sdr; la->fp = fp; la->dev = dev; la->ftype = sdr->ftype;
        /* the engine starts from the hand-over state of sdracquisition (src/sdracq.c:51-56): only
         * a channel that has not tracked yet can be taken over                                    */
        if (cnt != 0 || sdr->trk.remcode != 0.0 || sdr->trk.remcarr != 0.0 || sdr->trk.corrn > SDRB200_MAXCORRN) la->dead = 1;
    }
    pthread_mutex_unlock(&g_la_lock);
    if (la && !la->dead && !la->eng) {
        sdrb200_trk_chan_t c;
        memset(&c, 0, sizeof(c));
        la->code = (int8_t *)malloc((size_t)sdr->clen);
        la->rec = (sdrb200_trk_rec_t *)malloc(sizeof(sdrb200_trk_rec_t) * LA_CHUNK);
        if (!la->code || !la->rec) { la->dead = 1; return la; }
        for (i = 0; i < sdr->clen; i++) la->code[i] = (int8_t)sdr->code[i];
        c.f_sf = sdr->f_sf; c.f_if = sdr->f_if; c.f_cf = sdr->f_cf; c.foffset = sdr->foffset;
        c.dtype = sdr->dtype; c.clen = sdr->clen; c.crate = sdr->crate; c.ctime = sdr->ctime; c.nsamp = sdr->nsamp;
        c.prn = sdr->nav.sdreph.prn;                     /* the bit-sync rule tests this, src/sdrnav.c:203 */
        c.corrn = sdr->trk.corrn;
        for (i = 0; i < sdr->trk.corrn; i++) c.corrp[i] = sdr->trk.corrp[i];
        c.ne = sdr->trk.ne; c.nl = sdr->trk.nl; c.loopms = sdr->trk.loopms; c.navrate = sdr->nav.rate;
        c.prm1.dllw2 = sdr->trk.prm1.dllw2; c.prm1.dllaw = sdr->trk.prm1.dllaw; c.prm1.pllw2 = sdr->trk.prm1.pllw2;
        c.prm1.pllaw = sdr->trk.prm1.pllaw; c.prm1.fllw = sdr->trk.prm1.fllw;
        c.prm2.dllw2 = sdr->trk.prm2.dllw2; c.prm2.dllaw = sdr->trk.prm2.dllaw; c.prm2.pllw2 = sdr->trk.prm2.pllw2;
        c.prm2.pllaw = sdr->trk.prm2.pllaw; c.prm2.fllw = sdr->trk.prm2.fllw;
        c.code = la->code; c.buffloc = buffloc; c.acqfreq = sdr->acq.acqfreq;
        if (sdr->trk.carrfreq != sdr->acq.acqfreq || sdr->trk.codefreq != sdr->crate ||
            sdrb200_trk_create(&la->eng, 1, &c) != SDRB200_OK) { la->eng = NULL; la->dead = 1; }
    }
    return la;
}

/* the per-period path: one correlator call on the host's copy of the window */
static int track_one_period(sdrch_t *sdr, uint64_t buffloc, uint64_t cnt, char *data)
{
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
    if (rc) { fail("sdrtracking", rc); return rc; }      /* flagtrk stays OFF */
    /* cosine sums go to trk.QQ, sine sums to trk.II (argument order of src/sdrtrk.c:40-43) */
    memcpy(sdr->trk.QQ, II, sizeof(double) * nt);
    memcpy(sdr->trk.II, QQ, sizeof(double) * nt);
    sdr->trk.remcode = remc;
    sdr->trk.remcarr = remp;
    sdrnavigation(sdr, buffloc, cnt);
    sdr->flagtrk = ON;
    return 0;
}

uint64_t sdrtracking(sdrch_t *sdr, uint64_t buffloc, uint64_t cnt)   /* src/sdrtrk.c:15-54 */
{
    char *data;
    uint64_t bufflocnow, have;
    la_t *la = NULL;
    int dev;

    sdr->flagtrk = OFF;
    g_cur_sdr = NULL; g_cur_rec = NULL;
    data = (char *)sdrmalloc(sizeof(char) * (sdr->nsamp + 100) * sdr->dtype);
    mlock(hreadmtx);
    have = sdrstat.fendbuffsize * sdrstat.buffcnt;
    bufflocnow = have - sdr->nsamp;
    unmlock(hreadmtx);
    dev = device_of(sdr);

    if (data && bufflocnow > buffloc) {
        if (sdrb200_sdrh_lookahead && dev >= 0) la = la_of(sdr, buffloc, cnt, dev);
        if (la && !la->dead) {
            if (la->pos >= la->nrec) {                   /* records used up: run the engine on what has arrived */
                sdrb200_stream_t *ms = mirror_update(dev, sdr->ftype, sdr->dtype, buffloc, have);
                int n = 0;
                int rc = SDRB200_OK;
                la->pos = 0; la->nrec = 0;
                if (la->final || !ms) la->dead = 1;
                else rc = sdrb200_trk_run(la->eng, ms, LA_CHUNK, la->rec, &n);
                if (!la->dead && rc == SDRB200_ERANGE && n > 0) {
                    /* the engine stopped inside the chunk (a period it cannot run: the loops diverged, e.g. on
                     * a dead signal): its records up to there are served, then the per-period path takes over */
                    la->nrec = n; la->final = 1;
                } else if (!la->dead && rc != SDRB200_OK) {
                    fail("sdrtracking(lookahead)", rc);
                    la->dead = 1;
                } else if (!la->dead) la->nrec = n;
            }
            if (!la->dead && la->nrec > 0) {
                const sdrb200_trk_rec_t *r = &la->rec[la->pos];
                if (r->cnt != cnt || r->buffloc != buffloc) la->dead = 1;      /* the caller is somewhere else */
                else {
                    const int nt = 1 + 2 * sdr->trk.corrn;
                    sdr->currnsamp = r->currnsamp;
                    memcpy(sdr->trk.oldI, sdr->trk.II, 1 + 2 * sdr->trk.corrn * sizeof(double));
                    memcpy(sdr->trk.oldQ, sdr->trk.QQ, 1 + 2 * sdr->trk.corrn * sizeof(double));
                    sdr->trk.oldremcode = sdr->trk.remcode;
                    sdr->trk.oldremcarr = sdr->trk.remcarr;
                    memcpy(sdr->trk.II, r->II, sizeof(double) * nt);
                    memcpy(sdr->trk.QQ, r->QQ, sizeof(double) * nt);
                    sdr->trk.remcode = r->remcode;
                    sdr->trk.remcarr = r->remcarr;
                    sdrnavigation(sdr, buffloc, cnt);
                    sdr->flagtrk = ON;
                    g_cur_sdr = sdr; g_cur_rec = r;
                    la->pos++;
                    __atomic_add_fetch(&g_la_served, 1ULL, __ATOMIC_RELAXED);
                }
            } else if (!la->dead) {
                sleepms(1);                              /* the window has not arrived on the device yet */
            }
        }
        if (!la || la->dead) {
            if (track_one_period(sdr, buffloc, cnt, data)) sleepms(1);
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
    if (g_cur_sdr == sdr && g_cur_rec && g_cur_rec->flagloopfilter) {
        /* the device ran this period's loop filter on the same sums: take its state, so that the
         * host's frequencies are the ones the engine correlates the next periods with           */
        const sdrb200_trk_rec_t *r = g_cur_rec;
        (void)prm; (void)dt;
        t->carrNco = r->carrNco; t->carrfreq = r->carrfreq; t->carrErr = r->carrErr; t->freqErr = r->freqErr;
        return;
    }
    if (g_cur_sdr == sdr && g_cur_rec) {                  /* the host filters where the device did not: part ways */
        int i;
        for (i = 0; i < MAXSAT; i++) if (g_la[i].key == sdr) g_la[i].dead = 1;
    }
    {
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
}
void dll(sdrch_t *sdr, sdrtrkprm_t *prm, double dt)      /* src/sdrtrk.c:133-150 */
{
    sdrtrk_t *t = &sdr->trk;
    if (g_cur_sdr == sdr && g_cur_rec && g_cur_rec->flagloopfilter) {
        const sdrb200_trk_rec_t *r = g_cur_rec;
        (void)prm; (void)dt;
        t->codeNco = r->codeNco; t->codefreq = r->codefreq; t->codeErr = r->codeErr;
        return;
    }
    {
    const double ie = t->sumI[t->ne], il = t->sumI[t->nl], qe = t->sumQ[t->ne], ql = t->sumQ[t->nl];
    const double cerr = (sqrt(ie * ie + qe * qe) - sqrt(il * il + ql * ql)) /
                        (sqrt(ie * ie + qe * qe) + sqrt(il * il + ql * ql));
    t->codeNco += prm->dllaw * (cerr - t->codeErr) + prm->dllw2 * dt * cerr;
    t->codefreq = sdr->crate - t->codeNco + (t->carrfreq - sdr->f_if - sdr->foffset) / (sdr->f_cf / sdr->crate);
    t->codeErr = cerr;
    }
}
