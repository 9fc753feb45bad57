/* replay.c - deterministic single-threaded replay of the reference's channel
 * thread (TEST INFRASTRUCTURE; compiled only where /root/reference exists).
 *
 * It is compiled against the reference's own sdr.h and calls the reference's
 * own functions (initsdrch, sdracquisition, sdrtracking, cumsumcorr, pll, dll,
 * setobsdata, clearcumsumcorr); the only thing restated here is the loop body
 * of sdrthread() (src/sdrmain.c:247-315) with the sleeps, plotting and log
 * file removed, and the file front end's ring-buffer feed
 * (file_pushtomembuf, src/sdrrcv.c:469-495) with memory as the source.
 *
 * The same source is linked twice (oracle/Makefile):
 *   libreplay_ref.so   hot symbols from the reference objects   -> the oracle
 *   libreplay_b200.so  hot symbols from libsdrb200_sdrh.so      -> drop-in proof
 *
 * Link-level adjustments (no reference source is modified):
 *   -Wl,--wrap=sleepms      the 2 s ACQSLEEP (src/sdracq.c:58) and 1 ms
 *                           starvation sleep (src/sdrtrk.c:50) become no-ops
 *   -Wl,--wrap=pcorrelator  counts integrations used before the early exit
 */
#include "sdr.h"
#include "harness.h"

#ifndef OH_IMPL_NAME
#define OH_IMPL_NAME "reference"
#endif

const char *oh_impl(void) { return OH_IMPL_NAME; }

/* ---- link-time wrappers ------------------------------------------------ */
void __wrap_sleepms(int ms) { (void)ms; }

static int g_pcorr_calls = 0;
#ifndef OH_NO_PCORR_WRAP
extern void __real_pcorrelator(const char *data, int dtype, double ti, int n,
                               double *freq, int nfreq, double crate, int m,
                               cpx_t *codex, double *P);
void __wrap_pcorrelator(const char *data, int dtype, double ti, int n,
                        double *freq, int nfreq, double crate, int m,
                        cpx_t *codex, double *P)
{
    g_pcorr_calls++;
    __real_pcorrelator(data, dtype, ti, n, freq, nfreq, crate, m, codex, P);
}
#else
extern int sdrb200_sdrh_last_nintg(void);
#endif

/* ---- ring-buffer feed (memory-backed file front end) ------------------- */
static const signed char *g_src = NULL;
static uint64_t g_srclen = 0;       /* samples */
static uint64_t g_ringblocks = 0;   /* blocks allocated in sdrstat.buff */
static int g_handles = 0;

static void ring_release(void)
{
    free(sdrstat.buff);
    sdrstat.buff = NULL;
    g_ringblocks = 0;
}

static int ring_setup(const oh_cfg_t *cfg, const signed char *src, uint64_t nsamples)
{
    uint64_t blocks = (nsamples + FILE_BUFFSIZE - 1) / FILE_BUFFSIZE + 2;
    if (blocks > MEMBUFFLEN) blocks = MEMBUFFLEN;
    ring_release();
#ifdef OH_NO_PCORR_WRAP
    /* drop-in build: a new "recording" starts - what quitsdr()/startsdr() mean for a receiver.  The
     * product keeps device-side state per channel and a mirror of the ring: tell it to start over */
    { extern void sdrb200_sdrh_release(const void *); sdrb200_sdrh_release(NULL); }
#endif
    sdrstat.buff = (unsigned char *)malloc((size_t)cfg->dtype * FILE_BUFFSIZE * blocks);
    if (!sdrstat.buff) return -1;
    g_ringblocks = blocks;
    g_src = src; g_srclen = nsamples;
    sdrstat.fendbuffsize = FILE_BUFFSIZE;
    sdrstat.buffsize = FILE_BUFFSIZE * MEMBUFFLEN;
    sdrstat.buffcnt = 0;
    sdrstat.stopflag = 0;
    return 0;
}

/* push blocks until buffcnt == target (what file_pushtomembuf does per call) */
static void ring_feed(const oh_cfg_t *cfg, uint64_t target)
{
    const size_t blk = (size_t)cfg->dtype * FILE_BUFFSIZE;
    while (sdrstat.buffcnt < target) {
        uint64_t b = sdrstat.buffcnt;
        unsigned char *dst = &sdrstat.buff[(b % MEMBUFFLEN) % g_ringblocks * blk];
        uint64_t s0 = b * FILE_BUFFSIZE;
        size_t have = 0;
        if (s0 < g_srclen) {
            uint64_t ns = g_srclen - s0;
            if (ns > FILE_BUFFSIZE) ns = FILE_BUFFSIZE;
            have = (size_t)ns * cfg->dtype;
            memcpy(dst, g_src + s0 * cfg->dtype, have);
        }
        if (have < blk) memset(dst + have, 0, blk - have);
        sdrstat.buffcnt++;
    }
}

static void setup_ini(const oh_cfg_t *cfg)
{
    memset(&sdrini, 0, sizeof(sdrini));
    sdrini.fend = FEND_FILE;
    sdrini.f_cf[0] = cfg->f_cf; sdrini.f_sf[0] = cfg->f_sf; sdrini.f_if[0] = cfg->f_if;
    sdrini.dtype[0] = cfg->dtype;
    sdrini.useif1 = 1;
    sdrini.trkcorrn = cfg->corrn; sdrini.trkcorrd = cfg->corrd; sdrini.trkcorrp = cfg->corrp;
    sdrini.trkdllb[0] = cfg->dllb[0]; sdrini.trkdllb[1] = cfg->dllb[1];
    sdrini.trkpllb[0] = cfg->pllb[0]; sdrini.trkpllb[1] = cfg->pllb[1];
    sdrini.trkfllb[0] = cfg->fllb[0]; sdrini.trkfllb[1] = cfg->fllb[1];
    if (!g_handles) { openhandles(); g_handles = 1; }
}

static int sys_of(int ctype)
{
    if (ctype == CTYPE_G1) return SYS_GLO;
    if (ctype == CTYPE_L1SBAS) return SYS_SBS;
    return SYS_GPS;
}

/* initsdrch + optional override of the compile-time acquisition grid,
 * following the grid formula of src/sdrinit.c:633-635 */
static int channel_init(const oh_cfg_t *cfg, int chno, int prn, int ctype, sdrch_t *sdr)
{
    int i;
    memset(sdr, 0, sizeof(*sdr));
    if (initsdrch(chno, sys_of(ctype), prn, ctype, cfg->dtype, FTYPE1,
                  cfg->f_cf, cfg->f_sf, cfg->f_if, sdr) < 0) return -1;
    if (cfg->acq_intg > 0) sdr->acq.intg = cfg->acq_intg;
    if (cfg->acq_step > 0.0 && cfg->acq_hband > 0.0) {
        sdr->acq.hband = cfg->acq_hband;
        sdr->acq.step = cfg->acq_step;
        sdr->acq.nfreq = 2 * (int)(cfg->acq_hband / cfg->acq_step) + 1;
        free(sdr->acq.freq);
        sdr->acq.freq = (double *)malloc(sizeof(double) * sdr->acq.nfreq);
        if (!sdr->acq.freq) return -1;
        for (i = 0; i < sdr->acq.nfreq; i++)
            sdr->acq.freq[i] = sdr->f_if + ((i - (sdr->acq.nfreq - 1) / 2) * sdr->acq.step)
                               + sdr->foffset;
    }
    return 0;
}

static uint64_t do_acquire(sdrch_t *sdr, oh_acq_t *o, double *power_out)
{
    size_t np = (size_t)sdr->nsamp * sdr->acq.nfreq;
    double *power = (double *)calloc(sizeof(double), np);   /* src/sdrmain.c:252 */
    uint64_t buffloc;
    g_pcorr_calls = 0;
    buffloc = sdracquisition(sdr, power);
    o->prn = sdr->prn; o->ctype = sdr->ctype;
    o->acquired = sdr->flagacq;
    o->freqi = sdr->acq.freqi; o->codei = sdr->acq.acqcodei;
    o->acqfreq = sdr->acq.acqfreq; o->peakr = sdr->acq.peakr; o->cn0 = sdr->acq.cn0;
    o->maxp = power[(size_t)sdr->acq.freqi * sdr->nsamp + sdr->acq.acqcodei];
#ifndef OH_NO_PCORR_WRAP
    o->nintg_used = g_pcorr_calls;
#else
    o->nintg_used = sdrb200_sdrh_last_nintg();
#endif
    o->buffloc = buffloc;
    if (power_out) memcpy(power_out, power, sizeof(double) * np);
    free(power);
    return buffloc;
}

int oh_acquire(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
               uint64_t avail, int nprn, const int *prn, const int *ctype,
               oh_acq_t *out, double *power_out)
{
    int i;
    if (avail % FILE_BUFFSIZE || avail > nsamples) return -2;
    setup_ini(cfg);
    if (ring_setup(cfg, if_data, nsamples) < 0) return -1;
    ring_feed(cfg, avail / FILE_BUFFSIZE);
    for (i = 0; i < nprn; i++) {
        sdrch_t *sdr = &sdrch[0];
        memset(&out[i], 0, sizeof(out[i]));
        if (channel_init(cfg, 1, prn[i], ctype[i], sdr) < 0) { ring_release(); return -3; }
        do_acquire(sdr, &out[i],
                   power_out ? power_out + (size_t)i * sdr->nsamp * sdr->acq.nfreq : NULL);
        freesdrch(sdr);
    }
    ring_release();
    return 0;
}

/* one code period of one channel: the body of the while loop of sdrthread()
 * (src/sdrmain.c:264-312) plus the record; returns 1 if a period was tracked.
 * rec points at the channel's own row of records.                                */
/* optional: the reference's own tracking log (writelog, src/sdrout.c:412-436), one file for the
 * channel whose number is g_log_ch, written where sdrthread writes it (src/sdrmain.c:307-308) */
static FILE *g_log_fp = NULL;
static int   g_log_ch = -1;
int oh_set_logfile(const char *path, int chno)
{
    if (g_log_fp) { fclose(g_log_fp); g_log_fp = NULL; }
    g_log_ch = -1;
    if (!path) return 0;
    g_log_fp = fopen(path, "w");
    if (!g_log_fp) return -1;
    g_log_ch = chno;
    return 0;
}

/* the interpolation stocks setobsdata fills (src/sdrtrk.c:160-209), kept after a run for oh_last_stocks */
static oh_stocks_t g_stocks[MAXSAT];
static void keep_stocks(const sdrch_t *sdr, int i)
{
    oh_stocks_t *o = &g_stocks[i];
    memcpy(o->L, sdr->trk.L, sizeof(o->L)); memcpy(o->D, sdr->trk.D, sizeof(o->D));
    memcpy(o->S, sdr->trk.S, sizeof(o->S)); memcpy(o->remcout, sdr->trk.remcout, sizeof(o->remcout));
    memcpy(o->codei, sdr->trk.codei, sizeof(o->codei)); memcpy(o->cntout, sdr->trk.cntout, sizeof(o->cntout));
    memcpy(o->codeisum, sdr->trk.codeisum, sizeof(o->codeisum));
}
int oh_last_stocks(int i, oh_stocks_t *out)
{
    if (i < 0 || i >= MAXSAT || !out) return -1;
    *out = g_stocks[i];
    return 0;
}
/* what syncthread does with the stocks of one channel (src/sdrsync.c:96-121): carrier phase and
 * Doppler interpolated at the reference sample, with the reference's own uint64todouble/interp1 */
int oh_interp_obs(const oh_stocks_t *st, uint64_t sampbase, double samprefd, double *L, double *D)
{
    double codeid[OBSINTERPN];
    oh_stocks_t t = *st;
    uint64todouble(t.codei, sampbase, OBSINTERPN, codeid);
    *L = interp1(codeid, t.L, OBSINTERPN, samprefd);
    *D = interp1(codeid, t.D, OBSINTERPN, samprefd);
    return 0;
}

static int track_period(sdrch_t *sdr, uint64_t *buffloc, uint64_t *cnt, uint64_t *loopcnt,
                        oh_trkrec_t *rec, int *ntracked)
{
    oh_trkrec_t *r;
    int k, flagobs = 0;
    sdrtracking(sdr, (*buffloc), (*cnt));
    if (!sdr->flagtrk) return 0;                 /* cannot happen after the feed */

    cumsumcorr(&sdr->trk, sdr->nav.ocode[sdr->nav.ocodei]);
    sdr->trk.flagloopfilter = 0;
    flagobs = 0;
    if (!sdr->nav.flagsync) {
        pll(sdr, &sdr->trk.prm1, sdr->ctime);
        dll(sdr, &sdr->trk.prm1, sdr->ctime);
        sdr->trk.flagloopfilter = 1;
    } else if (sdr->nav.swloop) {
        pll(sdr, &sdr->trk.prm2, (double)sdr->trk.loopms / 1000);
        dll(sdr, &sdr->trk.prm2, (double)sdr->trk.loopms / 1000);
        sdr->trk.flagloopfilter = 2;
        if ((*loopcnt) % (SNSMOOTHMS / sdr->trk.loopms) == 0) {
            setobsdata(sdr, (*buffloc), (*cnt), &sdr->trk, 1); flagobs = 2;
        } else {
            setobsdata(sdr, (*buffloc), (*cnt), &sdr->trk, 0); flagobs = 1;
        }
        (*loopcnt)++;
    }

    r = &rec[*ntracked];
    r->cnt = (*cnt); r->buffloc = (*buffloc); r->currnsamp = sdr->currnsamp;
    r->flagsync = sdr->nav.flagsync; r->flagsyncf = sdr->nav.flagsyncf;
    r->flagtow = sdr->nav.flagtow; r->flagdec = sdr->nav.flagdec;
    r->flagloopfilter = sdr->trk.flagloopfilter;
    r->swsync = sdr->nav.swsync; r->swloop = sdr->nav.swloop;
    r->synci = sdr->nav.synci; r->biti = sdr->nav.biti; r->bit = sdr->nav.bit;
    for (k = 0; k < 1 + 2 * sdr->trk.corrn && k < OH_MAXTAPS; k++) {
        r->II[k] = sdr->trk.II[k]; r->QQ[k] = sdr->trk.QQ[k];
    }
    r->sumI0 = sdr->trk.sumI[0]; r->sumQ0 = sdr->trk.sumQ[0];
    r->codefreq = sdr->trk.codefreq; r->codeErr = sdr->trk.codeErr; r->codeNco = sdr->trk.codeNco;
    r->carrfreq = sdr->trk.carrfreq; r->carrErr = sdr->trk.carrErr; r->carrNco = sdr->trk.carrNco;
    r->freqErr = sdr->trk.freqErr;
    r->remcode = sdr->trk.remcode; r->remcarr = sdr->trk.remcarr;
    r->L0 = sdr->trk.L[0];
    r->D0 = sdr->trk.D[0]; r->S0 = sdr->trk.S[0]; r->remcout0 = sdr->trk.remcout[0];
    r->Isum = sdr->trk.Isum; r->obs_codei = sdr->trk.codei[0]; r->obs_cnt = sdr->trk.cntout[0];
    r->flagobs = flagobs; r->polarity = sdr->nav.polarity;

    if (g_log_fp && sdr->no == g_log_ch) {
        if (*ntracked == 0) writelog_header(g_log_fp, &sdr->trk);
        writelog(g_log_fp, &sdr->trk, &sdr->nav);
    }
    if (sdr->trk.flagloopfilter) clearcumsumcorr(&sdr->trk);
    (*cnt)++;
    (*buffloc) += sdr->currnsamp;
    sdr->trk.buffloc = (*buffloc);
    (*ntracked)++;
    return 1;
}

int oh_acq_track(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
                 uint64_t avail, int nprn, const int *prn, const int *ctype,
                 int nperiods, oh_acq_t *acq, oh_trkrec_t *rec, int *ntracked)
{
    int i, active;
    uint64_t *buffloc, *cnt, *loopcnt;
    if (avail % FILE_BUFFSIZE || avail > nsamples || nprn > MAXSAT) return -2;
    setup_ini(cfg);
    if (ring_setup(cfg, if_data, nsamples) < 0) return -1;
    ring_feed(cfg, avail / FILE_BUFFSIZE);
    buffloc = (uint64_t *)calloc(sizeof(uint64_t), 3 * (size_t)nprn);
    cnt = buffloc + nprn; loopcnt = cnt + nprn;
    memset(rec, 0, sizeof(oh_trkrec_t) * (size_t)nprn * nperiods);

    for (i = 0; i < nprn; i++) {
        sdrch_t *sdr = &sdrch[i];
        memset(&acq[i], 0, sizeof(acq[i]));
        ntracked[i] = 0;
        if (channel_init(cfg, i + 1, prn[i], ctype[i], sdr) < 0) { ring_release(); free(buffloc); return -3; }
        buffloc[i] = do_acquire(sdr, &acq[i], NULL);
    }

    /* lock-step tracking: one code period per channel per sweep */
    do {
        active = 0;
        for (i = 0; i < nprn; i++) {
            sdrch_t *sdr = &sdrch[i];
            uint64_t need;
            if (!sdr->flagacq || ntracked[i] >= nperiods) continue;
            /* stop when the stream cannot supply another full window */
            if (buffloc[i] + 2 * (uint64_t)sdr->nsamp + 200 > nsamples) continue;
            need = (buffloc[i] + 2 * (uint64_t)sdr->nsamp) / FILE_BUFFSIZE + 2;
            if (need > sdrstat.buffcnt) ring_feed(cfg, need);
            active |= track_period(sdr, &buffloc[i], &cnt[i], &loopcnt[i], rec + (size_t)i * nperiods, &ntracked[i]);
        }
    } while (active);

    for (i = 0; i < nprn; i++) { keep_stocks(&sdrch[i], i); freesdrch(&sdrch[i]); }
    if (g_log_fp) fflush(g_log_fp);
    ring_release();
    free(buffloc);
    return 0;
}

/* The same run with the reference's threading model: after the (sequential) acquisitions
 * the whole stream is in the ring and every acquired channel is tracked by its OWN pthread,
 * as sdrthread() does (src/sdrmain.c:168-175, 231-315), all of them calling sdrtracking /
 * cumsumcorr / pll / dll / setobsdata concurrently.                                       */
typedef struct {
    sdrch_t *sdr; uint64_t buffloc, cnt, loopcnt, nsamples;
    oh_trkrec_t *rec; int *ntracked; int nperiods;
} mt_arg_t;

static void *mt_channel(void *p)
{
    mt_arg_t *a = (mt_arg_t *)p;
    while (*a->ntracked < a->nperiods &&
           a->buffloc + 2 * (uint64_t)a->sdr->nsamp + 200 <= a->nsamples) {
        if (!track_period(a->sdr, &a->buffloc, &a->cnt, &a->loopcnt, a->rec, a->ntracked)) break;
    }
    return NULL;
}

int oh_acq_track_mt(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
                    uint64_t avail, int nprn, const int *prn, const int *ctype,
                    int nperiods, oh_acq_t *acq, oh_trkrec_t *rec, int *ntracked)
{
    int i;
    pthread_t th[MAXSAT];
    mt_arg_t arg[MAXSAT];
    if (avail % FILE_BUFFSIZE || avail > nsamples || nprn > MAXSAT) return -2;
    if ((nsamples + FILE_BUFFSIZE - 1) / FILE_BUFFSIZE + 2 > MEMBUFFLEN) return -4;   /* ring must hold it all */
    setup_ini(cfg);
    if (ring_setup(cfg, if_data, nsamples) < 0) return -1;
    ring_feed(cfg, avail / FILE_BUFFSIZE);
    memset(rec, 0, sizeof(oh_trkrec_t) * (size_t)nprn * nperiods);
    for (i = 0; i < nprn; i++) {
        sdrch_t *sdr = &sdrch[i];
        memset(&acq[i], 0, sizeof(acq[i]));
        ntracked[i] = 0;
        if (channel_init(cfg, i + 1, prn[i], ctype[i], sdr) < 0) { ring_release(); return -3; }
        memset(&arg[i], 0, sizeof(arg[i]));
        arg[i].buffloc = do_acquire(sdr, &acq[i], NULL);
        arg[i].sdr = sdr; arg[i].nsamples = nsamples; arg[i].rec = rec + (size_t)i * nperiods;
        arg[i].ntracked = &ntracked[i]; arg[i].nperiods = nperiods;
    }
    ring_feed(cfg, (nsamples + FILE_BUFFSIZE - 1) / FILE_BUFFSIZE + 1);
    for (i = 0; i < nprn; i++)
        if (sdrch[i].flagacq) pthread_create(&th[i], NULL, mt_channel, &arg[i]);
    for (i = 0; i < nprn; i++)
        if (sdrch[i].flagacq) pthread_join(th[i], NULL);
    for (i = 0; i < nprn; i++) freesdrch(&sdrch[i]);
    ring_release();
    return 0;
}

/* ---- single-function pass-throughs ------------------------------------- */
void oh_correlator(const signed char *data, int dtype, double ti, int n,
                   double freq, double phi0, double crate, double coff,
                   const int *s, int ns, double *II, double *QQ,
                   double *remc, double *remp, const short *code, int coden)
{
    correlator((const char *)data, dtype, ti, n, freq, phi0, crate, coff,
               (int *)s, ns, II, QQ, remc, remp, (short *)code, coden);
}

double oh_mixcarr(const signed char *data, int dtype, double ti, int n,
                  double freq, double phi0, short *II, short *QQ)
{
    return mixcarr((const char *)data, dtype, ti, n, freq, phi0, II, QQ);
}

double oh_rescode(const short *code, int len, double coff, int smax,
                  double ci, int n, short *rcode)
{
    return rescode(code, len, coff, smax, ci, n, rcode);
}

void oh_pcorrelator(const signed char *data, int dtype, double ti, int n,
                    const double *freq, int nfreq, double crate, int m,
                    const float *codex, double *P)
{
    if (!g_handles) { openhandles(); g_handles = 1; }
    pcorrelator((const char *)data, dtype, ti, n, (double *)freq, nfreq, crate, m,
                (cpx_t *)codex, P);
}

void oh_cpxfft(float *cpx, int n, int inverse)
{
    if (!g_handles) { openhandles(); g_handles = 1; }
    if (inverse) cpxifft(NULL, (cpx_t *)cpx, n);
    else         cpxfft(NULL, (cpx_t *)cpx, n);
}

void oh_cpxconv(float *cpxa, const float *cpxb, int m, int n, int flagsum, double *conv)
{
    if (!g_handles) { openhandles(); g_handles = 1; }
    cpxconv(NULL, NULL, (cpx_t *)cpxa, (cpx_t *)cpxb, m, n, flagsum, conv);
}

void oh_cpxpspec(float *cpx, int n, int flagsum, double *pspec)
{
    if (!g_handles) { openhandles(); g_handles = 1; }
    cpxpspec(NULL, (cpx_t *)cpx, n, flagsum, pspec);
}

int oh_gencode(int prn, int ctype, short *code, double *crate)
{
    int len = 0;
    double cr = 0.0;
    short *c = gencode(prn, ctype, &len, &cr);
    if (!c) return -1;
    memcpy(code, c, sizeof(short) * len);
    if (crate) *crate = cr;
    free(c);
    return len;
}

int oh_codefft(int prn, int ctype, double f_sf, float *xcode, int *nsamp, int *nfft)
{
    sdrch_t tmp;
    oh_cfg_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.f_sf = f_sf; cfg.f_cf = 1575.42e6; cfg.dtype = 1;
    cfg.corrn = 1; cfg.corrd = 1; cfg.corrp = 1;
    setup_ini(&cfg);
    memset(&tmp, 0, sizeof(tmp));
    if (initsdrch(1, sys_of(ctype), prn, ctype, 1, FTYPE1, cfg.f_cf, f_sf, 0.0, &tmp) < 0) return -1;
    if (nsamp) *nsamp = tmp.nsamp;
    if (nfft) *nfft = tmp.acq.nfft;
    if (xcode) memcpy(xcode, tmp.xcode, sizeof(float) * 2 * (size_t)tmp.acq.nfft);
    freesdrch(&tmp);
    return 0;
}

/* unit-level hook for the frame synchronisation of one code type: the reference's own initnavstruct()
 * (flen, addflen, prelen, prebits, src/sdrinit.c:489-568), predecodefec() and findpreamble()
 * (src/sdrnav.c:288-411) on a caller-supplied nav.fbits register of flen + addflen entries.
 * Returns findpreamble()'s result, *polarity as it left it; -1 on a wrong length / init failure. */
int oh_ref_framesync(int ctype, int prn, const int *fbits, int n, int *polarity)
{
    sdrnav_t nav;
    int r;
    memset(&nav, 0, sizeof(nav));
    if (initnavstruct(sys_of(ctype), ctype, prn, &nav) < 0) return -1;
    if (n != nav.flen + nav.addflen) return -1;
    memcpy(nav.fbits, fbits, sizeof(int) * (size_t)n);
    predecodefec(&nav);
    r = findpreamble(&nav);
    if (polarity) *polarity = nav.polarity;
    free(nav.bitsync); free(nav.fbits); free(nav.fbitsdec); free(nav.ocode);
    return r;
}
