/* port_harness.c - oracle/harness.h on top of the CPU restatement
 * (TEST INFRASTRUCTURE).  Same entry points as oracle/replay.c, which drives the
 * reference's own objects; the channel loop below restates the body of the
 * reference's per-channel thread, src/sdrmain.c:247-315, without sleeps. */
#include <stdlib.h>
#include <string.h>
#include "port.h"
#include "../harness.h"

#define BLK 65536                                  /* FILE_BUFFSIZE, src/sdr.h:137 */

const char *oh_impl(void) { return "port"; }

static void to_pcfg(const oh_cfg_t *a, p_cfg_t *b)
{
    memset(b, 0, sizeof(*b));
    b->f_sf = a->f_sf; b->f_if = a->f_if; b->f_cf = a->f_cf; b->dtype = a->dtype;
    b->corrn = a->corrn; b->corrd = a->corrd; b->corrp = a->corrp;
    memcpy(b->dllb, a->dllb, sizeof(b->dllb));
    memcpy(b->pllb, a->pllb, sizeof(b->pllb));
    memcpy(b->fllb, a->fllb, sizeof(b->fllb));
    b->acq_intg = a->acq_intg; b->acq_hband = a->acq_hband; b->acq_step = a->acq_step;
}

static uint64_t acquire_one(p_chan_t *c, const signed char *if_data, uint64_t avail,
                            oh_acq_t *o, double *power_out)
{
    const size_t np = (size_t)c->nsamp * c->nfreq;
    double *power = (double *)calloc(np, sizeof(double));          /* src/sdrmain.c:252 */
    uint64_t buffloc = p_acquisition(c, if_data, avail, power);
    o->prn = c->prn; o->ctype = c->ctype; o->acquired = c->flagacq;
    o->freqi = c->freqi; o->codei = c->acqcodei; o->nintg_used = c->nintg_used;
    o->acqfreq = c->acqfreq; o->peakr = c->peakr; o->cn0 = c->cn0;
    o->maxp = power[(size_t)c->freqi * c->nsamp + c->acqcodei];
    o->buffloc = buffloc;
    if (power_out) memcpy(power_out, power, sizeof(double) * np);
    free(power);
    return buffloc;
}

int oh_acquire(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
               uint64_t avail, int nprn, const int *prn, const int *ctype,
               oh_acq_t *out, double *power_out)
{
    p_cfg_t pc;
    int i;
    if (avail % BLK || avail > nsamples) return -2;
    to_pcfg(cfg, &pc);
    for (i = 0; i < nprn; i++) {
        p_chan_t c;
        memset(&out[i], 0, sizeof(out[i]));
        if (p_chan_init(&c, &pc, prn[i], ctype[i]) < 0) return -3;
        acquire_one(&c, if_data, avail, &out[i],
                    power_out ? power_out + (size_t)i * c.nsamp * c.nfreq : NULL);
        p_chan_free(&c);
    }
    return 0;
}

int oh_acq_track(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
                 uint64_t avail, int nprn, const int *prn, const int *ctype,
                 int nperiods, oh_acq_t *acq, oh_trkrec_t *rec, int *ntracked)
{
    p_cfg_t pc;
    p_chan_t *ch;
    uint64_t *buffloc, *cnt;
    int i, k, active, flagobs = 0;
    if (avail % BLK || avail > nsamples) return -2;
    to_pcfg(cfg, &pc);
    ch = (p_chan_t *)calloc((size_t)nprn, sizeof(p_chan_t));
    buffloc = (uint64_t *)calloc(2 * (size_t)nprn, sizeof(uint64_t));
    if (!ch || !buffloc) { free(ch); free(buffloc); return -1; }
    cnt = buffloc + nprn;
    memset(rec, 0, sizeof(oh_trkrec_t) * (size_t)nprn * nperiods);
    for (i = 0; i < nprn; i++) {
        memset(&acq[i], 0, sizeof(acq[i]));
        ntracked[i] = 0;
        if (p_chan_init(&ch[i], &pc, prn[i], ctype[i]) < 0) { free(ch); free(buffloc); return -3; }
        buffloc[i] = acquire_one(&ch[i], if_data, avail, &acq[i], NULL);
    }
    do {
        active = 0;
        for (i = 0; i < nprn; i++) {
            p_chan_t *c = &ch[i];
            oh_trkrec_t *r;
            if (!c->flagacq || ntracked[i] >= nperiods) continue;
            if (buffloc[i] + 2 * (uint64_t)c->nsamp + 200 > nsamples) continue;
            p_tracking(c, if_data, buffloc[i], cnt[i]);
            active = 1;
            /* src/sdrmain.c:269-280 */
            p_cumsumcorr(c, 1);                              /* ocode[] is all ones */
            c->flagloopfilter = 0;
            flagobs = 0;
            if (!c->flagsync) {
                p_pll(c, &c->prm1, c->ctime); p_dll(c, &c->prm1, c->ctime);
                c->flagloopfilter = 1;
            } else if (c->swloop) {
                p_pll(c, &c->prm2, (double)c->loopms / 1000); p_dll(c, &c->prm2, (double)c->loopms / 1000);
                c->flagloopfilter = 2;
                /* src/sdrmain.c:285-304: SNSMOOTHMS = 100, src/sdr.h:198 */
                flagobs = (c->loopcnt % (100 / c->loopms) == 0) ? 2 : 1;
                p_setobsdata(c, buffloc[i], cnt[i], flagobs == 2);
                c->loopcnt++;
            }
            r = &rec[(size_t)i * nperiods + ntracked[i]];
            r->cnt = cnt[i]; r->buffloc = buffloc[i]; r->currnsamp = c->currnsamp;
            r->flagsync = c->flagsync; r->flagloopfilter = c->flagloopfilter;
            r->flagsyncf = c->flagsyncf; r->flagtow = c->flagtow; r->polarity = c->polarity;
            r->swsync = c->swsync; r->swloop = c->swloop; r->synci = c->synci;
            r->biti = c->biti; r->bit = c->bit;
            for (k = 0; k < 1 + 2 * c->corrn; k++) { r->II[k] = c->II[k]; r->QQ[k] = c->QQ[k]; }
            r->sumI0 = c->sumI[0]; r->sumQ0 = c->sumQ[0];
            r->codefreq = c->codefreq; r->codeErr = c->codeErr; r->codeNco = c->codeNco;
            r->carrfreq = c->carrfreq; r->carrErr = c->carrErr; r->carrNco = c->carrNco;
            r->freqErr = c->freqErr; r->remcode = c->remcode; r->remcarr = c->remcarr;
            r->L0 = c->L0; r->D0 = c->D0; r->S0 = c->S0; r->remcout0 = c->remcout0; r->Isum = c->Isum;
            r->obs_codei = c->obs_codei; r->obs_cnt = c->obs_cnt; r->flagobs = flagobs;
            if (c->flagloopfilter) p_clearcumsumcorr(c);     /* src/sdrmain.c:310 */
            cnt[i]++;
            buffloc[i] += (uint64_t)c->currnsamp;
            ntracked[i]++;
        }
    } while (active);
    for (i = 0; i < nprn; i++) p_chan_free(&ch[i]);
    free(ch); free(buffloc);
    return 0;
}

void oh_correlator(const signed char *data, int dtype, double ti, int n, double freq, double phi0,
                   double crate, double coff, const int *s, int ns, double *II, double *QQ,
                   double *remc, double *remp, const short *code, int coden)
{
    p_correlator(data, dtype, ti, n, freq, phi0, crate, coff, s, ns, II, QQ, remc, remp, code, coden);
}
double oh_mixcarr(const signed char *data, int dtype, double ti, int n, double freq, double phi0,
                  short *II, short *QQ)
{
    return p_mixcarr(data, dtype, ti, n, freq, phi0, II, QQ);
}
double oh_rescode(const short *code, int len, double coff, int smax, double ci, int n, short *rcode)
{
    return p_rescode(code, len, coff, smax, ci, n, rcode);
}
void oh_pcorrelator(const signed char *data, int dtype, double ti, int n, const double *freq,
                    int nfreq, double crate, int m, const float *codex, double *P)
{
    p_pcorrelator(data, dtype, ti, n, freq, nfreq, crate, m, codex, P);
}
void oh_cpxfft(float *cpx, int n, int inverse) { p_fft(cpx, n, inverse); }
void oh_cpxconv(float *a, const float *b, int m, int n, int flagsum, double *conv)
{
    p_cpxconv(a, b, m, n, flagsum, conv);
}
void oh_cpxpspec(float *cpx, int n, int flagsum, double *pspec) { p_cpxpspec(cpx, n, flagsum, pspec); }
int oh_gencode(int prn, int ctype, short *code, double *crate) { return p_gencode(prn, ctype, code, crate); }

int oh_codefft(int prn, int ctype, double f_sf, float *xcode, int *nsamp, int *nfft)
{
    p_cfg_t pc;
    p_chan_t c;
    memset(&pc, 0, sizeof(pc));
    pc.f_sf = f_sf; pc.f_cf = 1575.42e6; pc.dtype = 1; pc.corrn = 1; pc.corrd = 1; pc.corrp = 1;
    if (p_chan_init(&c, &pc, prn, ctype) < 0) return -1;
    if (nsamp) *nsamp = c.nsamp;
    if (nfft) *nfft = c.nfft;
    if (xcode) memcpy(xcode, c.xcode, sizeof(float) * 2 * (size_t)c.nfft);
    p_chan_free(&c);
    return 0;
}

/* unit-level hook: the port's GLONASS string sync on a 200-unit register (pinned against the
 * reference's findpreamble() by tests/test_oracle_port.py through oh_ref_framesync)        */
int oh_port_framesync_g1(const int *reg200, int *polarity) { return p_frame_sync_g1(reg200, polarity); }
