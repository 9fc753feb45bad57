/* harness.h - flat C interface of the parity harness (TEST INFRASTRUCTURE).
 *
 * One interface, three implementations, so that tests can diff plain arrays:
 *   oracle/_ref/libreplay_ref.so   the reference's OWN sources (compiled where
 *                                  they lie under /root/reference) + the FFT shim,
 *                                  driven by oracle/replay.c
 *   oracle/_ref/libreplay_b200.so  the SAME replay.c and the reference's cold
 *                                  objects, but with every hot sdr.h symbol taken
 *                                  from libsdrb200_sdrh.so (the drop-in proof)
 *   oracle/liboracle_port.so       the CPU restatement in oracle/port/ (travels to
 *                                  the GPU box; no reference sources needed)
 *
 * replay.c restates the body of the reference's per-channel thread
 * (src/sdrmain.c:247-315) single-threaded and without sleeps so that a run is
 * deterministic.  Nothing here is product code.
 */
#ifndef ORACLE_HARNESS_H
#define ORACLE_HARNESS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OH_MAXTAPS   33          /* 1 + 2*corrn, corrn <= 16                 */
#define OH_CTYPE_L1CA   1        /* same values as sdr.h:205-212             */
#define OH_CTYPE_G1     20
#define OH_CTYPE_L1SBAS 27

/* front-end + tracking configuration (what the two INI files carry) */
typedef struct {
    double f_sf, f_if, f_cf;     /* [FEND] SF1, IF1, CF1                     */
    int    dtype;                /* 1 real, 2 I/Q                            */
    int    corrn, corrd, corrp;  /* [TRACK] CORRN, CORRD, CORRP              */
    double dllb[2], pllb[2], fllb[2];
    /* acquisition grid; 0 => the sdr.h defaults (src/sdr.h:140-149)         */
    int    acq_intg;
    double acq_hband, acq_step;
} oh_cfg_t;

/* one acquisition result (what sdracquisition leaves in sdrch_t.acq)        */
typedef struct {
    int      prn, ctype;
    int      acquired;           /* sdr->flagacq                              */
    int      freqi, codei;       /* acq.freqi, acq.acqcodei                   */
    int      nintg_used;         /* integrations consumed (early exit)        */
    double   acqfreq, peakr, cn0;
    double   maxp;               /* P[freqi*nsamp+codei]                      */
    uint64_t buffloc;            /* return value of sdracquisition            */
} oh_acq_t;

/* one tracking log record per channel per code period (superset of the
 * reference's writelog CSV row, src/sdrout.c:412-436) */
typedef struct {
    uint64_t cnt, buffloc;
    int      currnsamp;
    int      flagsync, flagsyncf, flagtow, flagdec, flagloopfilter;
    int      swsync, swloop, synci, biti, bit;
    double   II[OH_MAXTAPS], QQ[OH_MAXTAPS];   /* after cumsumcorr polarity   */
    double   sumI0, sumQ0;                     /* before clearcumsumcorr      */
    double   codefreq, codeErr, codeNco;
    double   carrfreq, carrErr, carrNco, freqErr;
    double   remcode, remcarr;
    double   L0;                               /* trk.L[0]                    */
    /* observables left by setobsdata (src/sdrtrk.c:160-209), newest slot [0] */
    double   D0, S0, remcout0, Isum;           /* trk.D[0], S[0], remcout[0], Isum */
    uint64_t obs_codei, obs_cnt;               /* trk.codei[0], cntout[0]     */
    int      flagobs;                          /* 0: not called this period, 1: called, 2: with snrflag */
    int      polarity;                         /* nav.polarity (set by findpreamble)                    */
} oh_trkrec_t;

/* which implementation answered: "reference", "b200-dropin" or "port" */
const char *oh_impl(void);

/* Acquisition of nprn channels on one IF stream.
 * if_data: int8 samples (dtype bytes per sample), nsamples long.
 * The search windows are the ones sdracquisition would use if the ring buffer
 * held exactly `avail` samples (buffloc = avail - (intg+1)*nsamp, src/sdracq.c:24-26);
 * `avail` must be a multiple of 65536 and <= nsamples.
 * power_out (optional, may be NULL): nprn * nfreq * nsamp doubles.
 * returns 0 on success. */
int oh_acquire(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
               uint64_t avail, int nprn, const int *prn, const int *ctype,
               oh_acq_t *out, double *power_out);

/* Acquire (as above) then track every acquired channel for nperiods code
 * periods, in lockstep, following src/sdrmain.c:264-315.
 * rec: nprn * nperiods records (rows of channels that did not acquire are
 * zero).  ntracked[i] = records written for channel i.  returns 0 on success. */
int oh_acq_track(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
                 uint64_t avail, int nprn, const int *prn, const int *ctype,
                 int nperiods, oh_acq_t *acq, oh_trkrec_t *rec, int *ntracked);

/* As oh_acq_track, but with the reference's threading model: one pthread per acquired
 * channel runs the tracking loop (src/sdrmain.c:168-175, 231-315); the whole stream must
 * fit the ring.  Only the builds of replay.c provide it (reference, b200 drop-in).        */
int oh_acq_track_mt(const oh_cfg_t *cfg, const signed char *if_data, uint64_t nsamples,
                    uint64_t avail, int nprn, const int *prn, const int *ctype,
                    int nperiods, oh_acq_t *acq, oh_trkrec_t *rec, int *ntracked);

/* ---- observables; replay builds only -------------------------------------------
 * the 80-deep interpolation stocks setobsdata shifts (src/sdrtrk.c:160-209, src/sdr.h:384-391),
 * newest first, as they stand after the last oh_acq_track run                                  */
#define OH_OBSINTERPN 80
typedef struct {
    double   L[OH_OBSINTERPN], D[OH_OBSINTERPN], S[OH_OBSINTERPN], remcout[OH_OBSINTERPN];
    uint64_t codei[OH_OBSINTERPN], cntout[OH_OBSINTERPN], codeisum[OH_OBSINTERPN];
} oh_stocks_t;
int oh_last_stocks(int channel_index, oh_stocks_t *out);
/* syncthread's use of the stocks (src/sdrsync.c:96-121) with the reference's uint64todouble/interp1 */
int oh_interp_obs(const oh_stocks_t *st, uint64_t sampbase, double samprefd, double *L, double *D);
/* write the reference's own tracking log (writelog, src/sdrout.c:386-436) of channel number
 * `chno` (1-based, sdr->no) during the next oh_acq_track; NULL closes it                      */
int oh_set_logfile(const char *path, int chno);

/* Thin pass-throughs to single hot functions, for unit-level parity. */
void oh_correlator(const signed char *data, int dtype, double ti, int n,
                   double freq, double phi0, double crate, double coff,
                   const int *s, int ns, double *II, double *QQ,
                   double *remc, double *remp, const short *code, int coden);
double oh_mixcarr(const signed char *data, int dtype, double ti, int n,
                  double freq, double phi0, short *II, short *QQ);
double oh_rescode(const short *code, int len, double coff, int smax,
                  double ci, int n, short *rcode);
void oh_pcorrelator(const signed char *data, int dtype, double ti, int n,
                    const double *freq, int nfreq, double crate, int m,
                    const float *codex, double *P);
void oh_cpxfft(float *cpx, int n, int inverse);
void oh_cpxconv(float *cpxa, const float *cpxb, int m, int n, int flagsum, double *conv);
void oh_cpxpspec(float *cpx, int n, int flagsum, double *pspec);
/* code table: returns length, writes +-1 shorts; crate out */
int  oh_gencode(int prn, int ctype, short *code, double *crate);
/* frequency-domain replica exactly as initsdrch builds it (src/sdrinit.c:645-653) */
int  oh_codefft(int prn, int ctype, double f_sf, float *xcode /* 2*nfft */, int *nsamp, int *nfft);

#ifdef __cplusplus
}
#endif
#endif
