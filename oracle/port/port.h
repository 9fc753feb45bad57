/* port.h - CPU restatement of the reference's two correlation hot paths.
 *
 * TEST INFRASTRUCTURE, not product.  Only tests/, __graft_entry__.smoke() and
 * the cpu_baseline / --impl reference legs of bench.py may link or call this.
 * It exists so that a parity checker travels to machines where the reference
 * sources are absent; it is itself pinned bit-for-bit against the reference
 * build (oracle/_ref/libreplay_ref.so) by tests/test_oracle_port.py and
 * against the committed vectors under tests/golden/.
 *
 * Every function names the reference lines (under src/) it follows.  FFTs go
 * through oracle/offt.c, the same double-precision DFT the reference build's
 * FFTW shim uses, so port and reference build agree to the last bit.
 */
#ifndef ORACLE_PORT_H
#define ORACLE_PORT_H

#include <stdint.h>

#define P_CDIV    32                 /* CDIV   src/sdrcmn.c:9  */
#define P_CMASK   0x1F               /* CMASK  src/sdrcmn.c:10 */
#define P_CSCALE  (1.0 / 32.0)       /* CSCALE src/sdrcmn.c:11 */
#define P_PI      3.1415926535897932 /* PI     src/sdr.h:103   */
#define P_DPI     (2.0 * P_PI)       /* DPI    src/sdr.h:104   */
#define P_ACQTH   3.0                /* ACQTH  src/sdr.h:147   */
#define P_NAVSYNCTH 50               /* NAVSYNCTH src/sdr.h:157 */
#define P_MAXTAPS 33

/* ---- DSP primitives (port_dsp.c) ---- */
double p_mixcarr(const signed char *data, int dtype, double ti, int n, double freq,
                 double phi0, short *I, short *Q);                  /* sdrcmn.c:633-669 */
double p_rescode(const short *code, int len, double coff, int smax, double ci, int n,
                 short *rcode);                                      /* sdrcmn.c:608-621 */
void   p_dots(const short *a1, const short *a2, const short *const *b, int nb, int n,
              double *d1, double *d2);                               /* sdrcmn.c:287-354 */
void   p_correlator(const signed char *data, int dtype, double ti, int n, double freq,
                    double phi0, double crate, double coff, const int *s, int ns,
                    double *II, double *QQ, double *remc, double *remp,
                    const short *code, int coden);                   /* sdrcmn.c:687-722 */
void   p_cpxcpx(const short *I, const short *Q, double scale, int n, float *cpx); /* :185-195 */
void   p_fft(float *cpx, int n, int inverse);                        /* sdrcmn.c:134-175 */
void   p_cpxconv(float *a, const float *b, int m, int n, int flagsum, double *conv); /* :228-251 */
void   p_cpxpspec(float *cpx, int n, int flagsum, double *pspec);    /* sdrcmn.c:261-276 */
void   p_pcorrelator(const signed char *data, int dtype, double ti, int n, const double *freq,
                     int nfreq, double crate, int m, const float *codex, double *P); /* :738-773 */
double p_maxvd(const double *d, int n, int exs, int exe, int *ind);  /* sdrcmn.c:461-476 */
double p_meanvd(const double *d, int n, int exs, int exe);           /* sdrcmn.c:487-497 */

/* ---- codes (port_code.c) ---- */
void   port_rtlsdr_exp(const unsigned char *buf, int n, signed char *out);            /* rcv/rtlsdr/rtlsdr.c:136-142 */
void   port_stereo_exp(const unsigned char *buf, int n, int dtype, signed char *out); /* rcv/stereo/stereo.c:163-204 */
int    port_gn3s_exp_v3(const unsigned char *buf, int n, int i_mode, signed char *out); /* rcv/gn3s/gn3s.cpp:143-177 */
int    p_gencode(int prn, int ctype, short *code, double *crate);    /* sdrcode.c:101-154,426-444,523-539 */

/* ---- channel model (port_acq.c / port_trk.c) ---- */
typedef struct {
    double pllb, dllb, fllb, dllw2, dllaw, pllw2, pllaw, fllw;       /* sdrtrkprm_t sdr.h:359-368 */
} p_loop_t;

typedef struct {
    /* constants: initsdrch, sdrinit.c:583-657 */
    int    prn, ctype, dtype, clen, nsamp, nsampchip;
    double f_cf, f_sf, f_if, foffset, crate, ctime, ti, ci;
    short  code[1023];
    float *xcode;                                  /* 2*nfft floats              */
    /* acquisition: sdracq_t sdr.h:344-356 */
    int    intg, nfreq, nfft, acqcodei, freqi;
    double *freq, acqfreq, cn0, peakr;
    int    flagacq, nintg_used;
    /* tracking: sdrtrk_t sdr.h:371-412 */
    int    corrn, corrp[16], ne, nl, loopms, currnsamp, flagloopfilter;
    p_loop_t prm1, prm2;
    double codefreq, carrfreq, remcode, remcarr, oldremcode, oldremcarr;
    double codeNco, codeErr, carrNco, carrErr, freqErr;
    double II[P_MAXTAPS], QQ[P_MAXTAPS], oldI[P_MAXTAPS], oldQ[P_MAXTAPS];
    double sumI[P_MAXTAPS], sumQ[P_MAXTAPS], oldsumI[P_MAXTAPS], oldsumQ[P_MAXTAPS];
    /* navigation bit sync: sdrnav_t subset, sdr.h:440-479 */
    int    rate, biti, ocodei, synci, flagsync, swsync, swreset, swloop, navcnt, bit;
    int    bitsync[32];
    double bitIP;
    /* observables: sdrtrk_t L/D/S/remcout/codei/cntout slot [0], Isum, sdr.h:384-400 */
    double L0, D0, S0, remcout0, Isum;
    uint64_t obs_codei, obs_cnt;
    int    flagremcarradd, loopcnt;
    /* frame synchronisation: sdrnav_t fbits (flen + addflen = 302 for L1CA), flagsyncf,
     * flagtow, polarity, firstsf, firstsfcnt (sdr.h:440-479) */
    int    fbits[302];
    int    flagsyncf, flagtow, polarity, flagpolarityadd;
    uint64_t firstsf, firstsfcnt;
} p_chan_t;
#define P_FBITS 302
int p_frame_sync_g1(const int *reg200, int *polarity);    /* port_trk.c */

typedef struct {                                   /* what the two INI files carry */
    double f_sf, f_if, f_cf;
    int    dtype, corrn, corrd, corrp;
    double dllb[2], pllb[2], fllb[2];
    int    acq_intg;
    double acq_hband, acq_step;
} p_cfg_t;

int      p_chan_init(p_chan_t *c, const p_cfg_t *cfg, int prn, int ctype);   /* sdrinit.c:385-480,583-657 */
void     p_chan_free(p_chan_t *c);
int      p_checkacquisition(const double *P, p_chan_t *c);                   /* sdracq.c:71-95 */
uint64_t p_acquisition(p_chan_t *c, const signed char *stream, uint64_t avail, double *power); /* sdracq.c:14-62 */
void     p_tracking(p_chan_t *c, const signed char *stream, uint64_t buffloc, uint64_t cnt);   /* sdrtrk.c:15-54 + sdrnav.c:15-41 */
void     p_cumsumcorr(p_chan_t *c, int polarity);                            /* sdrtrk.c:64-76 */
void     p_clearcumsumcorr(p_chan_t *c);                                     /* sdrtrk.c:77-86 */
void     p_pll(p_chan_t *c, const p_loop_t *prm, double dt);                 /* sdrtrk.c:94-125 */
void     p_dll(p_chan_t *c, const p_loop_t *prm, double dt);                 /* sdrtrk.c:133-150 */
void     p_setobsdata(p_chan_t *c, uint64_t buffloc, uint64_t cnt, int snrflag); /* sdrtrk.c:160-209 */

#endif
