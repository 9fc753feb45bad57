/* sdrb200.h - C ABI of the B200-native GNSS-SDRLIB correlation engine.
 *
 * libsdrb200.so exports exactly the symbols declared here: plain pointers and
 * sizes, no C++ or torch types.  Two layers sit on this header:
 *
 *   1. the batched engine (sdrb200_acq_*, sdrb200_trk_*): one launch sequence
 *      over the whole PRN x Doppler grid / over all channels, IF resident in
 *      HBM, only peak tuples / log records leave the GPU;
 *   2. single-call primitives (sdrb200_correlator, sdrb200_pcorrelator, ...)
 *      with HOST buffers, which libsdrb200_sdrh.so wraps under the reference's
 *      own sdr.h names (see include/sdrb200_sdrh.h) so that sdrmain.c /
 *      sdrinit.c / sdrspec.c link unchanged.
 *
 * Every entry point cites the reference interface it stands in for
 * (file:line under the reference's src/).  There is no CPU fallback: when no
 * CUDA device is usable every compute entry point returns SDRB200_ENODEV.
 */
#ifndef SDRB200_H
#define SDRB200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ status */
#define SDRB200_OK        0
#define SDRB200_EINVAL   -1      /* bad argument                              */
#define SDRB200_ENODEV   -2      /* no usable CUDA device (no CPU fallback)   */
#define SDRB200_ENOMEM   -3
#define SDRB200_ECUDA    -4      /* CUDA runtime error (see last_error)       */
#define SDRB200_ERANGE   -5      /* request outside the data held             */

const char *sdrb200_version(void);
const char *sdrb200_last_error(void);           /* thread-local message        */
int  sdrb200_device_count(void);
int  sdrb200_set_device(int dev);               /* for the calling thread      */
int  sdrb200_sm_count(void);
/* number of kernels this library launched in the calling process so far     */
uint64_t sdrb200_launch_count(void);

/* measurement helpers (bench.py): CUDA events on this library's own streams,
 * and an L2 flush (a write larger than the 126 MB L2) between timed iterations */
int  sdrb200_evt_create(void **evt);
void sdrb200_evt_destroy(void *evt);
int  sdrb200_evt_record(void *evt, void *cuda_stream);
int  sdrb200_evt_elapsed_ms(void *start, void *stop, float *ms);  /* waits for stop */
int  sdrb200_l2_flush(void *cuda_stream);
int  sdrb200_device_sync(void);

/* code types: same numeric values as the reference (src/sdr.h:205-212)      */
#define SDRB200_CTYPE_L1CA    1
#define SDRB200_CTYPE_G1      20
#define SDRB200_CTYPE_L1SBAS  27
#define SDRB200_DTYPE_REAL    1  /* DTYPEI  (src/sdr.h:126)                   */
#define SDRB200_DTYPE_IQ      2  /* DTYPEIQ (src/sdr.h:127)                   */

/* ------------------------------------------------------------------- codes
 * gencode(), src/sdrcode.c:523-539 (L1CA/SBAS 101-154; G1 426-444, which the
 * reference defines but never dispatches).  code: +-1, returns the length.   */
int sdrb200_gencode(int prn, int ctype, int8_t *code, double *crate);
/* the same tables generated on the GPU, one CTA per code (shift registers in two lanes, chips
 * combined by the CTA): codes [n][stride] in HOST memory, stride >= 1023.  Returns the code
 * length (1023 or 511).                                                                    */
int sdrb200_gencode_device(int n, const int *prn, int ctype, int8_t *codes, int stride);

/* --------------------------------------------------------- device IF stream
 * Device-resident counterpart of the reference's ring buffer sdrstat.buff and
 * rcvgetbuff() (src/sdrrcv.c:406-463, file front end 469-532): samples are
 * addressed by ABSOLUTE sample index; `capacity` samples are kept (ring).     */
typedef struct sdrb200_stream sdrb200_stream_t;
int  sdrb200_stream_create(sdrb200_stream_t **s, int dtype, uint64_t capacity);
void sdrb200_stream_destroy(sdrb200_stream_t *s);
/* append n samples of signed 8-bit IF from host memory (H2D copy)            */
int  sdrb200_stream_push_host(sdrb200_stream_t *s, const int8_t *host, uint64_t n);
/* append n samples of RTL-SDR offset-binary u8 I/Q, converted on the device
 * by the rule of rtlsdr_exp(), src/rcv/rtlsdr/rtlsdr.c:136-142               */
int  sdrb200_stream_push_host_rtlsdr(sdrb200_stream_t *s, const uint8_t *host, uint64_t n);
/* append n samples of a packed front-end format, ONE byte per sample, expanded on
 * the device by the reference's tables: stereo_exp() (src/rcv/stereo/stereo.c:163-204)
 * and gn3s_exp_v3() (src/rcv/gn3s/gn3s.cpp:143-177); FMT_RTLSDR forwards to the call
 * above (dtype bytes per sample)                                               */
#define SDRB200_FMT_RTLSDR      1
#define SDRB200_FMT_STEREO_FE1  2   /* MAX2769, 2 bit real      -> dtype 1 */
#define SDRB200_FMT_STEREO_FE2  3   /* MAX2112, 3+3 bit I/Q     -> dtype 2 */
#define SDRB200_FMT_GN3S_2BIT   4   /* GN3S v3 modes 32/38      -> dtype 1 */
#define SDRB200_FMT_GN3S_4BIT   5   /* GN3S v3 modes 36/42, I/Q -> dtype 2 */
int  sdrb200_stream_push_host_packed(sdrb200_stream_t *s, const uint8_t *host, uint64_t n, int format);
/* append n samples that are already in DEVICE memory (produced by another kernel or
 * written by a DMA engine); no counterpart in the reference, whose front ends fill a
 * host ring (src/sdrrcv.c:300-404)                                            */
int  sdrb200_stream_push_device(sdrb200_stream_t *s, const int8_t *d_src, uint64_t n);
/* advance the absolute index by n samples without storing any (content undefined)             */
int  sdrb200_stream_skip(sdrb200_stream_t *s, uint64_t n);
uint64_t sdrb200_stream_count(const sdrb200_stream_t *s);  /* samples pushed   */
/* copy n samples starting at absolute index `buffloc` back to the host
 * (the rcvgetbuff contract)                                                   */
int  sdrb200_stream_read(const sdrb200_stream_t *s, uint64_t buffloc, uint64_t n, int8_t *host);

/* -------------------------------------------------------------- acquisition
 * Batched counterpart of sdracquisition() (src/sdracq.c:14-62), pcorrelator()
 * (src/sdrcmn.c:738-773), cpxconv() (228-251) and checkacquisition()
 * (src/sdracq.c:71-95) over all PRNs x Doppler bins in one pass.              */
typedef struct {
    double f_sf;             /* sampling frequency (Hz)                        */
    int    dtype;            /* SDRB200_DTYPE_*                                */
    int    nsamp;            /* n: samples per code period (sdrch_t.nsamp)     */
    int    nsampchip;        /* samples per chip (sdrch_t.nsampchip)           */
    double ctime;            /* code period (s)                                */
    int    nfreq;            /* Doppler bins                                   */
    const double *freq;      /* [nfreq] absolute wipe-off frequencies (Hz)     */
    int    nprn;             /* replicas in the bank                           */
    int    clen;             /* chips per code                                 */
    double crate;            /* chip rate                                      */
    const int8_t *codes;     /* [nprn][clen] +-1                               */
    int    intg;             /* max non-coherent integrations (acq.intg)       */
    int    ncoh;             /* code periods per coherent block (1 = reference)*/
    double threshold;        /* peak-ratio threshold (ACQTH, src/sdr.h:147)    */
    int    grid_per_prn;     /* 0: the nfreq bins are shared by all replicas; 1: freq is [nprn][nfreq],
                              * replica p is searched over its own bins (GLONASS FDMA channels, each with
                              * its own carrier, src/sdrinit.c:633-635); windows of <= 8 sub-transforms */
} sdrb200_acq_cfg_t;

typedef struct {
    int    acquired;         /* flagacq                                        */
    int    freqi, codei;     /* acq.freqi, acq.acqcodei                        */
    int    nintg;            /* integrations consumed before the early exit    */
    double acqfreq;          /* freq[freqi]                                    */
    double peakr, cn0;       /* acq.peakr, acq.cn0                             */
    double maxp, maxp2, meanp;
} sdrb200_acq_result_t;

typedef struct sdrb200_acq sdrb200_acq_t;
int  sdrb200_acq_create(sdrb200_acq_t **plan, const sdrb200_acq_cfg_t *cfg);
/* Decomposition a plan for windows of nsamp*ncoh lags will use: transform length M = R sub-transforms of L points
 * (M is the smallest of 5*2^(k-3), 3*2^(k-2) and 2^k that holds the 2*nsamp*ncoh samples the reference transforms,
 * src/sdrinit.c:625, and has kernels), fused = 1 for the cluster kernel, 0 for the two-pass path.  Pure host
 * logic, no device needed.                                                                              */
int  sdrb200_acq_plan_shape(int nsamp, int ncoh, int grid_per_prn, int *M, int *L, int *R, int *fused);
void sdrb200_acq_destroy(sdrb200_acq_t *plan);
/* search windows start at absolute sample `buffloc0` of the stream and hop by
 * ncoh*nsamp per integration, as sdracquisition does; results: [nprn]         */
int  sdrb200_acq_run(sdrb200_acq_t *plan, const sdrb200_stream_t *s, uint64_t buffloc0,
                     sdrb200_acq_result_t *results);
/* same, IF taken from host memory: window = (intg+1)*ncoh*nsamp samples; the
 * H2D copy is part of the call (the end-to-end path)                          */
int  sdrb200_acq_run_host(sdrb200_acq_t *plan, const int8_t *host_if, uint64_t nsamples,
                          sdrb200_acq_result_t *results);
/* the same without the wait: H2D copy and the integration steps enqueued on the plan's stream; the
 * caller collects with sdrb200_acq_fetch().  With two plans a caller keeps one search's copy and
 * launches behind the other's - the batched form of the reference's one acquisition thread per
 * channel (src/sdrmain.c:144-149, src/sdracq.c:14-62)                                          */
int  sdrb200_acq_enqueue_host(sdrb200_acq_t *plan, const int8_t *host_if, uint64_t nsamples);
/* launch-only variant for timing: enqueue on the plan's stream, no sync, no
 * D2H; results are fetched with sdrb200_acq_fetch()                           */
int  sdrb200_acq_enqueue(sdrb200_acq_t *plan, const sdrb200_stream_t *s, uint64_t buffloc0);
int  sdrb200_acq_fetch(sdrb200_acq_t *plan, sdrb200_acq_result_t *results);
void *sdrb200_acq_cuda_stream(sdrb200_acq_t *plan);
/* Multi-GPU plumbing (north_star: PRN / Doppler slabs per GPU, NCCL only to gather the
 * per-shard peak tuples; rows of pcorrelator are independent, src/sdrcmn.c:759-768, the
 * per-PRN reduction is checkacquisition, src/sdracq.c:71-95).
 * results_device: the [nprn] tuples of the last enqueue in DEVICE memory (valid once the
 * plan's stream has passed them), so that one collective can gather them without a host
 * bounce; copy_results_device enqueues a copy into caller memory on the plan's stream.  */
const void *sdrb200_acq_results_device(sdrb200_acq_t *plan);
int  sdrb200_acq_copy_results_device(sdrb200_acq_t *plan, void *d_dst);
int  sdrb200_acq_sync(sdrb200_acq_t *plan);
/* Doppler slabs: a plan that holds only some bins of the grid cannot take the early exit of
 * sdracquisition (src/sdracq.c:29-43 tests the peak of the WHOLE grid after every
 * integration).  With history on the plan keeps the tuples of every step [intg][nprn];
 * the caller sets cfg.threshold out of reach and replays the early exit over all slabs.  */
int  sdrb200_acq_set_history(sdrb200_acq_t *plan, int on);
const void *sdrb200_acq_history_device(sdrb200_acq_t *plan);
int  sdrb200_acq_fetch_history(sdrb200_acq_t *plan, sdrb200_acq_result_t *hist);
/* per-kernel CUDA-event timing of the enqueued launches; kernel classes:
 * 0 forward (wipe-off + FFT), 1 row search (spectrum product, inverse FFT,
 * power, argmax), 2 combine (scratch path only, windows > 8 sub-transforms),
 * 3 peak statistics + decision                                                 */
int  sdrb200_acq_set_timing(sdrb200_acq_t *plan, int on);
int  sdrb200_acq_kernel_ms(sdrb200_acq_t *plan, double *ms4, long long *cnt4);
/* The [nprn][nfreq][nsamp] power matrix stays on chip unless it has to be
 * accumulated over several integrations (intg > 1).  keep_power(plan, 1) makes
 * the plan store it in HBM as well, so that it can be read back for plotting
 * (the `power` argument of sdracquisition, src/sdracq.c:14, sdrplot.c).        */
int  sdrb200_acq_keep_power(sdrb200_acq_t *plan, int on);
/* copy the accumulated power matrix of one PRN to the host: [nfreq][nsamp]
 * doubles; EINVAL when the matrix is not kept                                */
int  sdrb200_acq_power(sdrb200_acq_t *plan, int prn_index, double *power);

/* ----------------------------------------------------------------- tracking
 * Batched counterpart of sdrtracking() (src/sdrtrk.c:15-54), correlator()
 * (src/sdrcmn.c:687-722), cumsumcorr/pll/dll (src/sdrtrk.c:64-150) and the
 * bit-sync / nav-bit part of sdrnavigation() (src/sdrnav.c:15-41,198-282),
 * for all channels, with the loop filters closed on the device.              */
#define SDRB200_MAXCORRN 16
#define SDRB200_MAXTAPS  (1 + 2 * SDRB200_MAXCORRN)

typedef struct {             /* sdrtrkprm_t, src/sdr.h:359-368 (coefficients)  */
    double dllw2, dllaw, pllw2, pllaw, fllw;
} sdrb200_loop_t;

typedef struct {
    /* channel constants (sdrch_t, src/sdr.h:482-511) */
    double f_sf, f_if, f_cf, foffset;
    int    dtype;
    int    clen;  double crate, ctime;
    int    nsamp;
    int    prn;              /* selects the bit-sync rule (src/sdrnav.c:203)    */
    int    corrn;            /* trk.corrn                                       */
    int    corrp[SDRB200_MAXCORRN]; /* trk.corrp[] tap offsets (samples)        */
    int    ne, nl;           /* trk.ne, trk.nl                                  */
    int    loopms;           /* trk.loopms                                      */
    int    navrate;          /* nav.rate (code periods per data bit); frame sync on the device for 20 (GPS L1CA
                              * subframes), 10 (GLONASS G1 strings) and 2 (SBAS L1: Viterbi, preambles, CRC-24Q;
                              * src/sdrnav.c:288-411) */
    sdrb200_loop_t prm1, prm2;
    const int8_t *code;      /* [clen] +-1                                      */
    /* state handed over by acquisition (src/sdracq.c:51-56)                    */
    uint64_t buffloc;        /* first sample of the first code period           */
    double   acqfreq;        /* acq.acqfreq                                     */
} sdrb200_trk_chan_t;

typedef struct {             /* one record per channel per code period          */
    uint64_t cnt, buffloc;
    int32_t  currnsamp;
    int32_t  flagsync, flagloopfilter, swsync, swloop, synci, biti, bit;
    double   II[SDRB200_MAXTAPS], QQ[SDRB200_MAXTAPS];
    double   sumI0, sumQ0;
    double   codefreq, codeErr, codeNco;
    double   carrfreq, carrErr, carrNco, freqErr;
    double   remcode, remcarr;
    /* observables, newest slot of setobsdata() (src/sdrtrk.c:160-209; scheduling
     * src/sdrmain.c:285-304): carrier phase L[0] (cycles, cumulative), Doppler D[0]
     * (Hz), S[0], remcout[0] (samples), running |I| sum, codei[0], cntout[0].
     * flagobs: 0 = setobsdata not due this period, 1 = ran, 2 = ran with snrflag   */
    double   L0, D0, S0, remcout0, Isum;
    uint64_t obs_codei, obs_cnt;
    int32_t  flagobs;
    /* frame synchronisation of GPS L1CA channels (findpreamble + paritycheck,
     * src/sdrnav.c:43-58, 324-411): nav.flagsyncf, nav.polarity, nav.flagtow and the
     * reference sample / counter of the first subframe (nav.firstsf, nav.firstsfcnt)   */
    int32_t  flagsyncf, polarity, flagtow;
    uint64_t firstsf, firstsfcnt;
    /* cnt - nav.firstsfcnt at the last setobsdata call: trk.tow[0] = nav.firstsftow + obs_dcnt * ctime
     * (src/sdrtrk.c:170-171); firstsftow is the host's (decodenav)                                  */
    int64_t  obs_dcnt;
} sdrb200_trk_rec_t;

/* the interpolation stocks setobsdata keeps per channel (src/sdrtrk.c:160-209, src/sdr.h:384-391:
 * L, D, S, remcout, codei, cntout, codeisum; OBSINTERPN = 80 deep, newest first) - what
 * syncthread interpolates pseudorange, carrier phase and Doppler from (src/sdrsync.c:96-121).
 * tow[] is not kept on the device: it is nav.firstsftow + (cntout[] - nav.firstsfcnt) * ctime
 * (src/sdrtrk.c:170-171) and firstsftow comes from the host's decodenav.                    */
#define SDRB200_OBSINTERPN 80
typedef struct {
    double   L[SDRB200_OBSINTERPN], D[SDRB200_OBSINTERPN], S[SDRB200_OBSINTERPN], remcout[SDRB200_OBSINTERPN];
    uint64_t codei[SDRB200_OBSINTERPN], cntout[SDRB200_OBSINTERPN], codeisum[SDRB200_OBSINTERPN];
} sdrb200_trk_obs_t;

typedef struct sdrb200_trk sdrb200_trk_t;
int  sdrb200_trk_create(sdrb200_trk_t **eng, int nch, const sdrb200_trk_chan_t *chan);
void sdrb200_trk_destroy(sdrb200_trk_t *eng);
/* advance every channel by up to nperiods code periods (fewer if the stream
 * runs out); rec (host, may be NULL): [nch][nperiods]; nrec[nch] = written.    */
int  sdrb200_trk_run(sdrb200_trk_t *eng, const sdrb200_stream_t *s, int nperiods,
                     sdrb200_trk_rec_t *rec, int *nrec);
/* launch-only variant for timing (records stay on the device)                  */
int  sdrb200_trk_enqueue(sdrb200_trk_t *eng, const sdrb200_stream_t *s, int nperiods);
int  sdrb200_trk_sync(sdrb200_trk_t *eng);
int  sdrb200_trk_fetch_obs(sdrb200_trk_t *eng, int channel, sdrb200_trk_obs_t *out);
/* the reference's tracking log (writelog_header / writelog, src/sdrout.c:386-436) from a record:
 * one CSV line into buf (returns its length, or -1 if it does not fit).  corrn/corrd/corrp as in
 * sdrini (trk.corrx is rebuilt from them, src/sdrinit.c:455-468); firstsftow and flagdec come
 * from the host's decodenav (0 until a subframe is decoded); ctime = code period (s).        */
int  sdrb200_trk_log_header(int corrn, int corrd, char *buf, size_t cap);
int  sdrb200_trk_log_line(const sdrb200_trk_rec_t *rec, int corrn, double firstsftow, double ctime, int flagdec,
                          char *buf, size_t cap);
int  sdrb200_trk_reset(sdrb200_trk_t *eng);     /* back to the hand-over state   */
void *sdrb200_trk_cuda_stream(sdrb200_trk_t *eng);
/* One thread-block cluster tracks one channel: rank 0's control warps replay the NCOs and
 * run the loop filters (src/sdrtrk.c:31-47, src/sdrmain.c:264-312), the worker teams of
 * the cluster correlate whole code periods (correlator(), src/sdrcmn.c:687-722).  The
 * cluster size (1, 2, 4, 8) is chosen at the first launch so that all channels of THIS
 * engine are co-resident; a caller that runs several engines at once can set it
 * (0 = choose again).  ctrl_works: rank 0's worker team takes periods too.       */
int  sdrb200_trk_set_cluster(sdrb200_trk_t *eng, int cluster_size, int ctrl_works);
/* phase profile of channel 0 (clock cycles summed over the last run): go, NCO
 * tables, table conversion, wipe-off + prefix sums, tap sums, epilogue, periods */
int  sdrb200_trk_profile(sdrb200_trk_t *eng, int on, long long *out7);

/* ------------------------------------------- single-call primitives (host I/O)
 * Same argument meaning as the reference functions named; buffers are HOST
 * memory, the work runs on the GPU.  Return SDRB200_* (the sdr.h wrappers in
 * libsdrb200_sdrh.so print and return void like the reference).                */
int sdrb200_correlator(const int8_t *data, int dtype, double ti, int n, double freq,
                       double phi0, double crate, double coff, const int *s, int ns,
                       double *II, double *QQ, double *remc, double *remp,
                       const short *code, int coden);       /* sdrcmn.c:687-722 */
int sdrb200_mixcarr(const int8_t *data, int dtype, double ti, int n, double freq,
                    double phi0, short *II, short *QQ, double *rem); /* 633-669 */
int sdrb200_rescode(const short *code, int len, double coff, int smax, double ci,
                    int n, short *rcode, double *rem);               /* 608-621 */
int sdrb200_dot2x(const short *a1, const short *a2, const short *const *b, int nb,
                  int n, double *d1, double *d2);   /* dot_21/22/23, 287-354    */
int sdrb200_cpxfft(float *cpx, int n, int inverse);                  /* 134-175 */
int sdrb200_cpxcpx(const short *II, const short *QQ, double scale, int n,
                   float *cpx);                                      /* 185-195 */
int sdrb200_cpxconv(float *cpxa, const float *cpxb, int m, int n, int flagsum,
                    double *conv);                                   /* 228-251 */
int sdrb200_cpxpspec(float *cpx, int n, int flagsum, double *pspec); /* 261-276 */
int sdrb200_pcorrelator(const int8_t *data, int dtype, double ti, int n,
                        const double *freq, int nfreq, double crate, int m,
                        const float *codex, double *P);              /* 738-773 */
int sdrb200_checkacquisition(const double *P, int nsamp, int nfreq, int nsampchip,
                             double ctime, double *cn0, double *peakr, int *codei,
                             int *freqi);                      /* sdracq.c:71-95 */
/* Allocation-free variant of sdrb200_checkacquisition: same arguments and results.
 * It has NO counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid allocation).  Its scratch (device copy
 * of P, partial maxima, result record) is one static device array of
 * SDRB200_CHECKACQ_NOALLOC_MAXCELLS doubles (+8 KB; ~16 MB per GPU that loads the
 * library, held until exit); nsamp*nfreq beyond that returns SDRB200_ERANGE.  No
 * cudaMalloc/cudaFree/stream creation happens in the call (legacy default stream).
 * All calls share the array: overlapping calls on one GPU would race, so calls of
 * one process are serialised internally.                                        */
#define SDRB200_CHECKACQ_NOALLOC_MAXCELLS (2 * 1024 * 1024)
int sdrb200_checkacquisition_noalloc(const double *P, int nsamp, int nfreq, int nsampchip,
                                     double ctime, double *cn0, double *peakr, int *codei,
                                     int *freqi);

#ifdef __cplusplus
}
#endif
#endif /* SDRB200_H */
