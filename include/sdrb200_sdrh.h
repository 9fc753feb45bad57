/* sdrb200_sdrh.h - what libsdrb200_sdrh.so exports.
 *
 * libsdrb200_sdrh.so is built from sdrh/sdrh_shim.c AGAINST THE REFERENCE'S OWN
 * src/sdr.h and defines, with the reference's exact prototypes, the hot symbols
 * of that header (no declarations are repeated here - sdr.h is the contract):
 *
 *   acquisition  sdracquisition, checkacquisition            src/sdr.h:610-611
 *                pcorrelator                                  src/sdr.h:682-684
 *                cpxfft, cpxifft, cpxcpx, cpxconv, cpxpspec   src/sdr.h:650-659
 *   tracking     sdrtracking, cumsumcorr, clearcumsumcorr,
 *                pll, dll                                     src/sdr.h:615-619
 *                correlator, mixcarr, rescode                 src/sdr.h:667-668, 680-688
 *                dot_21, dot_22, dot_23                       src/sdr.h:660-666
 *
 * A maintainer of the reference links it instead of sdracq.o / sdrtrk.o and the
 * hot half of sdrcmn.o (INTEGRATION.md shows the Makefile lines).  Besides the
 * sdr.h names it exports five helpers:
 */
#ifndef SDRB200_SDRH_H
#define SDRB200_SDRH_H
#ifdef __cplusplus
extern "C" {
#endif

/* integrations consumed by the calling thread's last sdracquisition()           */
int  sdrb200_sdrh_last_nintg(void);
/* 1 (default): sdracquisition() copies the accumulated power matrix back into
 * the caller's `power` (the reference plots it, src/sdrmain.c:258-261); 0: skip
 * the 8*nfreq*nsamp-byte device-to-host copy                                    */
extern int sdrb200_sdrh_fill_power;
/* 1 (default): sdrtracking() runs the batched engine a chunk of code periods ahead on the samples
 * that have arrived (one mirror IF stream per GPU and front-end type, fed with rcvgetbuff) and
 * serves the per-period calls - and pll()/dll() - from its records; 0: one correlator call per
 * period.  Channel n runs on GPU (n - 1) % number of GPUs.                                  */
extern int sdrb200_sdrh_lookahead;
/* code periods sdrtracking() has served from the batched engine's records so far (all channels) */
unsigned long long sdrb200_sdrh_lookahead_served(void);
/* free the cached device plan of one channel (NULL: all); call before
 * freesdrch() if channels are re-initialised with different parameters          */
void sdrb200_sdrh_release(const void *sdrch);

#ifdef __cplusplus
}
#endif
#endif
