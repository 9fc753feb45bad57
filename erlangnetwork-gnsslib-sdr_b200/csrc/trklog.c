/* trklog.c - the reference's per-period tracking log from the batched engine's records.
 *
 * writelog_header / writelog, src/sdrout.c:386-436: one CSV line per code period with the counter
 * and time of the newest observation slot, prompt I/Q, the accumulated prompt sums, the 2*corrn+1
 * tap values in the order of their offsets (most early ... prompt ... most late), the loop-filter
 * states, the carrier phase and the status flags.  Every field comes from sdrb200_trk_rec_t;
 * nav.firstsftow and flagdec belong to the host's decodenav (src/sdrnav.c:60-83) and are passed in.
 */
#include <stdio.h>
#include <inttypes.h>
#include "sdrb200_internal.h"

/* order of the tap columns: ind[i] = 2*(n-i)-1 (early taps, farthest first), prompt, ind[n+i+1] = 2*(i+1) */
static void tap_order(int n, int *ind)
{
    int i;
    for (i = 0; i < 2 * n + 1; i++) ind[i] = 0;
    for (i = 0; i < n; i++) { ind[i] = 2 * (n - i) - 1; ind[n + i + 1] = 2 * (i + 1); }
}

int sdrb200_trk_log_header(int corrn, int corrd, char *buf, size_t cap)
{
    int ind[2 * SDRB200_MAXCORRN + 1], i, len = 0, w;
    if (!buf || corrn < 1 || corrn > SDRB200_MAXCORRN) return -1;
    tap_order(corrn, ind);
#define PUT(...) do { w = snprintf(buf + len, len < (int)cap ? cap - (size_t)len : 0, __VA_ARGS__); if (w < 0) return -1; len += w; } while (0)
    PUT("Cnt,Tow,IP,QP,sumI,sumQ");
    for (i = 0; i < 2 * corrn + 1; i++) {
        /* trk.corrx: [2i-1] = -corrd*i, [2i] = +corrd*i (src/sdrinit.c:452-456) */
        const int k = ind[i];
        const int x = (k == 0) ? 0 : ((k & 1) ? -corrd * ((k + 1) / 2) : corrd * (k / 2));
        PUT(",I(%d)", x);
    }
    PUT(",Code Freq,Code Err,Code NCO");
    PUT(",Carr Freq,Carr Err,Carr NCO");
    PUT(",Freq Err,Carrier Phase");
    PUT(",FlagSync,FlagSyncf,FlagTOW,FlagDec,FlagLoopFilter,swsync");
    PUT("\n");
    return (len < (int)cap) ? len : -1;
}

int sdrb200_trk_log_line(const sdrb200_trk_rec_t *r, int corrn, double firstsftow, double ctime, int flagdec,
                         char *buf, size_t cap)
{
    int ind[2 * SDRB200_MAXCORRN + 1], i, len = 0, w;
    /* trk.tow[0] as setobsdata left it (src/sdrtrk.c:170-171); 0 before the first call */
    const double tow = (r->obs_cnt || r->obs_codei) ? firstsftow + (double)r->obs_dcnt * ctime : 0.0;
    if (!r || !buf || corrn < 1 || corrn > SDRB200_MAXCORRN) return -1;
    tap_order(corrn, ind);
    PUT("%" PRIu64 ",%f,%f,%f", (uint64_t)r->obs_cnt, tow, r->II[0], r->QQ[0]);
    PUT(",%f,%f", r->sumI0, r->sumQ0);
    for (i = 0; i < 2 * corrn + 1; i++) PUT(",%f", r->II[ind[i]]);
    PUT(",%f,%f,%f", r->codefreq, r->codeErr, r->codeNco);
    PUT(",%f,%f,%f", r->carrfreq, r->carrErr, r->carrNco);
    PUT(",%f,%f", r->freqErr, r->L0);
    PUT(",%d,%d,%d,%d,%d,%d", r->flagsync, r->flagsyncf, r->flagtow, flagdec, r->flagloopfilter, r->swsync);
    PUT("\n");
#undef PUT
    return (len < (int)cap) ? len : -1;
}
