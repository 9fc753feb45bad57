/* port_trk.c - one code period of tracking, loop filters and bit sync,
 * restated for the CPU (TEST INFRASTRUCTURE; see port.h). */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "port.h"

/* checksync, src/sdrnav.c:198-233 (overlay code is all ones for L1CA/SBAS/G1,
 * src/sdrinit.c:520-558) */
static int bit_sync_step(double IP, double IPold, p_chan_t *c)
{
    int i;
    if (c->prn > 5) {                                        /* keyed on PRN only, :203 */
        int corr = 0;
        memmove(&c->bitsync[0], &c->bitsync[1], sizeof(int) * (size_t)(c->rate - 1));
        c->bitsync[c->rate - 1] = (IP < 0 ? -1 : 1);
        for (i = 0; i < c->rate; i++) corr += c->bitsync[i];
        if (abs(corr) == c->rate) { c->synci = c->biti; return 1; }
    } else if (IPold * IP < 0) {
        int best, at = 0;
        c->bitsync[c->biti] += 1;
        best = c->bitsync[0];                                /* maxvi, src/sdrcmn.c:417-431 */
        for (i = 1; i < c->rate; i++) if (best < c->bitsync[i]) { best = c->bitsync[i]; at = i; }
        c->synci = at;
        if (best > P_NAVSYNCTH) {
            c->synci--;
            if (c->synci < 0) c->synci = c->rate - 1;
            return 1;
        }
    }
    return 0;
}

/* paritycheck_l1ca, src/sdrnav_gps.c:141-166: bits[0..1] = D29*, D30* of the previous word,
 * bits[2..31] the word, all +-1; the sum of differences is the reference's own test      */
static int parity_l1ca(const int *b)
{
    int p[6], i, stat = 0;
    p[0] = b[0]*b[2]*b[3]*b[4]*b[6]*b[7]*b[11]*b[12]*b[13]*b[14]*b[15]*b[18]*b[19]*b[21]*b[24];
    p[1] = b[1]*b[3]*b[4]*b[5]*b[7]*b[8]*b[12]*b[13]*b[14]*b[15]*b[16]*b[19]*b[20]*b[22]*b[25];
    p[2] = b[0]*b[2]*b[4]*b[5]*b[6]*b[8]*b[9]*b[13]*b[14]*b[15]*b[16]*b[17]*b[20]*b[21]*b[23];
    p[3] = b[1]*b[3]*b[5]*b[6]*b[7]*b[9]*b[10]*b[14]*b[15]*b[16]*b[17]*b[18]*b[21]*b[22]*b[24];
    p[4] = b[1]*b[2]*b[4]*b[6]*b[7]*b[8]*b[10]*b[11]*b[15]*b[16]*b[17]*b[18]*b[19]*b[22]*b[23]*b[25];
    p[5] = b[0]*b[4]*b[6]*b[7]*b[9]*b[10]*b[11]*b[12]*b[14]*b[16]*b[20]*b[23]*b[24]*b[25];
    for (i = 0; i < 6; i++) stat += p[i] - b[26 + i];
    return stat == 0;
}

/* findpreamble + paritycheck for L1CA, src/sdrnav.c:373-411, 324-345 (predecodefec is a copy
 * for L1CA, :296-299).  fbits: 2 + 300 bits, oldest first.                                */
static int frame_sync_l1ca(p_chan_t *c)
{
    static const int pre[8] = {1, -1, -1, -1, 1, -1, 1, 1};              /* src/sdrinit.c:492 */
    int bits[P_FBITS], corr = 0, i, j, stat = 0;
    for (i = 0; i < 8; i++) corr += c->fbits[2 + i] * pre[i];
    if (abs(corr) != 8) return 0;
    c->polarity = corr > 0 ? 1 : -1;                                      /* set before the parity test */
    for (i = 0; i < P_FBITS; i++) bits[i] = c->polarity * c->fbits[i];
    for (i = 0; i < 10; i++) {
        if (bits[i * 30 + 1] == -1) for (j = 2; j < 26; j++) bits[i * 30 + j] *= -1;
        stat += parity_l1ca(&bits[i * 30]);
    }
    return stat == 10;
}

/* findpreamble for GLONASS G1, src/sdrnav.c:394-400: the time mark pre_g1 (src/sdrinit.c:494-496)
 * at the END of the flen = 200 register, i.e. the last 200 entries of fbits here; paritycheck()
 * accepts every G1 string (:357-359), predecodefec is a copy (:296-299)                     */
int p_frame_sync_g1(const int *reg200, int *polarity)
{
    static const int pre[30] = {-1,-1,-1,-1,-1, 1, 1, 1,-1,-1, 1,-1,-1,-1, 1,-1, 1,-1, 1, 1, 1, 1,-1, 1, 1,-1, 1,-1,-1, 1};
    int corr = 0, i;
    for (i = 0; i < 30; i++) corr += reg200[200 - 30 + i] * pre[i];
    if (abs(corr) != 30) return 0;
    *polarity = corr > 0 ? 1 : -1;
    return 1;
}

/* checkbit, src/sdrnav.c:241-282, including the frame-bit shift register (:272-275) */
static void bit_decide(double IP, p_chan_t *c)
{
    const int diffi = c->biti - c->synci;
    c->swreset = 0; c->swsync = 0;
    if (diffi == 1 || diffi == -c->rate + 1) { c->bitIP = IP; c->swreset = 1; c->navcnt = 1; }
    else c->bitIP += IP;
    c->swloop = (c->navcnt % c->loopms == 0) ? 1 : 0;
    if (diffi == 0) {
        c->bit = (c->bitIP < 0) ? -1 : 1;
        memmove(&c->fbits[0], &c->fbits[1], sizeof(int) * (P_FBITS - 1));
        c->fbits[P_FBITS - 1] = c->bit;
        c->swsync = 1;
    }
    c->navcnt++;
}

/* sdrtracking (src/sdrtrk.c:15-54) followed by the bit-sync part of
 * sdrnavigation (src/sdrnav.c:19-41) */
void p_tracking(p_chan_t *c, const signed char *stream, uint64_t buffloc, uint64_t cnt)
{
    const int nt = 1 + 2 * c->corrn;
    double carr[P_MAXTAPS], quad[P_MAXTAPS];
    c->currnsamp = (int)((c->clen - c->remcode) / (c->codefreq / c->f_sf));       /* :31-32 */
    /* memcpy of 1+2*corrn*8 bytes (:35-36): 2*corrn doubles plus the first byte of the last */
    memcpy(c->oldI, c->II, 1 + 2 * (size_t)c->corrn * sizeof(double));
    memcpy(c->oldQ, c->QQ, 1 + 2 * (size_t)c->corrn * sizeof(double));
    c->oldremcode = c->remcode;
    c->oldremcarr = c->remcarr;
    /* the call passes trk.QQ as the cosine output and trk.II as the sine output (:40-43) */
    p_correlator(stream + buffloc * c->dtype, c->dtype, c->ti, c->currnsamp, c->carrfreq,
                 c->oldremcarr, c->codefreq, c->oldremcode, c->corrp, c->corrn, carr, quad,
                 &c->remcode, &c->remcarr, c->code, c->clen);
    memcpy(c->QQ, carr, sizeof(double) * nt);
    memcpy(c->II, quad, sizeof(double) * nt);

    c->biti = (int)(cnt % (uint64_t)c->rate);
    c->ocodei = c->biti - c->synci - 1;
    if (c->ocodei < 0) c->ocodei += c->rate;
    if (c->rate == 1 && cnt > 2000 / (c->ctime * 1000)) { c->synci = 0; c->flagsync = 1; }
    if (!c->flagsync && cnt > 2000 / (c->ctime * 1000))
        c->flagsync = bit_sync_step(c->II[0], c->oldI[0], c);
    if (c->flagsync) {
        bit_decide(c->II[0], c);
        /* frame synchronisation, src/sdrnav.c:43-58 (GPS L1CA, GLONASS G1; what follows it - decodenav and
         * its reset when no TOW decodes, :60-83 - is outside this path)                     */
        if (c->swsync && ((c->ctype == 1 && c->rate == 20) || (c->ctype == 20 && c->rate == 10))) {
            if (!c->flagtow) c->flagsyncf = (c->ctype == 1) ? frame_sync_l1ca(c) : p_frame_sync_g1(&c->fbits[P_FBITS - 200], &c->polarity);
            if (c->flagsyncf && !c->flagtow) { c->firstsf = buffloc; c->firstsfcnt = cnt; c->flagtow = 1; }
        }
    }
}

/* src/sdrtrk.c:64-76 */
void p_cumsumcorr(p_chan_t *c, int polarity)
{
    int i;
    for (i = 0; i < 1 + 2 * c->corrn; i++) {
        c->II[i] *= polarity; c->QQ[i] *= polarity;
        c->oldsumI[i] += c->oldI[i]; c->oldsumQ[i] += c->oldQ[i];
        c->sumI[i] += c->II[i];      c->sumQ[i] += c->QQ[i];
    }
}

/* src/sdrtrk.c:77-86 */
void p_clearcumsumcorr(p_chan_t *c)
{
    int i;
    for (i = 0; i < 1 + 2 * c->corrn; i++) c->oldsumI[i] = c->oldsumQ[i] = c->sumI[i] = c->sumQ[i] = 0;
}

/* src/sdrtrk.c:94-125 */
void p_pll(p_chan_t *c, const p_loop_t *prm, double dt)
{
    const double IP = c->sumI[0], QP = c->sumQ[0], oIP = c->oldsumI[0], oQP = c->oldsumQ[0];
    double carrErr, freqErr, f1, f2;
    carrErr = (IP > 0) ? atan2(QP, IP) / P_PI : atan2(-QP, -IP) / P_PI;
    f1 = (IP == 0) ? P_PI / 2 : atan(QP / IP);
    f2 = (oIP == 0) ? P_PI / 2 : atan(oQP / oIP);
    freqErr = f1 - f2;
    if (freqErr > P_PI / 2) freqErr = P_PI - freqErr;
    if (freqErr < -P_PI / 2) freqErr = -P_PI - freqErr;
    c->carrNco += prm->pllaw * (carrErr - c->carrErr) + prm->pllw2 * dt * carrErr + prm->fllw * dt * freqErr;
    c->carrfreq = c->acqfreq + c->carrNco;
    c->carrErr = carrErr;
    c->freqErr = freqErr;
}

/* src/sdrtrk.c:133-150 */
void p_dll(p_chan_t *c, const p_loop_t *prm, double dt)
{
    const double IE = c->sumI[c->ne], IL = c->sumI[c->nl], QE = c->sumQ[c->ne], QL = c->sumQ[c->nl];
    const double codeErr = (sqrt(IE * IE + QE * QE) - sqrt(IL * IL + QL * QL)) /
                           (sqrt(IE * IE + QE * QE) + sqrt(IL * IL + QL * QL));
    c->codeNco += prm->dllaw * (codeErr - c->codeErr) + prm->dllw2 * dt * codeErr;
    c->codefreq = c->crate - c->codeNco + (c->carrfreq - c->f_if - c->foffset) / (c->f_cf / c->crate);
    c->codeErr = codeErr;
}

/* setobsdata(), src/sdrtrk.c:160-209 - only slot [0] of the interpolation stocks is
 * kept (the shiftdata() calls at :163-168 and :199-201 move history that the hot
 * path never reads back).  The +0.5 cycle polarity term (:186-195) waits for
 * nav.flagsyncf, which is set by frame synchronisation (src/sdrnav.c:46) - outside
 * this path; the replay's random data bits never raise it.                      */
void p_setobsdata(p_chan_t *c, uint64_t buffloc, uint64_t cnt, int snrflag)
{
    const double DPI = 2.0 * 3.1415926535897932;            /* src/sdr.h:103-104 */
    c->obs_codei = buffloc;                                  /* :172 */
    c->obs_cnt = cnt;                                        /* :173 */
    c->remcout0 = c->oldremcode * c->f_sf / c->codefreq;     /* :174 */
    c->D0 = -(c->carrfreq - c->f_if - c->foffset);           /* :177 */
    if (!c->flagremcarradd) {                                /* :180-184 */
        c->L0 -= c->remcarr / DPI;
        c->flagremcarradd = 1;
    }
    if (c->flagsyncf && !c->flagpolarityadd) {               /* :186-195 */
        if (c->polarity == 1) c->L0 += 0.5;
        c->flagpolarityadd = 1;
    }
    c->L0 += c->D0 * (c->loopms * c->currnsamp / c->f_sf);   /* :197 */
    c->Isum += fabs(c->sumI[0]);                             /* :199 */
    if (snrflag) {
        c->S0 = 10 * log(c->Isum / 100.0 / 100.0) + log(500.0) + 5;   /* :206 */
        c->Isum = 0;
    }
}
