/* trk_run_kernel.cuh - k_trk_cl: the batched tracking kernel, one thread-block CLUSTER per
 * channel (included by trk.cu after the epilogue functions).
 *
 * Every CTA has the same warp layout:
 *   warp 0   carrier NCO lane   (control; active in cluster rank 0 only): the serial chain only
 *   warp 1   code NCO lane      (control; rank 0 only)
 *   warp 2   epilogue warp      (control; rank 0 only): cumsumcorr, bit/frame sync, pll/dll,
 *            observables, record - everything around correlator() in sdrtracking() and the
 *            sdrthread loop body (src/sdrtrk.c:31-47, src/sdrmain.c:264-312)
 *   warp 3/4 carrier / code table pushers (control; rank 0 only): turn the lanes' raw run
 *            records into integer segments, one lane per record, and write them into the
 *            worker team's slot
 *   warps 5.. worker team: correlator() of one whole code period (src/sdrcmn.c:687-722)
 * Code period q of the channel is correlated by the worker team of cluster rank
 * r0 + q % W (W teams; r0 = 1 when rank 0 keeps its team idle so that its NCO lanes have the
 * SM to themselves, else 0).  Each team owns two table slots, so period q may be handed out
 * once period q - 2W has reported back.
 *
 * Why this shape: the NCO phases are a serial floating-point recurrence over periods
 * (remcode/remcarr feed the next period), the correlation of a period is not.  After bit
 * sync the loop filters touch carrfreq/codefreq only every loopms-th period
 * (src/sdrnav.c:241-262, src/sdrmain.c:271-280), so in between the NCO lanes run ahead of
 * the correlators and consecutive periods are correlated concurrently on different SMs.
 */
#ifndef TRK_PROF_THREAD
#define TRK_PROF_THREAD 416                  /* worker whose cycle stamps sdrb200_trk_profile reports */
#endif
#define TRK_T0      160                      /* first worker thread                         */
#define TRK_RING    16                       /* periods in flight: result / hand-over ring   */

/* header of a table slot in a worker CTA, written remotely by the control warps */
struct trk_slot_t {
    int32_t ncarr;                           /* carrier segments (carrier lane)             */
    int32_t ncode;                           /* code segments                               */
    int32_t n;                               /* currnsamp                                   */
    int32_t stop;                            /* no more periods: the team leaves            */
    unsigned long long boff;                 /* ring offset of the period's first sample    */
    int32_t q, pad;
};

/* what the NCO lanes hand to the epilogue warp for period q (slot q % TRK_RING) */
struct trk_hand_t {
    int32_t n, go, quiet, pad;
    double  remc, remp;
};

struct trk_cl_smem_t {
    trk_smem_t t;                            /* worker side: tables (2 slots x {carrier, code}), accumulators */
    int32_t taps[SDRB200_MAXCORRN];
    /* ---- worker side */
    unsigned long long fullh[2];             /* mbarriers per slot: header written (1 arrival),      */
    unsigned long long fullc[2];             /*   carrier table written (32), code table written (32) */
    unsigned long long fullk[2];
    trk_slot_t slot[2];
    /* ---- control side (rank 0) */
    unsigned long long done[TRK_RING];       /* mbarriers: tap sums of period q arrived (32 arrivals) */
    unsigned int res[TRK_RING][2 * SDRB200_MAXTAPS + 2];   /* tap sums mod 2^32 (I.., Q..), err */
    trk_hand_t hand[TRK_RING];
    volatile int seq_code;                   /* periods whose n/go the code lane has published   */
    volatile int seq_remc;                   /* ... whose remc is in hand[]                     */
    volatile int seq_remp;                   /* ... whose remp is in hand[] (carrier lane)      */
    volatile int epi_count;                  /* periods the epilogue has finished               */
    volatile int loops_count;                /* periods whose loop filters have run: lf is theirs */
    volatile int hdr_seq;                    /* periods whose n / boff are in rhdr[1]            */
    trk_lf_t lf;                             /* what the lanes need after the loop filters of a period */
    int32_t pad0;
    nco_raw_t raw[2][2][TRK_MAXSEG];         /* [lane][period parity]: run records of the walks  */
    struct { int32_t ns, n, q, pad; unsigned long long boff; } rhdr[2][2];
    volatile int raw_seq[2];                 /* periods whose raw table the lane has finished    */
    volatile int pushed[2];                  /* periods the pusher has written to their team     */
    nco_steptab_t tab[2];                    /* per-binade increments of the current carrier / code step */
    trk_state_t st;
    unsigned int vdec[2 * 756];              /* decision words of the SBAS Viterbi decoder (trk_frame_sync_sbas) */
    sdrb200_trk_rec_t rec;
    trk_chan_dev_t chc;                      /* the channel's constants (code excluded: t.codetab) */
};

/* will the epilogue of the period that starts with these counters leave carrfreq/codefreq alone?
 * Once bit sync is reached (flagsync never drops) the loop filters run only when checkbit()
 * raises swloop, a function of counters (src/sdrnav.c:241-262, src/sdrmain.c:271-280).
 * Advances navcnt the way checkbit does.                                                    */
__device__ __forceinline__ int trk_quiet_forecast(unsigned long long cnt, int synci, int *navcnt, int rate, int loopms)
{
    const int biti = (cnt >> 32) ? (int)(cnt % (unsigned long long)rate) : (int)((uint32_t)cnt % (uint32_t)rate);
    const int diffi = biti - synci;
    const int nc = (diffi == 1 || diffi == -rate + 1) ? 1 : *navcnt;
    *navcnt = nc + 1;
    return (nc % loopms != 0) ? 1 : 0;
}

/* per-binade table of `step` (nco_exact.h, nco_walk_t), built by one warp: lane l fills the
 * binades l+1 and l+33 (1..53); done only when the loop filters have changed the frequency          */
__device__ __noinline__ void trk_steptab_warp(nco_steptab_t *t, double step, int lane)
{
    const uint64_t b = nco_bits(step);
    const int ed = (int)((b >> 52) & 0x7FF);
    const bool ok = step != 0.0 && ed > 0 && ed < 0x7FF;      /* entries depend on |step| only */
    unsigned long long ties = 0ull;
    for (int s = lane + 1; s <= 53; s += 32) {
        double d = 0.0, inv = 0.0; int tie = 0; uint64_t Dn = 0;
        if (ok && !nco_steptab_entry(step, s, &d, &inv, &tie, &Dn)) { d = 0.0; inv = 0.0; }
        t->di[s][0] = d; t->di[s][1] = inv; t->Dn[s] = Dn;
        if (tie) ties |= 1ull << s;
    }
    for (int o = 16; o > 0; o >>= 1) ties |= __shfl_xor_sync(0xffffffffu, ties, o);
    if (lane == 0) {
        t->ties = ties; t->ed = ok ? ed : 0; t->pad = 0;
        t->di[0][0] = 0.0; t->di[0][1] = 0.0; t->Dn[0] = 0;
    }
    __syncwarp();
}
__device__ __noinline__ int trk_walk_raw(double x, double step, int count, double wrap, const nco_steptab_t *tab,
                                         nco_raw_t *raw, double *x_end)
{
    return nco_walk_raw(x, step, count, wrap, tab, raw, TRK_MAXSEG, x_end);
}

/* pusher warp: records of period q -> segments in the team's slot; returns 1 on a range error */
__device__ __forceinline__ int trk_push_raw(const nco_raw_t *raw, int ns, const nco_steptab_t *tab, uint32_t dst_cluster_addr, int lane)
{
    int bad = 0;
    for (int i = lane; i < ns; i += 32) {
        nco_iseg_t o;
        bad |= nco_raw_to_iseg(&raw[i], tab, &o);
        const uint32_t a = dst_cluster_addr + 32u * (uint32_t)i;
        cl_st_v4(a, make_uint4((uint32_t)o.k0, (uint32_t)o.shift, (uint32_t)o.X0, (uint32_t)(o.X0 >> 32)));
        cl_st_v4(a + 16u, make_uint4((uint32_t)(uint64_t)o.D, (uint32_t)((uint64_t)o.D >> 32), (uint32_t)o.neg, 0u));
    }
    return bad;
}

template <int DTYPE, int NT>
__global__ void __launch_bounds__(NT, (NT >= 1024) ? 1 : 2)
k_trk_cl(const trk_chan_dev_t *__restrict__ chans, trk_state_t *__restrict__ states,
         const int8_t *__restrict__ ifbuf, uint64_t capacity, uint64_t avail, int nperiods, int maxn, int qcap,
         sdrb200_trk_rec_t *__restrict__ recs, int *__restrict__ nrec, int *__restrict__ err, int ctrl_works,
         long long *__restrict__ prof, trk_obsstock_t *__restrict__ obsst)
{
    constexpr int NW = NT - TRK_T0, R = TRK_RING;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    trk_cl_smem_t *sm = (trk_cl_smem_t *)smem_raw;
    const trkd_work_t wk = trkd_work_layout((unsigned int *)(smem_raw + ((sizeof(trk_cl_smem_t) + 15) & ~(size_t)15)), NW / 32);
    const int rank = (int)cl_rank(), CL = (int)cl_size();
    const int chno = blockIdx.x / CL;
    const int r0 = (ctrl_works || CL == 1) ? 0 : 1;     /* first rank whose worker team takes periods */
    const int W = CL - r0;
    const trk_chan_dev_t *chg = &chans[chno];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool ctrl = (rank == 0);
    trk_state_t *st = &sm->st;
    /* optional cycle profile of channel 0 (sdrb200_trk_profile): slot i accumulates the cycles since the last stamp */
    const bool pf = (prof != nullptr) && chno == 0 && lane == 0;
    long long pt0 = 0;
#define TRK_PF(i) do { if (pf) { const long long t1_ = trk_clock(); atomicAdd((unsigned long long *)&prof[i], (unsigned long long)(t1_ - pt0)); pt0 = t1_; } } while (0)
#define TRK_PF0() do { if (pf) pt0 = trk_clock(); } while (0)

    /* ---- set-up: every CTA holds the code and taps; rank 0 also the channel state */
    for (int i = threadIdx.x; i < chg->clen; i += NT) sm->t.codetab[i] = chg->code[i];
    if (threadIdx.x < chg->corrn) sm->taps[threadIdx.x] = chg->corrp[threadIdx.x];
    trk_init_smem(&sm->t);
    __syncthreads();
    trkd_init<DTYPE>(wk, &sm->t, chg->clen, sm->taps, chg->corrn, NT);
    if (ctrl) {
        const uint32_t *src = (const uint32_t *)&states[chno];
        uint32_t *dst = (uint32_t *)st;
        for (int i = threadIdx.x; i < (int)(sizeof(trk_state_t) / 4); i += NT) dst[i] = src[i];
        /* record slots beyond the channel's taps are never written: keep them zero */
        for (int i = threadIdx.x; i < (int)(sizeof(sdrb200_trk_rec_t) / 4); i += NT) ((uint32_t *)&sm->rec)[i] = 0u;
        for (int i = threadIdx.x; i < (int)(offsetof(trk_chan_dev_t, code) / 4); i += NT)
            ((uint32_t *)&sm->chc)[i] = ((const uint32_t *)chg)[i];
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; i++) { cl_mbar_init(&sm->fullh[i], 1); cl_mbar_init(&sm->fullc[i], 32); cl_mbar_init(&sm->fullk[i], 32); }
        for (int i = 0; i < R; i++) cl_mbar_init(&sm->done[i], 32);
        sm->seq_code = 0; sm->seq_remc = 0; sm->seq_remp = 0; sm->epi_count = 0;
        sm->raw_seq[0] = 0; sm->raw_seq[1] = 0; sm->pushed[0] = 0; sm->pushed[1] = 0; sm->loops_count = 0; sm->hdr_seq = 0;
        cl_fence_mbar_init();
    }
    __syncthreads();
    cl_sync_all();
    const trk_chan_dev_t *ch = &sm->chc;                 /* valid in rank 0 only */
    const int ns = chg->corrn;
    const int smax = sm->taps[ns - 1];

    if (ctrl && wid == 1) {
        /* ================= code NCO lane: decides n / go of every period, walks the code table */
        double codefreq = 0.0, remc = 0.0;
        unsigned long long buffloc = 0ull, boff = 0ull, cnt = 0ull;
        int navcnt = 0, synci = 0, flagsync = 0, prev_quiet = 0, q = 0;
        double tab_step = -1.0;                               /* step the table sm->tab[1] was built for */
        for (;; q++) {
            int n = 0, go = 0, quiet = 0, nseg = 0;
            TRK_PF0();
            if (lane == 0) {
                if (q < nperiods) {
                    if (q == 0) {
                        codefreq = st->codefreq; remc = st->remcode; cnt = st->cnt; navcnt = st->navcnt;
                        synci = st->synci; flagsync = st->flagsync;
                        buffloc = st->buffloc; boff = buffloc % capacity;
                    } else if (!prev_quiet) {
                        cl_flag_wait_ge(&sm->loops_count, q);        /* the loop filters of period q-1 have run */
                        codefreq = sm->lf.codefreq; cnt = sm->lf.cnt; navcnt = sm->lf.navcnt;
                        synci = sm->lf.synci; flagsync = sm->lf.flagsync;
                    }
                    n = nco_currnsamp(ch->clen, remc, codefreq, ch->f_sf);            /* src/sdrtrk.c:31-32 */
                    go = (n > 0 && n <= maxn && buffloc + (unsigned long long)n <= avail) ? 1 : 0;
                    /* a period that cannot run for another reason than "the stream ends here" is an error:
                     * 2 = the ring no longer holds the window (the channel lags by more than the capacity),
                     * 3 = currnsamp outside the CTA's work area (codefreq ran away)                        */
                    if (go && avail - buffloc > capacity) { go = 0; atomicMax(err, 2); }
                    if (!go && (n <= 0 || n > maxn)) { atomicMax(err, 3); }
                    if (go && flagsync && ch->navrate > 1) quiet = trk_quiet_forecast(cnt, synci, &navcnt, ch->navrate, ch->loopms);
                }
                /* flow control: the team's slot is free, the rings' slots are consumed */
                if (q >= 2 * W) cl_mbar_wait(&sm->done[(q - 2 * W) % R], (uint32_t)(((q - 2 * W) / R) & 1));
                if (q >= R) cl_flag_wait_ge(&sm->epi_count, q - R + 1);
                sm->hand[q % R].n = n; sm->hand[q % R].go = go; sm->hand[q % R].quiet = quiet;
                cl_flag_set(&sm->seq_code, q + 1);
                if (q >= 2) cl_flag_wait_ge(&sm->pushed[1], q - 1);                     /* rhdr / raw buffer q & 1 is free */
                sm->rhdr[1][q & 1].n = go ? n : -1; sm->rhdr[1][q & 1].q = q; sm->rhdr[1][q & 1].boff = boff;
                cl_flag_set(&sm->hdr_seq, q + 1);                                       /* the pusher sends the header (or leaves) */
                if (pf && quiet) atomicAdd((unsigned long long *)&prof[10], 1ull);
            }
            TRK_PF(1);                                       /* code lane: waits + period decisions */
            go = __shfl_sync(0xffffffffu, go, 0);
            if (!go) break;
            n = __shfl_sync(0xffffffffu, n, 0);
            double remc2 = 0.0;
            {
                const double ci_all = __shfl_sync(0xffffffffu, __dmul_rn(ch->ti, codefreq), 0);
                if (ci_all != tab_step) {                     /* the pusher has finished with the old table: period q-1 is done */
                    trk_steptab_warp(&sm->tab[1], ci_all, lane); tab_step = ci_all;
                }
            }
            if (lane == 0) {
                const double ci = __dmul_rn(ch->ti, codefreq);                          /* ti*crate, sdrcmn.c:709 */
                const double c0 = nco_code_init(remc, smax, ci, ch->clen);
                double xe;
                nseg = trk_walk_raw(c0, ci, n + 2 * smax, (double)ch->clen, &sm->tab[1], sm->raw[1][q & 1], &xe);
                if (nseg < 0) { nseg = 0; atomicMax(err, 1); }
                remc2 = nco_code_rem(xe, smax, ci);
                sm->hand[q % R].remc = remc2;
                cl_flag_set(&sm->seq_remc, q + 1);
                sm->rhdr[1][q & 1].ns = nseg;
                cl_flag_set(&sm->raw_seq[1], q + 1);
                remc = remc2; buffloc += (unsigned long long)n; cnt++;
                boff += (unsigned long long)n; if (boff >= capacity) boff -= capacity; if (boff >= capacity) boff %= capacity;
                prev_quiet = quiet;
            }
            __syncwarp();
            TRK_PF(0);                                       /* code lane: table + walk */
        }
        /* no more periods (rhdr[1][q & 1].n = -1 told the pusher to leave) */
        if (lane == 0) cl_flag_wait_ge(&sm->pushed[1], q);   /* every real period is in its slot */
        __syncwarp();
        /* the teams leave: W more hand-outs that carry only the stop flag (this warp arrives for both lanes) */
        for (int j = 0; j < W; j++) {
            const int qq = q + j;
            if (lane == 0 && j > 0 && qq >= 2 * W) cl_mbar_wait(&sm->done[(qq - 2 * W) % R], (uint32_t)(((qq - 2 * W) / R) & 1));
            __syncwarp();
            const uint32_t tr = (uint32_t)(r0 + qq % W);
            const int slot = (qq / W) & 1;
            if (lane == 0) {
                cl_st_u32(cl_map(&sm->slot[slot], tr) + offsetof(trk_slot_t, stop), 1u);
                cl_mbar_arrive_remote(cl_map(&sm->fullh[slot], tr));
            }
            __syncwarp();
        }
    } else if (ctrl && wid == 0) {
        /* ================= carrier NCO lane */
        double carrfreq = 0.0, remp = 0.0, tab_step = -1.0;
        int prev_quiet = 0, qlast = 0;
        for (int q = 0;; q++) {
            int n = 0, go = 0, quiet = 0, nseg = 0;
            double xe = 0.0;
            TRK_PF0();
            if (lane == 0) {
                cl_flag_wait_ge(&sm->seq_code, q + 1);
                n = sm->hand[q % R].n; go = sm->hand[q % R].go; quiet = sm->hand[q % R].quiet;
            }
            go = __shfl_sync(0xffffffffu, go, 0);
            if (!go) break;
            if (lane == 0) {
                if (q == 0) { carrfreq = st->carrfreq; remp = st->remcarr; }
                else if (!prev_quiet) { cl_flag_wait_ge(&sm->loops_count, q); carrfreq = sm->lf.carrfreq; }
            }
            TRK_PF(4);                                       /* carrier lane: waits */
            {
                double phi, ps;
                nco_carrier_init(carrfreq, remp, ch->ti, &phi, &ps);
                ps = __shfl_sync(0xffffffffu, ps, 0);
                if (ps != tab_step) { trk_steptab_warp(&sm->tab[0], ps, lane); tab_step = ps; }
            }
            if (lane == 0) {
                double phi, ps;
                nco_carrier_init(carrfreq, remp, ch->ti, &phi, &ps);
                if (q >= 2) cl_flag_wait_ge(&sm->pushed[0], q - 1);                     /* raw buffer q & 1 is free */
                nseg = trk_walk_raw(phi, ps, n, 0.0, &sm->tab[0], sm->raw[0][q & 1], &xe);
                if (nseg < 0) { nseg = 0; atomicMax(err, 1); }
                sm->rhdr[0][q & 1].ns = nseg; sm->rhdr[0][q & 1].q = q;
                cl_flag_set(&sm->raw_seq[0], q + 1);
            }
            TRK_PF(2);                                       /* carrier lane: table + walk */
            if (lane == 0) {
                remp = trk_rem_call(xe);                       /* mixcarr's subtract loop, src/sdrcmn.c:665-668 */
                sm->hand[q % R].remp = remp;
                cl_flag_set(&sm->seq_remp, q + 1);
                prev_quiet = quiet;
            }
            __syncwarp();
            TRK_PF(3);                                       /* carrier lane: 2*pi reduction */
            qlast = q + 1;
        }
        if (lane == 0) {                                     /* the pusher leaves after the last real period */
            if (qlast >= 2) cl_flag_wait_ge(&sm->pushed[0], qlast - 1);
            sm->rhdr[0][qlast & 1].ns = -1;
            cl_flag_set(&sm->raw_seq[0], qlast + 1);
        }
        __syncwarp();
    } else if (ctrl && (wid == 3 || wid == 4)) {
        /* ================= table pushers: warp 3 carrier, warp 4 code */
        const int k = wid - 3;
        for (int q = 0;; q++) {
            TRK_PF0();
            const uint32_t tr = (uint32_t)(r0 + q % W);
            const int slot = (q / W) & 1;
            if (k == 1) {                                    /* the header goes out as soon as n is known */
                cl_flag_wait_ge(&sm->hdr_seq, q + 1);
                if (sm->rhdr[1][q & 1].n < 0) break;
                if (lane == 0) {
                    const uint32_t h = cl_map(&sm->slot[slot], tr);
                    const unsigned long long bo = sm->rhdr[1][q & 1].boff;
                    cl_st_u32(h + offsetof(trk_slot_t, n), (uint32_t)sm->rhdr[1][q & 1].n);
                    cl_st_u32(h + offsetof(trk_slot_t, stop), 0u);
                    cl_st_u32(h + offsetof(trk_slot_t, boff), (uint32_t)bo);
                    cl_st_u32(h + offsetof(trk_slot_t, boff) + 4, (uint32_t)(bo >> 32));
                    cl_st_u32(h + offsetof(trk_slot_t, q), (uint32_t)q);
                    cl_mbar_arrive_remote(cl_map(&sm->fullh[slot], tr));
                }
                __syncwarp();
            }
            cl_flag_wait_ge(&sm->raw_seq[k], q + 1);
            const int nsg = sm->rhdr[k][q & 1].ns;
            if (nsg < 0) break;
            if (trk_push_raw(sm->raw[k][q & 1], nsg, &sm->tab[k], cl_map(&sm->t.segs[2 * slot + k][0], tr), lane)) atomicMax(err, 1);
            if (lane == 0) {
                const uint32_t h = cl_map(&sm->slot[slot], tr);
                cl_st_u32(h + (k == 0 ? offsetof(trk_slot_t, ncarr) : offsetof(trk_slot_t, ncode)), (uint32_t)nsg);
            }
            cl_mbar_arrive_remote(cl_map(k == 0 ? &sm->fullc[slot] : &sm->fullk[slot], tr));
            __syncwarp();
            if (lane == 0) cl_flag_set(&sm->pushed[k], q + 1);
            TRK_PF(11);                                      /* pushers (both add to one slot) */
        }
    } else if (ctrl && wid == 2) {
        /* ================= epilogue warp: periods in order, as their tap sums arrive */
        int donep = 0, errw = 0;
        for (int q = 0; q < nperiods; q++) {
            TRK_PF0();
            cl_flag_wait_ge(&sm->seq_code, q + 1);
            const trk_hand_t *hd = &sm->hand[q % R];
            if (!hd->go) break;
            cl_mbar_wait(&sm->done[q % R], (uint32_t)((q / R) & 1));
            cl_flag_wait_ge(&sm->seq_remc, q + 1);
            cl_flag_wait_ge(&sm->seq_remp, q + 1);
            const unsigned int *res = sm->res[q % R];
            TRK_PF(5);                                       /* epilogue warp: waiting for the period's results */
            if (lane == 0) { st->currnsamp = hd->n; if (res[2 * SDRB200_MAXTAPS]) errw = 1; }
            __syncwarp();
            trk_epilogue(ch, st, res, res + SDRB200_MAXTAPS, hd->remc, hd->remp, recs ? &sm->rec : nullptr, lane,
                         (prof && chno == 0) ? prof : nullptr, &sm->lf, &sm->loops_count, q + 1, obsst ? &obsst[chno] : nullptr,
                         sm->vdec);
            if (recs) {
                const uint32_t *src = (const uint32_t *)&sm->rec;
                uint32_t *dst = (uint32_t *)&recs[(size_t)chno * nperiods + q];
                for (int i = lane; i < (int)(sizeof(sdrb200_trk_rec_t) / 4); i += 32) dst[i] = src[i];
            }
            __syncwarp();
            if (lane == 0) cl_flag_set(&sm->epi_count, q + 1);
            donep++;
        }
        __syncwarp();
        {
            uint32_t *dst = (uint32_t *)&states[chno];
            const uint32_t *src = (const uint32_t *)st;
            for (int i = lane; i < (int)(sizeof(trk_state_t) / 4); i += 32) dst[i] = src[i];
        }
        if (lane == 0) { nrec[chno] = donep; if (errw) atomicMax(err, 1); if (prof && chno == 0) prof[6] = donep; }
    } else if (wid >= TRK_T0 / 32 && rank >= r0) {
        /* ================= worker team: one whole code period per hand-out */
        const int t = threadIdx.x - TRK_T0;
        const bool pfw = (prof != nullptr) && chno == 0 && rank == r0 && t == TRK_PROF_THREAD;
        long long wt0 = pfw ? trk_clock() : 0;
        trk_wprof_t wpf; wpf.p = pfw ? prof : nullptr; wpf.t0 = wt0;
#define TRK_PFW(i) do { if (pfw) { const long long t1_ = trk_clock(); atomicAdd((unsigned long long *)&prof[i], (unsigned long long)(t1_ - wt0)); wt0 = t1_; wpf.t0 = t1_; } } while (0)
        for (int i = 0;; i++) {
            const int slot = i & 1;
            const uint32_t par = (uint32_t)((i >> 1) & 1);
            cl_mbar_wait(&sm->fullh[slot], par);
            const trk_slot_t *h = &sm->slot[slot];
            if (h->stop) break;
            const int n = h->n;
            const int8_t *data = ifbuf + (size_t)h->boff * DTYPE;
            /* the samples of the next period, and of the next one this team will see, towards L2 */
            {
                const size_t off = (size_t)t * TRK_SPAN;
                if (off < (size_t)n * DTYPE) {
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(data + (size_t)n * DTYPE + off));
                    if (W > 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(data + (size_t)n * DTYPE * W + off));
                }
            }
            TRK_PFW(9);                                      /* team: waiting for a hand-out */
            if (pfw) atomicAdd((unsigned long long *)&prof[31], 1ull);
            trkd_period<DTYPE, TRK_SPAN, NW, TRK_T0>(&sm->t, wk, data, n, ns, smax, i & 1, 0, 1,
                                           sm->t.segs[2 * slot], sm->t.segs[2 * slot + 1],
                                           [&]() -> int { cl_mbar_wait(&sm->fullc[slot], par); return h->ncarr; },
                                           [&]() -> int { cl_mbar_wait(&sm->fullk[slot], par); return h->ncode; }, &wpf);
            trk_bar_workers<NW>();                       /* sm->t.sum complete */
            TRK_PFW(7);                                      /* team: the period's dot products */
            if (t < 32) {
                const int q = h->q;
                const uint32_t dst = cl_map(&sm->res[q % R][0], 0);
                for (int k = lane; k < 2 * SDRB200_MAXTAPS; k += 32) cl_st_u32(dst + 4u * (uint32_t)k, sm->t.sum[k]);
                if (lane == 0) cl_st_u32(dst + 4u * (2 * SDRB200_MAXTAPS), (uint32_t)sm->t.err);
                cl_mbar_arrive_remote(cl_map(&sm->done[q % R], 0));
            }
            TRK_PFW(8);                                      /* team: chip boundaries + tap sums + report */
        }
    }
#undef TRK_PF
#undef TRK_PF0
#undef TRK_PFW
    /* nobody leaves while a peer may still write into (or wait on) its shared memory */
    __syncthreads();
    cl_sync_all();
}
