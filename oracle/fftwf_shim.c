/* fftwf_shim.c - ORACLE-ONLY stand-in for the seven FFTW entry points the
 * reference links against (test infrastructure, never product code).
 *
 * The reference calls, and only calls (src/sdrcmn.c:116,125,140-145,164-169;
 * src/sdrrcv.c:21):
 *     fftwf_malloc, fftwf_free, fftwf_init_threads, fftwf_plan_with_nthreads,
 *     fftwf_plan_dft_1d (in place, FFTW_ESTIMATE), fftwf_execute_dft,
 *     fftwf_destroy_plan.
 * Real FFTW (FFTW/fftw3 @ c69d6c3c) cannot be built in this image, so these
 * symbols are provided on top of oracle/offt.c, which evaluates the DFT that
 * FFTW documents (forward sign -1, backward +1, unnormalised) in double
 * precision and rounds once to float.  Rows of any report that used this shim
 * are labelled "reference + shim FFT (not FFTW)".
 *
 * The prototypes below restate the FFTW 3 C ABI for these seven symbols
 * (lib/fftw3/api/fftw3.h); the header itself is not needed to build the shim.
 */
#include <stdlib.h>
#include <time.h>
#include <pthread.h>
#include "offt.h"

typedef float fftwf_complex_t[2];
struct fftwf_plan_s { offt_plan *impl; };
typedef struct fftwf_plan_s *fftwf_plan_t;

/* tiny plan cache: the reference creates and destroys a plan per call
 * (src/sdrcmn.c:141-143), which with FFTW_ESTIMATE is cheap; recomputing the
 * twiddle table per call here would dominate the run time instead. */
#define NCACHE 16
static struct { int n, sign; offt_plan *p; } g_cache[NCACHE];
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;

static offt_plan *cached_plan(int n, int sign)
{
    int i, slot = -1;
    offt_plan *p = NULL;
    pthread_mutex_lock(&g_lock);
    for (i = 0; i < NCACHE; i++) {
        if (g_cache[i].p && g_cache[i].n == n && g_cache[i].sign == sign) {
            p = g_cache[i].p; break;
        }
        if (!g_cache[i].p && slot < 0) slot = i;
    }
    if (!p) {
        p = offt_create(n, sign);
        if (p && slot >= 0) { g_cache[slot].n = n; g_cache[slot].sign = sign; g_cache[slot].p = p; }
        /* cache full: the plan leaks by design (bounded: test processes only) */
    }
    pthread_mutex_unlock(&g_lock);
    return p;
}

void *fftwf_malloc(size_t n)
{
    void *p = NULL;
    if (posix_memalign(&p, 64, n ? n : 64)) return NULL;
    return p;
}

void fftwf_free(void *p) { free(p); }

int fftwf_init_threads(void) { return 1; }

void fftwf_plan_with_nthreads(int nthreads) { (void)nthreads; }

fftwf_plan_t fftwf_plan_dft_1d(int n, fftwf_complex_t *in, fftwf_complex_t *out,
                               int sign, unsigned flags)
{
    fftwf_plan_t pl;
    (void)in; (void)out; (void)flags;
    pl = (fftwf_plan_t)malloc(sizeof(*pl));
    if (!pl) return NULL;
    pl->impl = cached_plan(n, sign);
    if (!pl->impl) { free(pl); return NULL; }
    return pl;
}

/* seconds this process has spent inside fftwf_execute_dft (all threads added up); bench.py reports the
 * share of the CPU arm that is FFT time, i.e. what a real FFTW build would shrink            */
static double g_fft_seconds = 0.0;
double oracle_fft_seconds(int reset)
{
    double v;
    pthread_mutex_lock(&g_lock);
    v = g_fft_seconds;
    if (reset) g_fft_seconds = 0.0;
    pthread_mutex_unlock(&g_lock);
    return v;
}

void fftwf_execute_dft(const fftwf_plan_t pl, fftwf_complex_t *in, fftwf_complex_t *out)
{
    struct timespec t0, t1;
    /* the reference only ever transforms in place (in == out) */
    if (!pl || !pl->impl) return;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    if (in != out) {
        int i, n = offt_length(pl->impl);
        for (i = 0; i < n; i++) { out[i][0] = in[i][0]; out[i][1] = in[i][1]; }
    }
    offt_exec_f32(pl->impl, (float *)out);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    pthread_mutex_lock(&g_lock);
    g_fft_seconds += (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    pthread_mutex_unlock(&g_lock);
}

void fftwf_destroy_plan(fftwf_plan_t pl) { free(pl); }
