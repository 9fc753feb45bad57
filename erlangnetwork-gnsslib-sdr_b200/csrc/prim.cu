/* prim.cu - single-call primitives with HOST buffers: the per-function surface
 * of the reference's sdr.h that sits beside the batched engines.
 *
 *   sdrb200_mixcarr      <-> mixcarr()      src/sdrcmn.c:633-669
 *   sdrb200_rescode      <-> rescode()      src/sdrcmn.c:608-621
 *   sdrb200_dot2x        <-> dot_21/22/23   src/sdrcmn.c:287-354
 *   sdrb200_cpxcpx       <-> cpxcpx()       src/sdrcmn.c:185-195
 *   sdrb200_cpxfft       <-> cpxfft/cpxifft src/sdrcmn.c:134-175   (ANY length)
 *   sdrb200_cpxconv      <-> cpxconv()      src/sdrcmn.c:228-251
 *   sdrb200_cpxpspec     <-> cpxpspec()     src/sdrcmn.c:261-276
 *   sdrb200_pcorrelator  <-> pcorrelator()  src/sdrcmn.c:738-773
 *   sdrb200_checkacquisition <-> checkacquisition() src/sdracq.c:71-95
 *
 * The reference transforms arbitrary lengths through FFTW (nfft = 2*nsamp is
 * 32736 = 2^5*3*11*31 at 16.368 MS/s).  Here a length that is not one of the
 * CTA transform sizes goes through Bluestein's chirp-z identity on top of the
 * power-of-two transforms of fft_device.cuh; chirps are generated in double
 * with the quadratic phase reduced in integer arithmetic.
 * Every call: H2D of the arguments on the calling thread's stream, kernels,
 * D2H of the results, one stream synchronise.  No CPU arithmetic on the data
 * (the NCO segment tables are O(100) scalar steps on the host, see nco_exact.h).
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <vector>
#include <mutex>
#include "sdrb200_internal.h"
#include "nco_exact.h"
#include "fft_device.cuh"

extern "C" int sdrb200_trk_lut_ready(void);

static __constant__ short p_cos[32] = {32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6,0,6,12,18,23,27,30,31};
static __constant__ short p_sin[32] = {0,6,12,18,23,27,30,31,32,31,30,27,23,18,12,6,0,-6,-12,-18,-23,-27,-30,-31,-32,-31,-30,-27,-23,-18,-12,-6};

#define LAUNCH_CHECK() do { sdrb200_count_launch(1); SDRB200_CUDA_TRY(cudaGetLastError()); } while (0)

/* ------------------------------------------------------------ NCO helpers */
/* segments describing `count` values of the sequence of nco_walk(), for any
 * count: the walk is restarted every `chunk` values so that the table of one
 * call never overflows                                                        */
static int walk_all(double x, double step, long long count, double wrap, std::vector<nco_seg_t> &segs,
                    double *x_end)
{
    nco_table_t *t = (nco_table_t *)malloc(sizeof(nco_table_t));
    if (!t) return SDRB200_ENOMEM;
    const int chunk = 8192;
    long long done = 0;
    while (done < count) {
        int c = (int)((count - done < chunk) ? count - done : chunk);
        if (nco_walk(x, step, c, wrap, t)) { free(t); sdrb200_set_error("NCO table overflow"); return SDRB200_ERANGE; }
        for (int i = 0; i < t->nseg; i++) {
            nco_seg_t s = t->seg[i];
            if ((long long)s.k0 + done > 0x7fffffffLL) { free(t); return SDRB200_ERANGE; }
            s.k0 += (int32_t)done;
            segs.push_back(s);
        }
        x = t->x_end;
        done += c;
    }
    *x_end = x;
    free(t);
    return SDRB200_OK;
}

__device__ __forceinline__ double seg_value(const nco_seg_t *seg, int nseg, int k)
{
    int lo = 0, hi = nseg - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (seg[mid].k0 <= k) lo = mid; else hi = mid - 1;
    }
    return __dadd_rn(seg[lo].x0, __dmul_rn((double)(k - seg[lo].k0), seg[lo].d));
}

/* ------------------------------------------------------- mixcarr / rescode */
__global__ void k_mixcarr(const int8_t *__restrict__ data, int dtype, int n, const nco_seg_t *__restrict__ seg,
                          int nseg, short *__restrict__ I, short *__restrict__ Q)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const int idx = __double2int_rz(seg_value(seg, nseg, k)) & SDRB200_CMASK;
        const int c = p_cos[idx], s = p_sin[idx];
        if (dtype == 1) {
            const int x = data[k];
            I[k] = (short)(c * x); Q[k] = (short)(s * x);
        } else {
            const int x0 = data[2 * k], x1 = data[2 * k + 1];
            I[k] = (short)(c * x0 - s * x1); Q[k] = (short)(s * x0 + c * x1);
        }
    }
}

extern "C" int sdrb200_mixcarr(const int8_t *data, int dtype, double ti, int n, double freq,
                               double phi0, short *II, short *QQ, double *rem)
{
    if (!data || !II || !QQ || n < 1 || (dtype != 1 && dtype != 2)) { sdrb200_set_error("mixcarr: bad argument"); return SDRB200_EINVAL; }
    int rc = sdrb200_trk_lut_ready();
    if (rc) return rc;
    std::vector<nco_seg_t> segs;
    double phi, ps, x_end;
    nco_carrier_init(freq, phi0, ti, &phi, &ps);
    rc = walk_all(phi, ps, n, 0.0, segs, &x_end);
    if (rc) return rc;
    const size_t nb = (size_t)n * dtype, sb = segs.size() * sizeof(nco_seg_t);
    const size_t o_seg = (nb + 255) & ~(size_t)255, o_I = o_seg + ((sb + 255) & ~(size_t)255);
    const size_t o_Q = o_I + (((size_t)n * 2 + 255) & ~(size_t)255), tot = o_Q + (size_t)n * 2 + 256;
    sdrb200_ws_t *w = sdrb200_ws(tot, tot);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    memcpy(w->h_arena, data, nb);
    memcpy(w->h_arena + o_seg, segs.data(), sb);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, w->h_arena, o_seg + sb, cudaMemcpyHostToDevice, st));
    k_mixcarr<<<(n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256, 256, 0, st>>>(
        (const int8_t *)w->d_arena, dtype, n, (const nco_seg_t *)(w->d_arena + o_seg), (int)segs.size(),
        (short *)(w->d_arena + o_I), (short *)(w->d_arena + o_Q));
    LAUNCH_CHECK();
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena + o_I, w->d_arena + o_I, o_Q - o_I + (size_t)n * 2, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(II, w->h_arena + o_I, (size_t)n * 2);
    memcpy(QQ, w->h_arena + o_Q, (size_t)n * 2);
    if (rem) *rem = nco_carrier_rem(x_end);
    return SDRB200_OK;
}

__global__ void k_rescode(const short *__restrict__ code, int total, const nco_seg_t *__restrict__ seg, int nseg,
                          short *__restrict__ out)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < total; k += gridDim.x * blockDim.x)
        out[k] = code[__double2int_rz(seg_value(seg, nseg, k))];
}

extern "C" int sdrb200_rescode(const short *code, int len, double coff, int smax, double ci, int n,
                               short *rcode, double *rem)
{
    if (!code || !rcode || len < 1 || smax < 0 || n < 1 || !(ci > 0)) { sdrb200_set_error("rescode: bad argument"); return SDRB200_EINVAL; }
    std::vector<nco_seg_t> segs;
    const int total = n + 2 * smax;
    double x_end;
    int rc = walk_all(nco_code_init(coff, smax, ci, len), ci, total, (double)len, segs, &x_end);
    if (rc) return rc;
    const size_t cb = (size_t)len * 2, sb = segs.size() * sizeof(nco_seg_t), ob = (size_t)total * 2;
    const size_t o_seg = (cb + 255) & ~(size_t)255, o_out = o_seg + ((sb + 255) & ~(size_t)255);
    sdrb200_ws_t *w = sdrb200_ws(o_out + ob + 256, o_out + ob + 256);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    memcpy(w->h_arena, code, cb);
    memcpy(w->h_arena + o_seg, segs.data(), sb);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, w->h_arena, o_seg + sb, cudaMemcpyHostToDevice, st));
    k_rescode<<<(total + 255) / 256 > 1184 ? 1184 : (total + 255) / 256, 256, 0, st>>>(
        (const short *)w->d_arena, total, (const nco_seg_t *)(w->d_arena + o_seg), (int)segs.size(),
        (short *)(w->d_arena + o_out));
    LAUNCH_CHECK();
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena + o_out, w->d_arena + o_out, ob, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(rcode, w->h_arena + o_out, ob);
    if (rem) *rem = nco_code_rem(x_end, smax, ci);
    return SDRB200_OK;
}

/* ------------------------------------------------------------------ dots */
/* a1,a2 and the nb replicas live in one device array at element offsets; sums
 * are exact integers (64-bit atomics across CTAs)                              */
__global__ void k_dot2x(const short *__restrict__ base, size_t o_a1, size_t o_a2, size_t o_b0, size_t o_b1,
                        size_t o_b2, int nb, int n, long long *__restrict__ out)
{
    long long acc[6] = {0, 0, 0, 0, 0, 0};
    const short *a1 = base + o_a1, *a2 = base + o_a2;
    const short *b[3] = {base + o_b0, base + o_b1, base + o_b2};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int x1 = a1[i], x2 = a2[i];
        for (int j = 0; j < nb; j++) { const int c = b[j][i]; acc[j] += x1 * c; acc[3 + j] += x2 * c; }
    }
    for (int j = 0; j < 6; j++) {
        long long v = acc[j];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v != 0) atomicAdd((unsigned long long *)&out[j], (unsigned long long)v);
    }
}

extern "C" int sdrb200_dot2x(const short *a1, const short *a2, const short *const *b, int nb, int n,
                             double *d1, double *d2)
{
    if (!a1 || !a2 || !b || nb < 1 || nb > 3 || n < 1 || !d1 || !d2) { sdrb200_set_error("dot2x: bad argument"); return SDRB200_EINVAL; }
    const size_t stride = ((size_t)n + 127) & ~(size_t)127;          /* elements */
    const size_t bytes = stride * 2 * (2 + nb), o_out = (bytes + 255) & ~(size_t)255;
    sdrb200_ws_t *w = sdrb200_ws(o_out + 256, o_out + 256);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    short *h = (short *)w->h_arena;
    memcpy(h, a1, (size_t)n * 2);
    memcpy(h + stride, a2, (size_t)n * 2);
    for (int j = 0; j < nb; j++) memcpy(h + stride * (2 + j), b[j], (size_t)n * 2);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, h, bytes, cudaMemcpyHostToDevice, st));
    SDRB200_CUDA_TRY(cudaMemsetAsync(w->d_arena + o_out, 0, 48, st));
    int blocks = (n + 255) / 256; if (blocks > 592) blocks = 592;
    k_dot2x<<<blocks, 256, 0, st>>>((const short *)w->d_arena, 0, stride, 2 * stride, 3 * stride, 4 * stride, nb, n,
                                    (long long *)(w->d_arena + o_out));
    LAUNCH_CHECK();
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena + o_out, w->d_arena + o_out, 48, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    const long long *r = (const long long *)(w->h_arena + o_out);
    for (int j = 0; j < nb; j++) { d1[j] = (double)r[j]; d2[j] = (double)r[3 + j]; }
    return SDRB200_OK;
}

/* ---------------------------------------------------------------- cpxcpx */
__global__ void k_cpxcpx(const short *__restrict__ I, const short *__restrict__ Q, float scale, int n,
                         float2 *__restrict__ out)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = make_float2((float)I[i] * scale, Q ? (float)Q[i] * scale : 0.0f);
}

extern "C" int sdrb200_cpxcpx(const short *II, const short *QQ, double scale, int n, float *cpx)
{
    if (!II || !cpx || n < 1) { sdrb200_set_error("cpxcpx: bad argument"); return SDRB200_EINVAL; }
    const size_t sb = ((size_t)n * 2 + 255) & ~(size_t)255, ob = (size_t)n * 8;
    sdrb200_ws_t *w = sdrb200_ws(2 * sb + ob + 256, 2 * sb + ob + 256);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    memcpy(w->h_arena, II, (size_t)n * 2);
    if (QQ) memcpy(w->h_arena + sb, QQ, (size_t)n * 2);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, w->h_arena, 2 * sb, cudaMemcpyHostToDevice, st));
    k_cpxcpx<<<(n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256, 256, 0, st>>>(
        (const short *)w->d_arena, QQ ? (const short *)(w->d_arena + sb) : nullptr, (float)scale, n,
        (float2 *)(w->d_arena + 2 * sb));
    LAUNCH_CHECK();
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena + 2 * sb, w->d_arena + 2 * sb, ob, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(cpx, w->h_arena + 2 * sb, ob);
    return SDRB200_OK;
}

/* ===================================================== any-length batched FFT */
static bool cta_size(int n) { return n >= 256 && n <= 16384 && (n & (n - 1)) == 0; }
static bool pow2(int n) { return n > 0 && (n & (n - 1)) == 0; }

/* M == L: one CTA per batch item, natural order in and out */
template <int L, int SIGN>
__global__ void __launch_bounds__(L / 16)
k_fft_cta(const float2 *__restrict__ in, float2 *__restrict__ out, const float2 *__restrict__ tw)
{
    extern __shared__ float2 fsm[];
    constexpr int T = L / 16;
    const float2 *x = in + (size_t)blockIdx.x * L;
    float2 *y = out + (size_t)blockIdx.x * L;
    float2 v[16];
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = x[threadIdx.x + e * T];
    cta_fft<L, SIGN>(v, fsm, tw, 1);
#pragma unroll
    for (int e = 0; e < 16; e++) y[threadIdx.x + e * T] = v[e];
}

/* M = R*L > 16384: grid (R, batch); CTA q builds the L-point sequence
 * W_M^(s b q) * sum_a x[L a + b] W_R^(s a q), transforms it and writes X[R j + q] */
template <int L, int SIGN>
__global__ void __launch_bounds__(L / 16)
k_fft_big(const float2 *__restrict__ in, float2 *__restrict__ out, const float2 *__restrict__ tw, int R)
{
    extern __shared__ float2 fsm[];
    constexpr int T = L / 16;
    const int q = blockIdx.x;
    const size_t M = (size_t)L * R;
    const float2 *x = in + (size_t)blockIdx.y * M;
    float2 *y = out + (size_t)blockIdx.y * M;
    float2 v[16];
#pragma unroll
    for (int e = 0; e < 16; e++) {
        const int b = threadIdx.x + e * T;
        float2 acc = x[b];
        for (int a = 1; a < R; a++) {
            float2 w = __ldg(&tw[(size_t)((a * q) % R) * L]);
            if (SIGN > 0) w.y = -w.y;
            acc = cadd(acc, cmul(x[(size_t)L * a + b], w));
        }
        float2 w = __ldg(&tw[(size_t)b * q]);
        if (SIGN > 0) w.y = -w.y;
        v[e] = cmul(acc, w);
    }
    cta_fft<L, SIGN>(v, fsm, tw, R);
#pragma unroll
    for (int e = 0; e < 16; e++) y[(size_t)(threadIdx.x + e * T) * R + q] = v[e];
}

__global__ void k_twiddle(float2 *tw, int M)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < M; k += gridDim.x * blockDim.x) {
        double s, c;
        sincospi(-2.0 * (double)k / (double)M, &s, &c);
        tw[k] = make_float2((float)c, (float)s);
    }
}
/* chirp w[k] = exp(-i*pi*k^2/n), phase reduced as (k^2 mod 2n) in integers */
__global__ void k_chirp(float2 *w, int n)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const unsigned long long r = ((unsigned long long)k * (unsigned long long)k) % (2ULL * (unsigned long long)n);
        double s, c;
        sincospi(-(double)r / (double)n, &s, &c);
        w[k] = make_float2((float)c, (float)s);
    }
}
/* Bluestein kernel sequence h[j] = conj(w[|j|]) wrapped into length M2 */
__global__ void k_bs_kernel(const float2 *__restrict__ w, int n, int M2, float2 *__restrict__ h)
{
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < M2; j += gridDim.x * blockDim.x) {
        float2 v = make_float2(0.f, 0.f);
        if (j < n) v = make_float2(w[j].x, -w[j].y);
        else if (M2 - j < n) v = make_float2(w[M2 - j].x, -w[M2 - j].y);
        h[j] = v;
    }
}
/* u[k] = x[k] * w[k]^(conj if inverse), zero-padded to M2; batch along y */
__global__ void k_bs_pre(const float2 *__restrict__ x, const float2 *__restrict__ w, int n, int M2, int inverse,
                         float2 *__restrict__ u)
{
    const float2 *xb = x + (size_t)blockIdx.y * n;
    float2 *ub = u + (size_t)blockIdx.y * M2;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < M2; k += gridDim.x * blockDim.x) {
        float2 v = make_float2(0.f, 0.f);
        if (k < n) { float2 c = w[k]; if (inverse) c.y = -c.y; v = cmul(xb[k], c); }
        ub[k] = v;
    }
}
/* U *= H (forward) or U *= conj(H) (inverse transform wanted) */
__global__ void k_bs_mul(float2 *__restrict__ U, const float2 *__restrict__ H, int M2, int inverse)
{
    float2 *ub = U + (size_t)blockIdx.y * M2;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < M2; k += gridDim.x * blockDim.x) {
        /* inverse: the kernel sequence is conj(h); h is even (h[j] = h[M2-j]) so its
         * spectrum is conj(H[(M2-k)%M2])                                           */
        float2 hk = inverse ? H[(M2 - k) & (M2 - 1)] : H[k];
        if (inverse) hk.y = -hk.y;
        ub[k] = cmul(ub[k], hk);
    }
}
/* X[k] = w[k]^(conj if inverse) * v[k] / M2 */
__global__ void k_bs_post(const float2 *__restrict__ v, const float2 *__restrict__ w, int n, int M2, int inverse,
                          float2 *__restrict__ X)
{
    const float2 *vb = v + (size_t)blockIdx.y * M2;
    float2 *xb = X + (size_t)blockIdx.y * n;
    const float inv = 1.0f / (float)M2;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        float2 c = w[k]; if (inverse) c.y = -c.y;
        float2 r = cmul(vb[k], c);
        xb[k] = make_float2(r.x * inv, r.y * inv);
    }
}

struct gfft_plan {
    int n = 0, M2 = 0, L = 0, R = 0;
    bool bluestein = false;
    float2 *tw = nullptr, *chirp = nullptr, *H = nullptr;
    void release() { cudaFree(tw); cudaFree(chirp); cudaFree(H); tw = chirp = H = nullptr; n = 0; }
};
struct gfft_cache { gfft_plan p[4]; int next = 0; ~gfft_cache() { for (auto &q : p) q.release(); } };
static thread_local gfft_cache g_fft;

template <int SIGN> static int pow2_fft(cudaStream_t st, const gfft_plan *p, const float2 *in, float2 *out, int batch)
{
#define GF_CASE(LL)                                                                                    \
    case LL: {                                                                                         \
        const size_t smem = sizeof(float2) * FFT_SMEM_ELEMS(LL);                                       \
        if (p->R == 1) {                                                                               \
            SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fft_cta<LL, SIGN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            k_fft_cta<LL, SIGN><<<batch, LL / 16, smem, st>>>(in, out, p->tw);                         \
        } else {                                                                                       \
            SDRB200_CUDA_TRY(cudaFuncSetAttribute(k_fft_big<LL, SIGN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            k_fft_big<LL, SIGN><<<dim3(p->R, batch), LL / 16, smem, st>>>(in, out, p->tw, p->R);       \
        }                                                                                              \
    } break;
    switch (p->L) {
        GF_CASE(256) GF_CASE(512) GF_CASE(1024) GF_CASE(2048) GF_CASE(4096) GF_CASE(8192) GF_CASE(16384)
    default: sdrb200_set_error("fft: unsupported sub-length %d", p->L); return SDRB200_EINVAL;
    }
#undef GF_CASE
    LAUNCH_CHECK();
    return SDRB200_OK;
}

static int gfft_get(cudaStream_t st, int n, gfft_plan **out)
{
    for (auto &q : g_fft.p) if (q.n == n) { *out = &q; return SDRB200_OK; }
    gfft_plan *p = &g_fft.p[g_fft.next];
    g_fft.next = (g_fft.next + 1) % 4;
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    p->release();
    p->bluestein = !(pow2(n) && n >= 256);
    p->M2 = n;
    if (p->bluestein) { p->M2 = 256; while (p->M2 < 2 * n - 1) p->M2 <<= 1; }
    p->L = p->M2 <= 8192 ? p->M2 : 8192;
    p->R = p->M2 / p->L;
    if (p->R > 512) { sdrb200_set_error("fft: length %d too long", n); return SDRB200_EINVAL; }
    SDRB200_CUDA_TRY(cudaMalloc((void **)&p->tw, sizeof(float2) * p->M2));
    k_twiddle<<<64, 256, 0, st>>>(p->tw, p->M2);
    LAUNCH_CHECK();
    if (p->bluestein) {
        SDRB200_CUDA_TRY(cudaMalloc((void **)&p->chirp, sizeof(float2) * n));
        SDRB200_CUDA_TRY(cudaMalloc((void **)&p->H, sizeof(float2) * 2 * (size_t)p->M2));
        k_chirp<<<64, 256, 0, st>>>(p->chirp, n);
        LAUNCH_CHECK();
        k_bs_kernel<<<64, 256, 0, st>>>(p->chirp, n, p->M2, p->H + p->M2);
        LAUNCH_CHECK();
        int rc = pow2_fft<-1>(st, p, p->H + p->M2, p->H, 1);
        if (rc) return rc;
    }
    p->n = n;
    *out = p;
    return SDRB200_OK;
}

/* batch transforms of length n, in place on d_x; scratch: 2*batch*M2 float2 when
 * Bluestein is needed, batch*n otherwise (out-of-place power-of-two pass)        */
static size_t gfft_scratch_elems(int n, int batch)
{
    int M2 = n;
    if (!(pow2(n) && n >= 256)) { M2 = 256; while (M2 < 2 * n - 1) M2 <<= 1; return 2 * (size_t)batch * M2; }
    return (size_t)batch * M2;
}
static int gfft_exec(cudaStream_t st, float2 *d_x, int n, int batch, int inverse, float2 *scratch)
{
    gfft_plan *p;
    int rc = gfft_get(st, n, &p);
    if (rc) return rc;
    if (!p->bluestein) {
        rc = inverse ? pow2_fft<+1>(st, p, d_x, scratch, batch) : pow2_fft<-1>(st, p, d_x, scratch, batch);
        if (rc) return rc;
        SDRB200_CUDA_TRY(cudaMemcpyAsync(d_x, scratch, sizeof(float2) * (size_t)batch * n, cudaMemcpyDeviceToDevice, st));
        return SDRB200_OK;
    }
    const int M2 = p->M2;
    float2 *u = scratch, *U = scratch + (size_t)batch * M2;
    dim3 g((M2 + 255) / 256 > 296 ? 296 : (M2 + 255) / 256, batch);
    k_bs_pre<<<g, 256, 0, st>>>(d_x, p->chirp, n, M2, inverse, u);
    LAUNCH_CHECK();
    rc = pow2_fft<-1>(st, p, u, U, batch);
    if (rc) return rc;
    k_bs_mul<<<g, 256, 0, st>>>(U, p->H, M2, inverse);
    LAUNCH_CHECK();
    rc = pow2_fft<+1>(st, p, U, u, batch);
    if (rc) return rc;
    k_bs_post<<<g, 256, 0, st>>>(u, p->chirp, n, M2, inverse, d_x);
    LAUNCH_CHECK();
    return SDRB200_OK;
}

extern "C" int sdrb200_cpxfft(float *cpx, int n, int inverse)
{
    if (!cpx || n < 1) { sdrb200_set_error("cpxfft: bad argument"); return SDRB200_EINVAL; }
    const size_t xb = ((size_t)n * 8 + 255) & ~(size_t)255;
    const size_t sb = gfft_scratch_elems(n, 1) * 8;
    sdrb200_ws_t *w = sdrb200_ws(xb + sb + 256, xb);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    memcpy(w->h_arena, cpx, (size_t)n * 8);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, w->h_arena, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    int rc = gfft_exec(st, (float2 *)w->d_arena, n, 1, inverse, (float2 *)(w->d_arena + xb));
    if (rc) return rc;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena, w->d_arena, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(cpx, w->h_arena, (size_t)n * 8);
    return SDRB200_OK;
}

/* ------------------------------------------------------ cpxconv / cpxpspec */
/* a = -(a * conj(b)), src/sdrcmn.c:236-240; b is broadcast over the batch */
__global__ void k_negmulconj(float2 *__restrict__ a, const float2 *__restrict__ b, int m)
{
    float2 *ab = a + (size_t)blockIdx.y * m;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const float2 p = ab[i], q = b[i];
        ab[i] = make_float2(-p.x * q.x - p.y * q.y, p.x * q.y - p.y * q.x);
    }
}
/* out[i] (+)= (re^2+im^2)/div in float, added into double (src/sdrcmn.c:244-250, 268-274) */
__global__ void k_power(const float2 *__restrict__ a, int stride, int n, float div, int flagsum, double *__restrict__ out)
{
    const float2 *ab = a + (size_t)blockIdx.y * stride;
    double *ob = out + (size_t)blockIdx.y * n;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float pw = (ab[i].x * ab[i].x + ab[i].y * ab[i].y) / div;
        ob[i] = flagsum ? ob[i] + (double)pw : (double)pw;
    }
}

extern "C" int sdrb200_cpxconv(float *cpxa, const float *cpxb, int m, int n, int flagsum, double *conv)
{
    if (!cpxa || !cpxb || !conv || m < 1 || n < 1 || n > m) { sdrb200_set_error("cpxconv: bad argument"); return SDRB200_EINVAL; }
    const size_t xb = ((size_t)m * 8 + 255) & ~(size_t)255, cb = ((size_t)n * 8 + 255) & ~(size_t)255;
    const size_t sb = gfft_scratch_elems(m, 1) * 8;
    sdrb200_ws_t *w = sdrb200_ws(2 * xb + cb + sb + 256, 2 * xb + cb);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    memcpy(w->h_arena, cpxa, (size_t)m * 8);
    memcpy(w->h_arena + xb, cpxb, (size_t)m * 8);
    if (flagsum) memcpy(w->h_arena + 2 * xb, conv, (size_t)n * 8);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, w->h_arena, 2 * xb + (flagsum ? (size_t)n * 8 : 0), cudaMemcpyHostToDevice, st));
    float2 *a = (float2 *)w->d_arena, *b = (float2 *)(w->d_arena + xb), *scr = (float2 *)(w->d_arena + 2 * xb + cb);
    double *d_conv = (double *)(w->d_arena + 2 * xb);
    int rc = gfft_exec(st, a, m, 1, 0, scr);
    if (rc) return rc;
    const int gb = (m + 255) / 256 > 1184 ? 1184 : (m + 255) / 256;
    k_negmulconj<<<dim3(gb, 1), 256, 0, st>>>(a, b, m);
    LAUNCH_CHECK();
    rc = gfft_exec(st, a, m, 1, 1, scr);
    if (rc) return rc;
    k_power<<<dim3(gb, 1), 256, 0, st>>>(a, m, n, (float)m * m, flagsum, d_conv);
    LAUNCH_CHECK();
    /* the reference leaves the inverse transform in cpxa (in-place contract) */
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena, w->d_arena, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena + 2 * xb, d_conv, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(cpxa, w->h_arena, (size_t)m * 8);
    memcpy(conv, w->h_arena + 2 * xb, (size_t)n * 8);
    return SDRB200_OK;
}

extern "C" int sdrb200_cpxpspec(float *cpx, int n, int flagsum, double *pspec)
{
    if (!cpx || !pspec || n < 1) { sdrb200_set_error("cpxpspec: bad argument"); return SDRB200_EINVAL; }
    const size_t xb = ((size_t)n * 8 + 255) & ~(size_t)255, sb = gfft_scratch_elems(n, 1) * 8;
    sdrb200_ws_t *w = sdrb200_ws(2 * xb + sb + 256, 2 * xb);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    memcpy(w->h_arena, cpx, (size_t)n * 8);
    if (flagsum) memcpy(w->h_arena + xb, pspec, (size_t)n * 8);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, w->h_arena, xb + (flagsum ? (size_t)n * 8 : 0), cudaMemcpyHostToDevice, st));
    float2 *a = (float2 *)w->d_arena;
    double *d_ps = (double *)(w->d_arena + xb);
    int rc = gfft_exec(st, a, n, 1, 0, (float2 *)(w->d_arena + 2 * xb));
    if (rc) return rc;
    k_power<<<dim3((n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256, 1), 256, 0, st>>>(a, n, n, 1.0f, flagsum, d_ps);
    LAUNCH_CHECK();
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena, w->d_arena, xb + (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(cpx, w->h_arena, (size_t)n * 8);
    memcpy(pspec, w->h_arena + xb, (size_t)n * 8);
    return SDRB200_OK;
}

/* ------------------------------------------------------------ pcorrelator */
/* wipe-off of the zero-padded window for a batch of Doppler bins: mixcarr with
 * phi0 = 0 (exact index table per bin) + cpxcpx scale, src/sdrcmn.c:756-764     */
__global__ void k_pc_mix(const int8_t *__restrict__ data, int dtype, int nlive, int m, const uint8_t *__restrict__ cidx,
                         float scale, float2 *__restrict__ x)
{
    const uint8_t *ci = cidx + (size_t)blockIdx.y * m;
    float2 *xb = x + (size_t)blockIdx.y * m;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < m; t += gridDim.x * blockDim.x) {
        float2 v = make_float2(0.f, 0.f);
        if (t < nlive) {
            const int idx = ci[t], c = p_cos[idx], s = p_sin[idx];
            if (dtype == 1) { const int d = data[t]; v = make_float2((float)(short)(c * d) * scale, (float)(short)(s * d) * scale); }
            else {
                const int d0 = data[2 * t], d1 = data[2 * t + 1];
                v = make_float2((float)(short)(c * d0 - s * d1) * scale, (float)(short)(s * d0 + c * d1) * scale);
            }
        }
        xb[t] = v;
    }
}

extern "C" int sdrb200_pcorrelator(const int8_t *data, int dtype, double ti, int n, const double *freq,
                                   int nfreq, double crate, int m, const float *codex, double *P)
{
    (void)crate;
    if (!data || !freq || !codex || !P || n < 1 || nfreq < 1 || m < 2 * n || (dtype != 1 && dtype != 2)) {
        sdrb200_set_error("pcorrelator: bad argument (m must be >= 2n)"); return SDRB200_EINVAL;
    }
    int rc = sdrb200_trk_lut_ready();
    if (rc) return rc;
    /* bins are processed in batches that keep the scratch within ~256 MB */
    const size_t per_bin = (gfft_scratch_elems(m, 1) + (size_t)m) * 8 + (size_t)m;
    int bb = (int)(((size_t)256 << 20) / per_bin); if (bb < 1) bb = 1; if (bb > nfreq) bb = nfreq;
    const size_t o_data = 0, nbytes = (size_t)2 * n * dtype;
    const size_t o_code = (nbytes + 255) & ~(size_t)255;
    const size_t o_cidx = o_code + (((size_t)m * 8 + 255) & ~(size_t)255);
    const size_t o_x = o_cidx + (((size_t)bb * m + 255) & ~(size_t)255);
    const size_t o_P = o_x + (((size_t)bb * m * 8 + 255) & ~(size_t)255);
    const size_t o_scr = o_P + (((size_t)bb * n * 8 + 255) & ~(size_t)255);
    const size_t d_tot = o_scr + gfft_scratch_elems(m, bb) * 8 + 256;
    const size_t h_tot = o_code + (size_t)m * 8 + (size_t)bb * m + (size_t)bb * n * 8 + 1024;
    sdrb200_ws_t *w = sdrb200_ws(d_tot, h_tot);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    char *h = w->h_arena;
    memcpy(h, data, nbytes);
    memcpy(h + o_code, codex, (size_t)m * 8);
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, h, o_code + (size_t)m * 8, cudaMemcpyHostToDevice, st));
    uint8_t *h_cidx = (uint8_t *)(h + o_code + (size_t)m * 8);
    double *h_P = (double *)(h + ((o_code + (size_t)m * 8 + (size_t)bb * m + 255) & ~(size_t)255));
    nco_table_t *tb = (nco_table_t *)malloc(sizeof(nco_table_t));
    if (!tb) return SDRB200_ENOMEM;
    const float scale = (float)(SDRB200_CSCALE / m);
    const int gb = (m + 255) / 256 > 296 ? 296 : (m + 255) / 256;
    for (int f0 = 0; f0 < nfreq && rc == SDRB200_OK; f0 += bb) {
        const int nb = (nfreq - f0 < bb) ? nfreq - f0 : bb;
        for (int j = 0; j < nb; j++) {
            double phi, ps;
            nco_carrier_init(freq[f0 + j], 0.0, ti, &phi, &ps);
            if (nco_walk(phi, ps, m, 0.0, tb)) { sdrb200_set_error("pcorrelator: NCO table overflow"); rc = SDRB200_ERANGE; break; }
            int32_t hint = 0;
            for (int t = 0; t < m; t++) h_cidx[(size_t)j * m + t] = (uint8_t)(((int)nco_value(tb, t, &hint)) & SDRB200_CMASK);
        }
        if (rc) break;
        memcpy(h_P, P + (size_t)f0 * n, (size_t)nb * n * 8);
        if (cudaMemcpyAsync(w->d_arena + o_cidx, h_cidx, (size_t)nb * m, cudaMemcpyHostToDevice, st) != cudaSuccess ||
            cudaMemcpyAsync(w->d_arena + o_P, h_P, (size_t)nb * n * 8, cudaMemcpyHostToDevice, st) != cudaSuccess) { rc = SDRB200_ECUDA; break; }
        float2 *x = (float2 *)(w->d_arena + o_x), *scr = (float2 *)(w->d_arena + o_scr);
        k_pc_mix<<<dim3(gb, nb), 256, 0, st>>>((const int8_t *)(w->d_arena + o_data), dtype, 2 * n, m,
                                              (const uint8_t *)(w->d_arena + o_cidx), scale, x);
        sdrb200_count_launch(1);
        if ((rc = gfft_exec(st, x, m, nb, 0, scr))) break;
        k_negmulconj<<<dim3(gb, nb), 256, 0, st>>>(x, (const float2 *)(w->d_arena + o_code), m);
        sdrb200_count_launch(1);
        if ((rc = gfft_exec(st, x, m, nb, 1, scr))) break;
        k_power<<<dim3(gb, nb), 256, 0, st>>>(x, m, n, (float)m * m, 1, (double *)(w->d_arena + o_P));
        sdrb200_count_launch(1);
        if (cudaMemcpyAsync(h_P, w->d_arena + o_P, (size_t)nb * n * 8, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
            cudaStreamSynchronize(st) != cudaSuccess) { sdrb200_set_error("pcorrelator: %s", cudaGetErrorString(cudaGetLastError())); rc = SDRB200_ECUDA; break; }
        memcpy(P + (size_t)f0 * n, h_P, (size_t)nb * n * 8);
    }
    free(tb);
    return rc;
}

/* ------------------------------------------------------- checkacquisition */
struct ca_out_t { double maxP, maxP2, sum; int flat, cnt; };

/* pass 1: global maxvd over nfreq*nsamp (first index wins) */
__global__ void k_ca_max(const double *__restrict__ P, long long total, unsigned long long *__restrict__ key_hi,
                         double *__restrict__ blockmax, long long *__restrict__ blockidx)
{
    __shared__ double s_v[32]; __shared__ long long s_i[32];
    double best = -1.0; long long bi = 0x7fffffffffffffffLL;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const double v = P[i];
        if (v > best) { best = v; bi = i; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        double ov = __shfl_xor_sync(0xffffffffu, best, o); long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { s_v[threadIdx.x >> 5] = best; s_i[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); w++)
            if (s_v[w] > best || (s_v[w] == best && s_i[w] < bi)) { best = s_v[w]; bi = s_i[w]; }
        blockmax[blockIdx.x] = best; blockidx[blockIdx.x] = bi;
    }
    (void)key_hi;
}
/* pass 2 (one CTA): finish the argmax, then masked mean / second max of the row */
__global__ void k_ca_row(const double *__restrict__ P, int nsamp, int nfreq, int nsampchip, const double *__restrict__ blockmax,
                         const long long *__restrict__ blockidx, int nblocks, ca_out_t *__restrict__ out)
{
    __shared__ double s_v[32]; __shared__ int s_c[32]; __shared__ double s_m[32];
    __shared__ long long s_flat; __shared__ double s_max;
    if (threadIdx.x == 0) {
        double best = blockmax[0]; long long bi = blockidx[0];
        for (int b = 1; b < nblocks; b++)
            if (blockmax[b] > best || (blockmax[b] == best && blockidx[b] < bi)) { best = blockmax[b]; bi = blockidx[b]; }
        s_flat = bi; s_max = best;
    }
    __syncthreads();
    const long long flat = s_flat;
    const int codei = (int)(flat % nsamp), freqi = (int)(flat / nsamp);
    int exs = codei - 2 * nsampchip; if (exs < 0) exs += nsamp;
    int exe = codei + 2 * nsampchip; if (exe >= nsamp) exe -= nsamp;
    const double *row = P + (size_t)freqi * nsamp;
    double sum = 0.0, m2 = -1.0; int cnt = 0;
    for (int i = threadIdx.x; i < nsamp; i += blockDim.x) {
        const bool inc = (exs <= exe) ? (i < exs || i > exe) : (i < exs && i > exe);
        if (inc) { const double v = row[i]; sum += v; cnt++; if (i >= 1 && v > m2) m2 = v; }
    }
    if (threadIdx.x == 0 && row[0] > m2) m2 = row[0];                     /* max=data[0] seed, sdrcmn.c:464 */
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        m2 = fmax(m2, __shfl_xor_sync(0xffffffffu, m2, o));
    }
    if ((threadIdx.x & 31) == 0) { s_v[threadIdx.x >> 5] = sum; s_c[threadIdx.x >> 5] = cnt; s_m[threadIdx.x >> 5] = m2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); w++) { sum += s_v[w]; cnt += s_c[w]; m2 = fmax(m2, s_m[w]); }
        out->maxP = s_max; out->maxP2 = m2; out->sum = sum; out->cnt = cnt; out->flat = (int)flat;
    }
    (void)nfreq;
}

extern "C" int sdrb200_checkacquisition(const double *P, int nsamp, int nfreq, int nsampchip, double ctime,
                                        double *cn0, double *peakr, int *codei, int *freqi)
{
    if (!P || nsamp < 1 || nfreq < 1 || !cn0 || !peakr || !codei || !freqi) { sdrb200_set_error("checkacquisition: bad argument"); return SDRB200_EINVAL; }
    const long long total = (long long)nsamp * nfreq;
    if (total > 0x7fffffffLL) { sdrb200_set_error("checkacquisition: grid too large"); return SDRB200_EINVAL; }
    const size_t pb = ((size_t)total * 8 + 255) & ~(size_t)255;
    const int nblocks = 296;
    sdrb200_ws_t *w = sdrb200_ws(pb + 16 * nblocks + 512, 512);
    if (!w) return SDRB200_ENODEV;
    cudaStream_t st = (cudaStream_t)w->cuda_stream;
    /* P is caller memory (pageable): copied straight from it */
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->d_arena, P, (size_t)total * 8, cudaMemcpyHostToDevice, st));
    double *bm = (double *)(w->d_arena + pb); long long *bi = (long long *)(bm + nblocks);
    ca_out_t *d_out = (ca_out_t *)(bi + nblocks);
    k_ca_max<<<nblocks, 256, 0, st>>>((const double *)w->d_arena, total, nullptr, bm, bi);
    LAUNCH_CHECK();
    k_ca_row<<<1, 1024, 0, st>>>((const double *)w->d_arena, nsamp, nfreq, nsampchip, bm, bi, nblocks, d_out);
    LAUNCH_CHECK();
    SDRB200_CUDA_TRY(cudaMemcpyAsync(w->h_arena, d_out, sizeof(ca_out_t), cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    const ca_out_t *o = (const ca_out_t *)w->h_arena;
    const double meanP = o->sum / (double)o->cnt;
    *cn0 = 10 * log10(o->maxP / meanP / ctime);
    *peakr = o->maxP / o->maxP2;
    *codei = o->flat % nsamp;
    *freqi = o->flat / nsamp;
    return SDRB200_OK;
}

/* ------------------------------------------- checkacquisition, allocation-free
 * sdrb200_checkacquisition_noalloc: same arguments, same kernels and same results as
 * sdrb200_checkacquisition above, for callers that must not allocate (CUDA graph
 * capture around them, serving loops, harnesses that forbid cudaMalloc).  No counterpart
 * in the reference.
 *
 * The scratch the two kernels need - the device copy of P, the per-block partial maxima
 * and the result record - lives in ONE file-scope device array of fixed size
 * (SDRB200_CHECKACQ_NOALLOC_MAXCELLS doubles + 8 KB, ~16 MB held by every GPU that loads
 * the library, until exit).  In host code `g_ca_scratch` is only a stand-in for the
 * symbol: its DEVICE address is fetched once per device with cudaGetSymbolAddress (that
 * device current) and kept in g_ca_addr[].  The call makes no cudaMalloc / cudaFree /
 * cudaMallocHost / stream or event creation: copies run on the legacy default stream from
 * and to the caller's (pageable) memory.  Every call shares the array and finds the
 * previous call's contents in it: nothing has to start at zero (both kernels write every
 * word they later read; the tail area is nevertheless cleared per call), and calls
 * that overlap on the same GPU would race on it.  Calls from several host threads of one
 * process are therefore serialised by a mutex here (each call ends with a stream
 * synchronise inside the lock); nothing else may use the array.                        */
#define CA_NOALLOC_BLOCKS 296
#define CA_NOALLOC_TAIL   8192          /* partial maxima (296 x 16 B) + ca_out_t, rounded up */
#define CA_NOALLOC_MAXDEV 64
static __device__ __align__(256) unsigned char g_ca_scratch[(size_t)SDRB200_CHECKACQ_NOALLOC_MAXCELLS * 8 + CA_NOALLOC_TAIL];
static unsigned char *g_ca_addr[CA_NOALLOC_MAXDEV];     /* device address of g_ca_scratch, per device */
static std::mutex g_ca_mtx;

extern "C" int sdrb200_checkacquisition_noalloc(const double *P, int nsamp, int nfreq, int nsampchip, double ctime,
                                                double *cn0, double *peakr, int *codei, int *freqi)
{
    if (!P || nsamp < 1 || nfreq < 1 || !cn0 || !peakr || !codei || !freqi) {
        sdrb200_set_error("checkacquisition_noalloc: bad argument"); return SDRB200_EINVAL;
    }
    const long long total = (long long)nsamp * nfreq;
    if (total > (long long)SDRB200_CHECKACQ_NOALLOC_MAXCELLS) {
        sdrb200_set_error("checkacquisition_noalloc: %lld cells exceed the static scratch of %d cells "
                          "(use sdrb200_checkacquisition)", total, SDRB200_CHECKACQ_NOALLOC_MAXCELLS);
        return SDRB200_ERANGE;
    }
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    std::lock_guard<std::mutex> lock(g_ca_mtx);
    int dev = 0;
    SDRB200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= CA_NOALLOC_MAXDEV) { sdrb200_set_error("checkacquisition_noalloc: device index %d out of range", dev); return SDRB200_EINVAL; }
    if (!g_ca_addr[dev]) {                              /* set-up, once per device, that device current */
        unsigned char *p = nullptr;
        SDRB200_CUDA_TRY(cudaGetSymbolAddress((void **)&p, g_ca_scratch));
        if (!p) { sdrb200_set_error("checkacquisition_noalloc: no device address for the scratch symbol"); return SDRB200_ECUDA; }
        g_ca_addr[dev] = p;
    }
    unsigned char *base = g_ca_addr[dev];
    const size_t pb = (size_t)SDRB200_CHECKACQ_NOALLOC_MAXCELLS * 8;
    double *d_P = (double *)base;
    double *bm = (double *)(base + pb); long long *bi = (long long *)(bm + CA_NOALLOC_BLOCKS);
    ca_out_t *d_out = (ca_out_t *)(bi + CA_NOALLOC_BLOCKS);
    cudaStream_t st = 0;                                /* legacy default stream: nothing to create */
    SDRB200_CUDA_TRY(cudaMemsetAsync(base + pb, 0, CA_NOALLOC_TAIL, st));
    SDRB200_CUDA_TRY(cudaMemcpyAsync(d_P, P, (size_t)total * 8, cudaMemcpyHostToDevice, st));
    k_ca_max<<<CA_NOALLOC_BLOCKS, 256, 0, st>>>(d_P, total, nullptr, bm, bi);
    LAUNCH_CHECK();
    k_ca_row<<<1, 1024, 0, st>>>(d_P, nsamp, nfreq, nsampchip, bm, bi, CA_NOALLOC_BLOCKS, d_out);
    LAUNCH_CHECK();
    ca_out_t o;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(&o, d_out, sizeof(ca_out_t), cudaMemcpyDeviceToHost, st));
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    const double meanP = o.sum / (double)o.cnt;
    *cn0 = 10 * log10(o.maxP / meanP / ctime);
    *peakr = o.maxP / o.maxP2;
    *codei = o.flat % nsamp;
    *freqi = o.flat / nsamp;
    return SDRB200_OK;
}
